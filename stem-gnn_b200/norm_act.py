"""BatchNorm1d + ReLU + dropout of the encoder layers as one fused op (K7).

`bn_act_drop(z, bn, relu, p, training)` computes what the reference's encode loop does with three
torch modules -- `z = self.norms[i](z); z = self.dropout(self.activation(z))`
(STEM-GNN/model/encoder.py:313-317) -- on the SAME `nn.BatchNorm1d` module (its parameters, running
statistics and `num_batches_tracked` are read and updated exactly like `F.batch_norm` does), so
`state_dict` keys and eval behaviour are unchanged.  Dropout uses the library's own counter-based
Philox stream (not torch's generator): same distribution, different masks.  No CPU path.
"""
import ctypes as C

import torch
import torch.nn as nn

from . import _lib

_seeds = {}          # device index -> uint64-as-int64 device tensor: the dropout stream position
_GOLDEN = 0x9E3779B97F4A7C15 - (1 << 64)      # odd increment of the seed (as a signed int64)


def _seed_tensor(device):
    key = device.index if device.index is not None else torch.cuda.current_device()
    t = _seeds.get(key)
    if t is None:
        t = torch.tensor([torch.initial_seed() & 0x7FFFFFFFFFFFFFFF], dtype=torch.int64, device=device)
        _seeds[key] = t
    return t


def manual_seed(seed: int, device=None):
    """Re-seed the dropout stream of `device` (default: current)."""
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    _seed_tensor(device).fill_(int(seed) & 0x7FFFFFFFFFFFFFFF)


def supported(z, bn, activation) -> bool:
    return (z.is_cuda and z.dim() == 2 and z.dtype == torch.float32 and z.size(1) % 4 == 0 and z.size(0) > 0
            and isinstance(bn, nn.BatchNorm1d) and (activation is None or isinstance(activation, nn.ReLU)))


def _ws(rows, feat, device):
    need = C.c_size_t(0)
    _lib.check(_lib.lib().stemgnn_bn_workspace(rows, feat, C.byref(need)))
    return _lib.workspace(need.value, device)


def _rows16(t):
    if t.stride(1) != 1 or t.stride(0) % 4 or t.data_ptr() % 16 or t.stride(0) < t.size(1):
        t = t.contiguous()
    return t


def batch_stats(x):
    """(mean [F], biased variance [F]) of the columns of x (K7 statistics pass)."""
    x = _rows16(x)
    n, f = x.shape
    mean = torch.empty(f, dtype=torch.float32, device=x.device)
    var = torch.empty_like(mean)
    ws = _ws(n, f, x.device)
    with _lib.traced("bn_stats", n * f * 4, 0, 2):
        _lib.check(_lib.lib().stemgnn_bn_stats(n, f, x.data_ptr(), x.stride(0), mean.data_ptr(), var.data_ptr(),
                                               ws.data_ptr(), ws.numel(), _lib.stream()))
    return mean, var


def train_stats(x, eps, momentum, num_batches_tracked, running_mean, running_var):
    """(mean, biased var, invstd) of a training batch and, in the same launches, nn.BatchNorm1d's running-statistics
    update (stemgnn_bn_train_stats).  momentum None = cumulative average over `num_batches_tracked` (already
    incremented by the caller)."""
    x = _rows16(x)
    n, f = x.shape
    mean = torch.empty(f, dtype=torch.float32, device=x.device)
    var, invstd = torch.empty_like(mean), torch.empty_like(mean)
    ws = _ws(n, f, x.device)
    with _lib.traced("bn_stats", n * f * 4, 0, 2):
        _lib.check(_lib.lib().stemgnn_bn_train_stats(
            n, f, x.data_ptr(), x.stride(0), float(eps), -1.0 if momentum is None else float(momentum),
            _lib.ptr(num_batches_tracked), _lib.ptr(running_mean), _lib.ptr(running_var), mean.data_ptr(),
            var.data_ptr(), invstd.data_ptr(), ws.data_ptr(), ws.numel(), _lib.stream()))
    return mean, var, invstd


class BatchSync:
    """Collectives that make the batch statistics those of ALL ranks' rows when one batch (graph) is
    split over several processes (the 1-D partition): `gather(v)` returns every rank's
    vector stacked in rank order [world, len(v)], `reduce(v)` the element-wise sum over ranks,
    `batch_rows` the global row count.  The defaults are torch.distributed calls on `group`."""

    def __init__(self, batch_rows, world, group=None, gather=None, reduce=None):
        self.batch_rows, self.world, self.group = int(batch_rows), int(world), group
        self._gather, self._reduce = gather, reduce

    def gather(self, v):
        if self._gather is not None:
            return self._gather(v)
        import torch.distributed as dist
        out = v.new_empty(self.world * v.numel())      # flat: gloo rejects a [world, len] output, nccl takes both
        dist.all_gather_into_tensor(out, v.contiguous().view(-1), group=self.group)
        return out.view(self.world, v.numel())

    def reduce(self, v):
        if self._reduce is not None:
            return self._reduce(v)
        import torch.distributed as dist
        dist.all_reduce(v, group=self.group)
        return v


def synced_batch_stats(x, sync: BatchSync):
    """Global (mean, biased variance) from every rank's local ones: combined in rank order on every
    rank (identical results everywhere): mean = sum n_r mean_r / N, var = sum n_r (var_r + (mean_r -
    mean)^2) / N."""
    mean, var = batch_stats(x)
    return merge_batch_stats(mean, var, x.size(0), sync)


def merge_batch_stats(mean, var, rows, sync: BatchSync):
    """The host side of `synced_batch_stats`: this rank's (mean, biased variance, row count) gathered
    and combined (plain tensor arithmetic on [world, F] -- runs on any device, tested over gloo)."""
    f = mean.numel()
    pack = torch.cat([mean, var, mean.new_full((1,), float(rows))])
    g = sync.gather(pack)
    w = (g[:, -1] / float(sync.batch_rows)).unsqueeze(1)
    gmean = (w * g[:, :f]).sum(0)
    gvar = (w * (g[:, f:2 * f] + (g[:, :f] - gmean) ** 2)).sum(0)
    return gmean, gvar


class _BnActDrop(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, mean, invstd, use_batch_stats, relu, p, seed, sync=None):
        x = _rows16(x)
        n, f = x.shape
        out = torch.empty((n, f), dtype=torch.float32, device=x.device)
        with _lib.traced("bn_act_forward", 2 * n * f * 4, 0, 1):
            _lib.check(_lib.lib().stemgnn_bn_act_forward(
                n, f, x.data_ptr(), x.stride(0), mean.data_ptr(), invstd.data_ptr(), gamma.data_ptr(),
                beta.data_ptr(), int(relu), float(p), _lib.ptr(seed), out.data_ptr(), out.stride(0), _lib.stream()))
        ctx.save_for_backward(x, gamma, beta, mean, invstd, seed)
        ctx.cfg = (bool(use_batch_stats), bool(relu), float(p))
        ctx.sync = sync if use_batch_stats else None
        return out

    @staticmethod
    def backward(ctx, gy):
        x, gamma, beta, mean, invstd, seed = ctx.saved_tensors
        use_batch_stats, relu, p = ctx.cfg
        gy = _rows16(gy)
        n, f = x.shape
        sum_g = torch.empty(f, dtype=torch.float32, device=x.device)
        sum_gx = torch.empty_like(sum_g)
        gx = torch.empty((n, f), dtype=torch.float32, device=x.device) if ctx.needs_input_grad[0] else None
        ws = _ws(n, f, x.device)
        L = _lib.lib()
        split = ctx.sync is not None and gx is not None     # batch split over ranks: sums, all-reduce, apply
        with _lib.traced("bn_act_backward", (5 if gx is not None else 2) * n * f * 4, 0, 3):
            _lib.check(L.stemgnn_bn_act_backward(
                n, f, x.data_ptr(), x.stride(0), gy.data_ptr(), gy.stride(0), mean.data_ptr(), invstd.data_ptr(),
                gamma.data_ptr(), beta.data_ptr(), int(relu), p, _lib.ptr(seed), int(use_batch_stats),
                sum_g.data_ptr(), sum_gx.data_ptr(), None if split else _lib.ptr(gx),
                gx.stride(0) if gx is not None and not split else 0, ws.data_ptr(), ws.numel(), _lib.stream()))
            if split:
                tot = ctx.sync.reduce(torch.cat([sum_g, sum_gx]))        # global sums enter g_x only;
                _lib.check(L.stemgnn_bn_act_backward_apply(               # gamma / beta keep this rank's share
                    n, f, x.data_ptr(), x.stride(0), gy.data_ptr(), gy.stride(0), mean.data_ptr(),
                    invstd.data_ptr(), gamma.data_ptr(), beta.data_ptr(), int(relu), p, _lib.ptr(seed),
                    tot[:f].data_ptr(), tot[f:].data_ptr(), ctx.sync.batch_rows, gx.data_ptr(), gx.stride(0),
                    _lib.stream()))
        return (gx, sum_gx if ctx.needs_input_grad[1] else None, sum_g if ctx.needs_input_grad[2] else None,
                None, None, None, None, None, None, None)


def bn_act_drop(z, bn: nn.BatchNorm1d, relu: bool, p: float, training: bool, sync: BatchSync = None):
    """dropout(relu(bn(z))) with bn's own parameters / buffers; `relu=False`, `p=0` give plain BatchNorm.
    `sync`: z holds only this rank's rows of the batch (partitioned graph) -- statistics and the
    backward's column sums are taken over all ranks' rows, like full-batch BatchNorm on one GPU."""
    use_batch = bn.training or bn.running_mean is None
    f = z.size(1)
    if use_batch and bn.training and (sync.batch_rows if sync is not None else z.size(0)) <= 1:
        # what nn.BatchNorm1d (the reference's module, model/encoder.py:313) does for a one-row batch
        raise ValueError(f"Expected more than 1 value per channel when training, got input size {tuple(z.shape)}")
    fused_tail = use_batch and sync is None and bn.training
    if fused_tail:
        # statistics, invstd and the running-statistics update in the two launches of the statistics pass
        with torch.no_grad():
            track = bn.track_running_stats and bn.running_mean is not None
            if track:
                bn.num_batches_tracked.add_(1)
            mean, var, invstd = train_stats(z.detach(), bn.eps, bn.momentum, bn.num_batches_tracked if track else None,
                                            bn.running_mean if track else None, bn.running_var if track else None)
    elif use_batch:
        with torch.no_grad():
            mean, var = synced_batch_stats(z.detach(), sync) if sync is not None else batch_stats(z.detach())
            if bn.training and bn.track_running_stats and bn.running_mean is not None:
                bn.num_batches_tracked.add_(1)
                if bn.momentum is None:
                    m = 1.0 / bn.num_batches_tracked.to(torch.float32)        # cumulative moving average
                else:
                    m = bn.momentum
                n = sync.batch_rows if sync is not None else z.size(0)
                bn.running_mean.mul_(1 - m).add_(mean * m)
                bn.running_var.mul_(1 - m).add_(var * (n / max(n - 1, 1)) * m)
    else:
        mean, var = bn.running_mean, bn.running_var
    if not fused_tail:
        with torch.no_grad():
            invstd = torch.rsqrt(var + bn.eps)
    gamma = bn.weight if bn.weight is not None else torch.ones(f, device=z.device)
    beta = bn.bias if bn.bias is not None else torch.zeros(f, device=z.device)
    p_eff = float(p) if training else 0.0
    seed = None
    if p_eff > 0.0:
        st = _seed_tensor(z.device)
        seed = st.clone()                 # this call's key, kept for the backward
        st.add_(_GOLDEN)                  # device-side advance: a replayed CUDA graph draws new masks
    return _BnActDrop.apply(z, gamma, beta, mean, invstd, use_batch, relu, p_eff, seed, sync)
