"""K7 (fused BatchNorm1d + ReLU + dropout) against torch's own modules on the same nn.BatchNorm1d."""
import copy

import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

import stemgnn_b200 as S

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _close(a, b, tol=1e-5):
    scale = b.abs().max().clamp_min(1e-30)
    assert ((a - b).abs().max() / scale).item() < tol, ((a - b).abs().max() / scale).item()


@pytest.mark.parametrize("n,f", [(1000, 256), (777, 36), (5, 4), (40000, 128), (3000, 520)])
@pytest.mark.parametrize("relu", [True, False])
def test_training_forward_backward_and_running_stats_match_torch(n, f, relu):
    g = torch.Generator().manual_seed(n + f)
    x = (torch.randn(n, f, generator=g) * 3 + 5).to(DEV)       # mean far from 0: the shifted sums matter
    go = torch.randn(n, f, generator=g).to(DEV)
    bn_a = nn.BatchNorm1d(f).to(DEV).train()
    with torch.no_grad():
        bn_a.weight.copy_(torch.rand(f, generator=g) + 0.5); bn_a.bias.copy_(torch.randn(f, generator=g))
    bn_b = copy.deepcopy(bn_a)
    xa = x.clone().requires_grad_(True); xb = x.clone().requires_grad_(True)
    ya = S.bn_act_drop(xa, bn_a, relu=relu, p=0.0, training=True)
    yb = bn_b(xb)
    yb = F.relu(yb) if relu else yb
    ya.backward(go); yb.backward(go)
    _close(ya.detach(), yb.detach())
    _close(xa.grad, xb.grad)
    _close(bn_a.weight.grad, bn_b.weight.grad); _close(bn_a.bias.grad, bn_b.bias.grad)
    _close(bn_a.running_mean, bn_b.running_mean); _close(bn_a.running_var, bn_b.running_var)
    assert int(bn_a.num_batches_tracked) == int(bn_b.num_batches_tracked) == 1


def test_eval_mode_uses_running_statistics():
    g = torch.Generator().manual_seed(1)
    n, f = 500, 64
    x = torch.randn(n, f, generator=g).to(DEV)
    bn = nn.BatchNorm1d(f).to(DEV)
    with torch.no_grad():
        bn.running_mean.copy_(torch.randn(f, generator=g)); bn.running_var.copy_(torch.rand(f, generator=g) + 0.3)
    bn.eval()
    xa = x.clone().requires_grad_(True); xb = x.clone().requires_grad_(True)
    ya = S.bn_act_drop(xa, bn, relu=True, p=0.5, training=False)     # dropout is off in eval
    yb = F.relu(bn(xb))
    _close(ya.detach(), yb.detach())
    ya.sum().backward(); yb.sum().backward()
    _close(xa.grad, xb.grad)
    assert int(bn.num_batches_tracked) == 0


def test_dropout_mask_statistics_and_backward_consistency():
    n, f, p = 4096, 256, 0.15
    x = torch.randn(n, f, device=DEV)
    bn = nn.BatchNorm1d(f).to(DEV).train()
    S.norm_act.manual_seed(1234)
    xa = x.clone().requires_grad_(True)
    y = S.bn_act_drop(xa, bn, relu=False, p=p, training=True)
    ref = F.batch_norm(x, None, None, bn.weight, bn.bias, True, 0.0, bn.eps)
    kept = y != 0
    frac = kept.float().mean().item()
    assert abs(frac - (1 - p)) < 0.005, frac
    _close(y[kept], ref[kept] / (1 - p))                      # kept entries are scaled by 1/(1-p)
    # rows and columns are hit evenly
    assert (kept.float().mean(0) - (1 - p)).abs().max().item() < 0.03
    assert (kept.float().mean(1) - (1 - p)).abs().max().item() < 0.12
    # the backward recomputes the SAME mask: d sum(y) / dx is zero through the dropped entries
    go = torch.randn(n, f, device=DEV)
    y.backward(go)
    xb = x.clone().requires_grad_(True)
    yb = F.batch_norm(xb, None, None, bn.weight, bn.bias, True, 0.0, bn.eps)
    yb.backward(go * kept / (1 - p))
    _close(xa.grad, xb.grad)
    # a second call draws a different mask; re-seeding reproduces the first
    y2 = S.bn_act_drop(x, bn, relu=False, p=p, training=True)
    assert not torch.equal(y2 != 0, kept)
    S.norm_act.manual_seed(1234)
    y3 = S.bn_act_drop(x, bn, relu=False, p=p, training=True)
    assert torch.equal(y3 != 0, kept)


def test_encoder_uses_fused_norm_and_matches_unfused_modules():
    g = torch.Generator().manual_seed(8)
    n, e, fin, hid = 2000, 12000, 64, 128
    ei = torch.randint(0, n, (2, e), generator=g).to(DEV)
    x = torch.randn(n, fin, generator=g).to(DEV)
    torch.manual_seed(0)
    enc = S.Encoder(fin, hid, nn.ReLU, 2, backbone="sage", normalize="batch", dropout=0.0).to(DEV).train()
    ref = copy.deepcopy(enc)
    S._lib.TRACE = []
    try:
        z = enc(x, ei)
        names = [r[0] for r in S._lib.TRACE]
    finally:
        S._lib.TRACE = None
    assert names.count("bn_act_forward") == 2
    # unfused: the same layers, torch BatchNorm / ReLU
    with torch.no_grad():
        h = ref.layers[0](x, ei); h = F.relu(ref.norms[0](h)); h = ref.norms[1](ref.layers[1](h, ei))
    _close(z.detach(), h, 2e-5)
    _close(enc.norms[0].running_var, ref.norms[0].running_var)


@pytest.mark.parametrize("world", [1, 2, 3])
@pytest.mark.parametrize("relu", [False, True])
def test_batch_split_over_ranks_equals_full_batch(world, relu):
    """SyncBN for the 1-D partition: every rank holds a block of rows; statistics are
    combined from all ranks' (mean, var, n), the backward's column sums are all-reduced.  The ranks
    run one after the other with the collectives replaced by their known results; outputs and
    gradients must equal the full-batch op's rows."""
    from stemgnn_b200.norm_act import BatchSync, batch_stats, bn_act_drop
    n, f = 1201, 64
    g_ = torch.Generator().manual_seed(world)
    z = (torch.randn(n, f, generator=g_) * 2 + 3).to(DEV)
    up = torch.randn(n, f, generator=g_).to(DEV)
    bn = torch.nn.BatchNorm1d(f).to(DEV).train()
    with torch.no_grad():
        bn.weight.uniform_(0.5, 1.5); bn.bias.uniform_(-0.5, 0.5)
    zf = z.clone().requires_grad_(True)
    full = bn_act_drop(zf, bn, relu, 0.0, True)
    full.backward(up)
    sum_g, sum_gx = bn.bias.grad.clone(), bn.weight.grad.clone()
    run_mean, run_var = bn.running_mean.clone(), bn.running_var.clone()
    bounds = [(r * n // world, (r + 1) * n // world) for r in range(world)]
    packs = torch.stack([torch.cat([*batch_stats(z[lo:hi]), torch.tensor([float(hi - lo)], device=DEV)])
                         for lo, hi in bounds])
    acc_g, acc_gx = torch.zeros(f, device=DEV), torch.zeros(f, device=DEV)
    for lo, hi in bounds:
        bn2 = torch.nn.BatchNorm1d(f).to(DEV).train()
        bn2.load_state_dict({k: (torch.zeros_like(v) if "running_mean" in k else torch.ones_like(v) if "running_var" in k
                                 else v.clone() if "num_batches" not in k else torch.zeros_like(v))
                             for k, v in bn.state_dict().items()})
        sync = BatchSync(n, world, gather=lambda v: packs, reduce=lambda v: torch.cat([sum_g, sum_gx]))
        zl = z[lo:hi].clone().requires_grad_(True)
        out = bn_act_drop(zl, bn2, relu, 0.0, True, sync=sync)
        out.backward(up[lo:hi])
        torch.testing.assert_close(out, full[lo:hi], rtol=1e-5, atol=1e-5)
        torch.testing.assert_close(zl.grad, zf.grad[lo:hi], rtol=1e-5, atol=1e-5)
        torch.testing.assert_close(bn2.running_mean, run_mean, rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(bn2.running_var, run_var, rtol=1e-5, atol=1e-6)
        acc_g += bn2.bias.grad; acc_gx += bn2.weight.grad     # parameter gradients are this rank's share
    torch.testing.assert_close(acc_g, sum_g, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(acc_gx, sum_gx, rtol=1e-4, atol=1e-4)


def test_one_row_batch_in_training_raises_like_torch():
    bn = nn.BatchNorm1d(8).to(DEV).train()
    x = torch.randn(1, 8, device=DEV)
    with pytest.raises(ValueError, match="more than 1 value per channel"):
        bn(x)                                                  # torch's own behaviour
    with pytest.raises(ValueError, match="more than 1 value per channel"):
        S.bn_act_drop(x, bn, relu=True, p=0.0, training=True)
    bn.eval()
    S.bn_act_drop(x, bn, relu=True, p=0.0, training=False)    # running statistics: fine


@pytest.mark.parametrize("momentum", [0.1, 0.3, None])
def test_running_statistics_over_several_batches_follow_torch(momentum):
    """stemgnn_bn_train_stats: invstd and the running-statistics update (momentum m, or the cumulative average of
    momentum=None driven by the module's device counter) done by the statistics kernel's tail, three batches in a
    row, against nn.BatchNorm1d itself."""
    f = 64
    g = torch.Generator().manual_seed(9)
    bn_a = nn.BatchNorm1d(f, momentum=momentum).to(DEV).train()
    bn_b = copy.deepcopy(bn_a)
    for step in range(3):
        x = (torch.randn(500 + 37 * step, f, generator=g) * (1 + step) + step).to(DEV)
        ya = S.bn_act_drop(x, bn_a, relu=False, p=0.0, training=True)
        yb = bn_b(x)
        _close(ya, yb, 1e-5)
        assert int(bn_a.num_batches_tracked) == int(bn_b.num_batches_tracked) == step + 1
        _close(bn_a.running_mean, bn_b.running_mean, 1e-5)      # the batch mean itself carries fp32 summation noise
        _close(bn_a.running_var, bn_b.running_var, 1e-5)
    bn_c = nn.BatchNorm1d(f, track_running_stats=False).to(DEV).train()      # no running statistics at all
    _close(S.bn_act_drop(x, bn_c, relu=False, p=0.0, training=True), bn_c(x), 1e-5)
