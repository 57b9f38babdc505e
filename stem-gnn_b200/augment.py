"""Graph augmentations of the pretrain step and the row gathers that feed it (K9): drop-in
`mask_feature` / `dropout_adj` with the signatures of torch_geometric.utils 2.3.0, which is what
the reference imports (STEM-GNN/pretrain.py:11,41-44), plus `gather_rows` for
`node_text_feat[data.x]` / `edge_text_feat[data.xe]` (pretrain.py:33-38).

The random draws come from `torch.rand` on the tensor's device, exactly where the PyG functions
draw them, so under the same torch seed the masks are the ones the reference would get on that
device.  `generator="philox"` switches dropout_adj to the in-kernel counter-based stream (no E-sized
uniform tensor)."""
import ctypes as C

import torch

from . import _lib


def _check_p(p):
    if p < 0. or p > 1.:
        raise ValueError(f"Dropout probability has to be between 0 and 1 (got {p}")


def gather_rows(src, ids):
    """src[ids] for a 2-D fp32 CUDA matrix and an int64 id vector."""
    if src.dtype != torch.float32 or src.dim() != 2:
        raise TypeError("gather_rows: src must be a 2-D float32 tensor")
    if ids.dtype != torch.int64 or ids.dim() != 1:
        raise TypeError("gather_rows: ids must be a 1-D int64 tensor")
    src, ids = src.contiguous(), ids.contiguous()
    n, f = ids.numel(), src.size(1)
    if f % 4:
        raise ValueError("gather_rows: the feature width must be a multiple of 4")
    out = torch.empty(n, f, dtype=torch.float32, device=src.device)
    with _lib.traced("gather_rows", 2 * n * f * 4 + n * 8, 0, 1):
        _lib.check(_lib.lib().stemgnn_gather_rows(_lib.ptr(src), f, src.size(0), _lib.ptr(ids), n, f,
                                                  _lib.ptr(out), f, _lib.stream()))
    return out


class _GatherRows(torch.autograd.Function):
    @staticmethod
    def forward(ctx, src, ids):
        ctx.save_for_backward(ids)
        ctx.n_src = src.size(0)
        return gather_rows(src, ids)

    @staticmethod
    def backward(ctx, g):
        # index_add_ of the gathered rows' gradients, as a pull: rows grouped by id with the stable
        # sort of K0, summed in position order by K1 (deterministic, no atomics)
        from .graph import CSR
        from .ops import _aggregate_call
        (ids,) = ctx.saved_tensors
        g = g.contiguous()
        pos = torch.arange(ids.numel(), dtype=torch.int64, device=ids.device)
        csr = CSR(torch.stack([pos, ids]), ctx.n_src, 1)
        gs = torch.empty(ctx.n_src, g.size(1), dtype=g.dtype, device=g.device)
        _aggregate_call(csr, g.size(0), g, gs)
        return gs, None


def _seed_tensor(device):
    return torch.randint(0, 2 ** 62, (1,), dtype=torch.int64, device=device)


def dropout_adj(edge_index, edge_attr=None, p=0.5, force_undirected=False, num_nodes=None,
                training=True, generator="torch", return_edge_id=False):
    """torch_geometric.utils.dropout_adj (2.3.0).  Returns (edge_index, edge_attr) -- and the int64
    original-edge-id vector as a third value when `return_edge_id` (how integer edge types follow)."""
    _check_p(p)
    if not training or p == 0.0:
        if return_edge_id:
            return edge_index, edge_attr, torch.arange(edge_index.size(1), device=edge_index.device)
        return edge_index, edge_attr
    if edge_index.dtype != torch.int64 or edge_index.dim() != 2 or edge_index.size(0) != 2:
        raise ValueError("edge_index must be an int64 tensor of shape [2, E]")
    L = _lib.lib()
    ei = edge_index.contiguous()
    dev, E = ei.device, ei.size(1)
    rnd = seed = None
    if generator == "torch":
        rnd = torch.rand(E, device=dev)           # same call, same stream position as PyG's
    elif generator == "philox":
        seed = _seed_tensor(dev)
    else:
        raise ValueError("generator must be 'torch' or 'philox'")
    need = C.c_size_t(0)
    _lib.check(L.stemgnn_dropout_adj_workspace(E, C.byref(need)))
    ws = _lib.workspace(need.value, dev)
    kept_dev = torch.zeros(1, dtype=torch.int64, device=dev)
    fu = int(bool(force_undirected))
    with _lib.traced("dropout_adj_count", E * 20, 0, 2):
        _lib.check(L.stemgnn_dropout_adj_count(_lib.ptr(ei), E, _lib.ptr(rnd), _lib.ptr(seed), p, fu,
                                               _lib.ptr(kept_dev), _lib.ptr(ws), need.value, _lib.stream()))
    kept = int(kept_dev.item())                   # the data-dependent size boolean indexing syncs on too
    m = kept * (2 if fu else 1)
    out = torch.empty(2, m, dtype=torch.int64, device=dev)
    want_id = edge_attr is not None or return_edge_id
    eid = torch.empty(m, dtype=torch.int64, device=dev) if want_id else None
    with _lib.traced("dropout_adj_write", E * 20 + m * (24 if want_id else 16), 0, 1):
        _lib.check(L.stemgnn_dropout_adj_write(_lib.ptr(ei), E, _lib.ptr(rnd), _lib.ptr(seed), p, fu, kept,
                                               _lib.ptr(out), _lib.ptr(eid), _lib.ptr(ws), need.value,
                                               _lib.stream()))
    if edge_attr is not None:
        if edge_attr.dtype == torch.float32 and edge_attr.dim() == 2 and edge_attr.size(1) % 4 == 0:
            edge_attr = _GatherRows.apply(edge_attr, eid)
        else:
            edge_attr = edge_attr[eid]            # integer edge types / odd widths: a plain index
    if return_edge_id:
        return out, edge_attr, eid
    return out, edge_attr


class _MaskFill(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, row_keep, col_keep, p_elem, fill, want_mask):
        L = _lib.lib()
        x = x.contiguous()
        n, f = x.shape
        out = torch.empty_like(x)
        seed = _seed_tensor(x.device) if p_elem > 0 else None
        mask = torch.empty(n, f, dtype=torch.bool, device=x.device) if want_mask else None
        with _lib.traced("mask_fill", 2 * n * f * 4, 0, 1):
            _lib.check(L.stemgnn_mask_fill(n, f, _lib.ptr(x), f, _lib.ptr(row_keep), _lib.ptr(col_keep),
                                           p_elem, _lib.ptr(seed), fill, _lib.ptr(out), f, _lib.ptr(mask),
                                           _lib.stream()))
        ctx.save_for_backward(row_keep, col_keep, seed)
        ctx.p_elem = p_elem
        ctx.mark_non_differentiable(mask) if mask is not None else None
        return (out, mask) if want_mask else out

    @staticmethod
    def backward(ctx, g, *unused):
        row_keep, col_keep, seed = ctx.saved_tensors
        g = g.contiguous()
        n, f = g.shape
        gx = torch.empty_like(g)
        _lib.check(_lib.lib().stemgnn_mask_fill(n, f, _lib.ptr(g), f, _lib.ptr(row_keep), _lib.ptr(col_keep),
                                                ctx.p_elem, _lib.ptr(seed), 0.0, _lib.ptr(gx), f, None,
                                                _lib.stream()))
        return gx, None, None, None, None, None


def mask_feature(x, p=0.5, mode="col", fill_value=0., training=True):
    """torch_geometric.utils.mask_feature (2.3.0): returns (masked x, bool mask)."""
    _check_p(p)
    if not training or p == 0.0:
        return x, torch.ones_like(x, dtype=torch.bool)
    assert mode in ["row", "col", "all"]
    if x.dim() != 2 or x.dtype != torch.float32:
        raise TypeError("mask_feature: x must be a 2-D float32 tensor")
    if not x.is_cuda:
        raise _lib.StemGnnError("mask_feature: x must be a CUDA tensor (no CPU fallback)")
    if x.size(1) % 4:
        raise ValueError("mask_feature: the feature width must be a multiple of 4")
    if mode == "row":
        mask = torch.rand(x.size(0), device=x.device) >= p
        return _MaskFill.apply(x, mask, None, 0.0, float(fill_value), False), mask.view(-1, 1)
    if mode == "col":
        mask = torch.rand(x.size(1), device=x.device) >= p
        return _MaskFill.apply(x, None, mask, 0.0, float(fill_value), False), mask.view(1, -1)
    out, mask = _MaskFill.apply(x, None, None, float(p), float(fill_value), True)
    return out, mask


def negative_sampling(edge_index, num_nodes=None, num_neg_samples=None, method="sparse",
                      force_undirected=False):
    """torch_geometric.utils.negative_sampling (2.3.0) as the reference calls it
    (`negative_sampling(pos_edge_index, z.size(0))`, model/pt_model.py:60): up to `num_neg_samples`
    (default: as many as there are edges) distinct node pairs (i, j), i != j, that are not edges.
    Both PyG methods ("sparse", "dense") map to the one device implementation (K12)."""
    if force_undirected:
        raise NotImplementedError("negative_sampling(force_undirected=True) is outside the hot path")
    assert method in ["sparse", "dense"]
    if edge_index.dtype != torch.int64 or edge_index.dim() != 2 or edge_index.size(0) != 2:
        raise ValueError("edge_index must be an int64 tensor of shape [2, E]")
    if not edge_index.is_cuda:
        raise _lib.StemGnnError("negative_sampling: edge_index must be a CUDA tensor (no CPU fallback)")
    L = _lib.lib()
    ei = edge_index.contiguous()
    dev, E = ei.device, ei.size(1)
    N = int(num_nodes) if num_nodes is not None else (int(ei.max()) + 1 if E else 0)
    population = N * N - N
    want = E if num_neg_samples is None else int(num_neg_samples)
    if population <= 0 or E >= population or want <= 0:
        return ei.new_empty((2, 0))
    want = min(want, population - E)
    prob = 1.0 - E / population
    out = count = None
    for attempt in range(3):                                      # PyG retries three times as well
        m = min(int(1.1 * (1.5 ** attempt) * want / prob) + 16, (1 << 30) - E - 1)
        need = C.c_size_t(0)
        _lib.check(L.stemgnn_negative_sampling_workspace(E, m, C.byref(need)))
        ws = _lib.workspace(need.value, dev)
        out = torch.empty(2, want, dtype=torch.int64, device=dev)
        cnt = torch.zeros(1, dtype=torch.int64, device=dev)
        seed = _seed_tensor(dev)
        with _lib.traced("negative_sampling", E * 16 + m * 24 + want * 16, 0, 6):
            _lib.check(L.stemgnn_negative_sampling(_lib.ptr(ei), E, N, m, want, _lib.ptr(seed), _lib.ptr(out),
                                                   _lib.ptr(cnt), _lib.ptr(ws), need.value, _lib.stream()))
        count = int(cnt.item())
        if count >= want:
            return out
    return out[:, :count]


def edge_concat_linear(z, edge_index, linear):
    """`linear(torch.cat([z[edge_index[0]], z[edge_index[1]]], dim=-1))` -- the topo_sem decoder input of
    model/pt_model.py:80-81 -- without the E x 2D concatenation: two row gathers (K9) feed ONE
    two-source tensor-core GEMM (K5) against the two halves of `linear.weight`."""
    from . import dense
    d = z.size(1)
    w = linear.weight
    if w.size(1) != 2 * d:
        raise ValueError("linear must take 2 * z.size(1) input features")
    a = _GatherRows.apply(z, edge_index[0].contiguous())
    b = _GatherRows.apply(z, edge_index[1].contiguous())
    if z.is_cuda and dense.supported(d, w.size(0)):
        return dense.linear_sum([a, b], [w[:, :d], w[:, d:]], linear.bias)
    return linear(torch.cat([a, b], dim=-1))
