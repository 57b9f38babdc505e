"""Dense layers of the conv on the tensor cores (K5/K6).

`linear_sum(srcs, weights, bias)` = sum_s srcs[s] @ weights[s].T + bias -- what MySAGEConv.forward
computes as `lin_l(out) + lin_r(x_r)` (reference STEM-GNN/model/encoder.py:84-88) -- with its
autograd: the backward-data products run through the same kernel (W given transposed), the
weight gradients through the split-K kernel.  fp32 in / fp32 out, 3xTF32 on tcgen05: within 1e-5
of torch's fp32 result.  No CPU or torch fallback: the C-ABI library must be built.
"""
import ctypes as C
from typing import Optional, Sequence

import torch

from . import _lib


def _check(t, name):
    if not t.is_cuda:
        raise _lib.StemGnnError(f"{name} must be a CUDA tensor (no CPU fallback)")
    if t.dtype != torch.float32:
        raise TypeError(f"{name} must be float32 (got {t.dtype})")
    return t


def _rowmajor(t):
    """A 2-D view whose rows are contiguous and 16-byte aligned; copies only if they are not."""
    if t.stride(1) != 1 or t.stride(0) % 4 or t.data_ptr() % 16 or t.stride(0) < t.size(1):
        t = t.contiguous()
    return t


K_SLAB = 768       # columns accumulated in TMEM per call once a product is cut into slabs (= K_SINGLE: same bound)
K_SINGLE = 768     # widest product done in one call (measured max error 7.5e-9 * K of the output scale)
# Beyond 1,536 columns in total (Cora's 1,436 + 1,436) slabs of 768 leave 1.03e-5 (measured, C1): the truncation of
# a slab's accumulator is relative to the slab's own partial sum, so the total goes like sqrt(slab) -- 256 there.
K_SLAB_WIDE, K_WIDE = 256, 1536


def linear_raw(srcs: Sequence[torch.Tensor], weights: Sequence[torch.Tensor],
               bias: Optional[torch.Tensor] = None, w_transposed: bool = False,
               out: Optional[torch.Tensor] = None, accumulate: bool = False,
               row_div: Optional[torch.Tensor] = None) -> torch.Tensor:
    """out[m, n] = sum_s srcs[s][m, :] . weights[s][n, :] (+ bias[n]); with w_transposed the
    weights are [k, n] and the product is srcs[s] @ weights[s].  One stemgnn_linear call while
    sum(k) <= K_SINGLE; wider products are cut into column slabs of K_SLAB whose partial products are
    added round-to-nearest by the epilogue (accumulate), keeping the 1e-5 bound."""
    ktot_all = sum(int(a.size(1)) for a in srcs)
    if ktot_all > K_SINGLE:
        slab = K_SLAB_WIDE if ktot_all > K_WIDE else K_SLAB
        pieces = []
        for a, w in zip(srcs, weights):
            for lo in range(0, a.size(1), slab):
                hi = min(a.size(1), lo + slab)
                pieces.append((a[:, lo:hi], w[lo:hi] if w_transposed else w[:, lo:hi]))
        pieces.sort(key=lambda pc: -pc[0].size(1))   # full slabs first, the remainders pair up
        calls, cur, cur_k = [], [], 0
        for pc in pieces:                       # at most two sources and `slab` columns per call
            k = pc[0].size(1)
            if cur and (len(cur) == 2 or cur_k + k > slab):
                calls.append(cur); cur, cur_k = [], 0
            cur.append(pc); cur_k += k
        calls.append(cur)
        for i, grp in enumerate(calls):
            out = linear_raw([g[0] for g in grp], [g[1] for g in grp], bias if i == 0 else None,
                             w_transposed, out, accumulate or i > 0)
        if row_div is not None:                 # the division belongs to the finished sum of the slabs
            out = out / row_div.view(-1, 1)
        return out
    ns = len(srcs)
    if ns not in (1, 2) or len(weights) != ns:
        raise ValueError("one or two (source, weight) pairs")
    srcs = [_rowmajor(_check(a, "src")) for a in srcs]
    weights = [_check(w, "weight") for w in weights]
    weights = [w if w.stride(1) == 1 else w.contiguous() for w in weights]
    m = srcs[0].size(0)
    n = weights[0].size(1 if w_transposed else 0)
    for a, w in zip(srcs, weights):
        k = w.size(0 if w_transposed else 1)
        if a.dim() != 2 or a.size(0) != m or a.size(1) != k or w.size(1 if w_transposed else 0) != n:
            raise ValueError("shape mismatch between sources and weights")
        if k % 4:
            raise ValueError("feature widths must be multiples of 4 (pad the layer)")
    if out is None:
        out = torch.empty((m, n), dtype=torch.float32, device=srcs[0].device)
    a = _lib.LinearArgs()
    a.m, a.n, a.n_src, a.w_transposed = m, n, ns, int(w_transposed)
    for s in range(ns):
        a.a[s], a.lda[s], a.k[s] = srcs[s].data_ptr(), srcs[s].stride(0), srcs[s].size(1)
        a.w[s], a.ldw[s] = weights[s].data_ptr(), weights[s].stride(0)
    a.bias = _lib.ptr(_check(bias, "bias").contiguous()) if bias is not None else None
    a.out, a.ldo = out.data_ptr(), out.stride(0)
    a.accumulate = int(accumulate)
    if row_div is not None:
        if row_div.numel() != m or row_div.dtype != torch.float32 or not row_div.is_contiguous():
            raise ValueError("row_div must be a contiguous float32 vector with one entry per row")
        a.row_div = row_div.data_ptr()
    L = _lib.lib()
    need = C.c_size_t(0)
    _lib.check(L.stemgnn_linear_workspace(C.byref(a), C.byref(need)))
    ws = _lib.workspace(need.value, out.device)
    ktot = sum(int(x.size(1)) for x in srcs)
    nbytes = m * ktot * 4 + m * n * 4 + n * ktot * 4
    with _lib.traced("linear", nbytes, 2 * m * n * ktot, 2):
        _lib.check(L.stemgnn_linear(C.byref(a), ws.data_ptr(), ws.numel(), _lib.stream()))
    return out


WGRAD_MAX_TILES = 64     # 256 x 256 weight tiles one stemgnn_linear_wgrad call takes (include/stemgnn_b200.h)


def linear_wgrad_raw(g: torch.Tensor, srcs: Sequence[torch.Tensor], want_bias: bool = False):
    """([g.T @ srcs[s] for s], g.sum(0) or None) in one stemgnn_linear_wgrad call; weight matrices of more than
    64 tiles of 256 x 256 are cut into blocks of output features (column slices of g), one call each."""
    ns = len(srcs)
    if ns not in (1, 2):
        raise ValueError("one or two sources")
    g = _rowmajor(_check(g, "grad"))
    srcs = [_rowmajor(_check(a, "src")) for a in srcs]
    m, n = g.shape
    kbt = -(-sum(int(a.size(1)) for a in srcs) // 256)
    if -(-n // 256) * kbt > WGRAD_MAX_TILES:
        if kbt > WGRAD_MAX_TILES:
            raise ValueError(f"sum of the source widths above {WGRAD_MAX_TILES * 256} columns")
        nb = (WGRAD_MAX_TILES // kbt) * 256
        parts = [linear_wgrad_raw(g[:, lo:min(n, lo + nb)], srcs, want_bias) for lo in range(0, n, nb)]
        gws = [torch.cat([p[0][s] for p in parts], dim=0) for s in range(ns)]
        return gws, (torch.cat([p[1] for p in parts]) if want_bias else None)
    if n % 4 or any(a.size(1) % 4 for a in srcs):
        raise ValueError("feature widths must be multiples of 4 (pad the layer)")
    if any(a.size(0) != m for a in srcs):
        raise ValueError("row counts differ")
    gws = [torch.empty((n, a.size(1)), dtype=torch.float32, device=g.device) for a in srcs]
    gb = torch.empty(n, dtype=torch.float32, device=g.device) if want_bias else None
    a = _lib.LinearWgradArgs()
    a.m, a.n, a.n_src = m, n, ns
    a.g, a.ldg = g.data_ptr(), g.stride(0)
    for s in range(ns):
        a.a[s], a.lda[s], a.k[s] = srcs[s].data_ptr(), srcs[s].stride(0), srcs[s].size(1)
        a.gw[s], a.ldgw[s] = gws[s].data_ptr(), gws[s].stride(0)
    a.gbias = _lib.ptr(gb)
    L = _lib.lib()
    need = C.c_size_t(0)
    _lib.check(L.stemgnn_linear_wgrad_workspace(C.byref(a), C.byref(need)))
    ws = _lib.workspace(need.value, g.device)
    ktot = sum(int(x.size(1)) for x in srcs)
    nbytes = ns * m * n * 4 + m * ktot * 4 + n * ktot * 4
    with _lib.traced("linear_wgrad", nbytes, 2 * m * n * ktot, 2 + (2 if want_bias else 0)):
        _lib.check(L.stemgnn_linear_wgrad(C.byref(a), ws.data_ptr(), ws.numel(), _lib.stream()))
    return gws, gb


class _LinearSum(torch.autograd.Function):
    """sum_s x_s @ W_s.T + b with the three gradient products on the tensor cores."""

    @staticmethod
    def forward(ctx, bias, *xw):
        ns = len(xw) // 2
        xs, ws = list(xw[:ns]), list(xw[ns:])
        ctx.save_for_backward(*xs, *ws)
        ctx.ns, ctx.has_bias = ns, bias is not None
        return linear_raw(xs, ws, bias)

    @staticmethod
    def backward(ctx, g):
        ns = ctx.ns
        saved = ctx.saved_tensors
        xs, ws = list(saved[:ns]), list(saved[ns:])
        g = _rowmajor(g)
        need_x = ctx.needs_input_grad[1:1 + ns]
        need_w = ctx.needs_input_grad[1 + ns:1 + 2 * ns]
        gxs = [linear_raw([g], [w], None, w_transposed=True) if nx else None for w, nx in zip(ws, need_x)]
        gws, gb = [None] * ns, None
        if any(need_w) or (ctx.has_bias and ctx.needs_input_grad[0]):
            gws, gb = linear_wgrad_raw(g, xs, want_bias=ctx.has_bias and ctx.needs_input_grad[0])
            gws = [gw if nw else None for gw, nw in zip(gws, need_w)]
        return (gb, *gxs, *gws)


def pad4(t: torch.Tensor, rows: bool = False) -> torch.Tensor:
    """Zero-pad the last dimension (and, with `rows`, the first) of a matrix / vector to a multiple of 4: the
    kernels read 16-byte row segments.  Differentiable (the gradient of the padding is dropped)."""
    rk = (-t.size(-1)) % 4
    rn = (-t.size(0)) % 4 if rows and t.dim() == 2 else 0
    if not (rk or rn):
        return t
    return torch.nn.functional.pad(t, (0, rk, 0, rn) if t.dim() == 2 else (0, rk))


def linear_sum(srcs: Sequence[torch.Tensor], weights: Sequence[torch.Tensor],
               bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """sum_s srcs[s] @ weights[s].T + bias, differentiable w.r.t. everything (1 or 2 sources).  Widths that are
    not multiples of 4 (Cora's 1,433 input features, BASELINE configs[0]) are zero-padded -- sources by columns,
    weights by columns and output rows -- so they run the same tensor-core kernels; the padded output columns are
    sliced off again (no torch GEMM fallback anywhere)."""
    n = weights[0].size(0)
    srcs = [pad4(a) for a in srcs]
    weights = [pad4(w, rows=True) for w in weights]
    if bias is not None:
        bias = pad4(bias)
    out = _LinearSum.apply(bias, *srcs, *weights)
    return out[:, :n] if out.size(1) != n else out


def supported(*widths: int) -> bool:
    """Layer widths the tensor-core path takes: any positive width (odd ones are padded by `linear_sum`)."""
    return all(w > 0 for w in widths)
