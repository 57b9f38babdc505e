"""Edge-wise inner product of InnerProductDecoder on the GPU (K8).

`edge_dot(z, edge_index, sigmoid)` = `(z[edge_index[0]] * z[edge_index[1]]).sum(dim=1)` (+ sigmoid),
the decode step of the reference's topology reconstruction loss (STEM-GNN/model/encoder.py:363-366,
model/pt_model.py:62-63), without the two E x F gathered matrices.  Its autograd is two weighted
sum-aggregations (K1) over the same edge list, weights = upstream gradient per edge:
    g_z[i] = sum_{e: dst(e)=i} g_e z[src(e)] + sum_{e: src(e)=i} g_e z[dst(e)]
in edge order per row (deterministic, no atomics).  No CPU path.
"""
import torch

from . import _lib
from .graph import structure_of


class _EdgeDot(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z, edge_index, sigmoid):
        if not z.is_cuda:
            raise _lib.StemGnnError("z must be a CUDA tensor (no CPU fallback)")
        if z.dtype != torch.float32 or z.dim() != 2:
            raise TypeError("z must be a float32 [N, F] tensor")
        if edge_index.dtype != torch.int64 or edge_index.dim() != 2 or edge_index.size(0) != 2:
            raise ValueError("edge_index must be an int64 tensor of shape [2, E]")
        if not edge_index.is_cuda:
            raise _lib.StemGnnError("edge_index must be a CUDA tensor (no CPU fallback)")
        if z.stride(1) != 1 or z.stride(0) < z.size(1):
            z = z.contiguous()
        ei = edge_index.contiguous()
        n, f = z.shape
        e = ei.size(1)
        out = torch.empty(e, dtype=torch.float32, device=z.device)
        with _lib.traced("edge_dot", 2 * e * f * 4 + e * 20, 2 * e * f, 1):
            _lib.check(_lib.lib().stemgnn_edge_dot(_lib.ptr(ei), e, n, f, _lib.ptr(z), z.stride(0), int(sigmoid),
                                                   _lib.ptr(out), _lib.stream()))
        ctx.save_for_backward(z, ei, out if sigmoid else None)
        ctx.sigmoid = bool(sigmoid)
        return out

    @staticmethod
    def backward(ctx, g):
        from .ops import _aggregate_call
        z, ei, y = ctx.saved_tensors
        if not ctx.needs_input_grad[0]:
            return None, None, None
        w = (g * y * (1 - y)) if ctx.sigmoid else g
        w = w.contiguous()
        zc = z.contiguous()
        n, f = zc.shape
        gs = structure_of(ei, n)
        g_dst = torch.empty((n, f), dtype=torch.float32, device=z.device)     # d/dz[dst]: gathers z[src]
        g_src = torch.empty_like(g_dst)                                        # d/dz[src]: gathers z[dst]
        _aggregate_call(gs.grouped_by(1), n, zc, g_dst, edge_weight=w)
        _aggregate_call(gs.grouped_by(0), n, zc, g_src, edge_weight=w)
        return g_dst.add_(g_src), None, None


def edge_dot(z: torch.Tensor, edge_index: torch.Tensor, sigmoid: bool = False) -> torch.Tensor:
    """[E] inner products of the end-point rows of every edge (differentiable w.r.t. z)."""
    return _EdgeDot.apply(z, edge_index, sigmoid)
