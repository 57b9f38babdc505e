"""One SAGE layer as ONE autograd node: out = lin_l(reduce_{j->i} relu(x_j)) + lin_r(x_i) (+ bias)
(reference STEM-GNN/model/encoder.py:72-97, the plain-message case: no edge features).

Composed of separate nodes (ops.aggregate -> dense.linear_sum) the backward pays two passes over [N, F]
matrices that carry no arithmetic of their own, and on a partitioned graph it cannot start the halo exchange
before the whole dense backward has run:

  * MeanAggregation's `g_agg / count[dst]` (K2a, stemgnn_row_scale) can be done by the epilogue of the
    backward-data product g @ W_l that produces g_agg (stemgnn_linear, row_div; off by default, see below);
  * autograd's accumulation of the two gradients of x (the aggregated path and the lin_r path): here the
    aggregation's epilogue adds the lin_r term to the finished row (stemgnn_aggregate, row_add);
  * on a PartitionedGraph the exchange of g_agg is started as soon as g_agg exists, and the weight-gradient
    product and g @ W_r run while the blocks are on the wire; in the overlapped mode (halo_mode="blocks") the
    forward likewise computes lin_r(x) + bias during the exchange of x and adds lin_l(agg) afterwards.

Same arithmetic per element as the composed form (one rounding per sum, IEEE division), so single-GPU results are
bit-identical to it; the split forward of the overlapped mode adds the two products in another order (that mode's
bar is 1e-5 anyway, partition.py).
"""
import os

import torch

from . import _lib, dense, ops
from .partition import PartitionedGraph

# The count division inside the epilogue of g @ W_l (stemgnn_linear row_div) saves the K2a pass (0.058 ms on the
# arxiv-shaped step) but __fdiv_rn per element makes the GEMM epilogue-bound: 0.543 -> 0.603 ms for the step's four
# K5 calls (profiles/r02_summary.md) -- a wash, so the separate pass stays the default; =1 switches the epilogue on.
_ROW_DIV_IN_EPILOGUE = os.environ.get("STEMGNN_ROW_DIV_EPILOGUE", "0") == "1"


class _beside_collective:
    """Dense products launched while a halo exchange is in flight: K5 / K6 are persistent one-CTA-per-SM kernels
    (~200 KB of shared memory each), and the SMs NCCL's copy CTAs sit on cannot take such a CTA, so with a grid of
    all SMs a statically dealt share of the tiles would wait for the collective to END (8 B200s: dense 3.5 -> 5.4 ms
    per step).  Inside this context the grids leave `pg.sm_reserve` SMs (NCCL runs up to 32 channels here) free."""

    def __init__(self, pg):
        self.n = int(getattr(pg, "sm_reserve", 32)) if (isinstance(pg, PartitionedGraph) and pg.world > 1
                                                        and pg._blocks is None and pg._gather is None) else 0

    def __enter__(self):
        if self.n:
            self.prev = _lib.lib().stemgnn_set_sm_reserve(self.n)

    def __exit__(self, *exc):
        if self.n:
            _lib.lib().stemgnn_set_sm_reserve(self.prev)
        return False


class _SageLayer(torch.autograd.Function):

    @staticmethod
    def forward(ctx, x, graph, reduce, w_l, w_r, bias):
        x = ops._f32c(x, "x")
        part = isinstance(graph, PartitionedGraph)
        mean = reduce == ops.REDUCE_MEAN
        agg = torch.empty_like(x)
        out = None
        if part:
            if x.size(0) != graph.num_nodes:
                raise ValueError(f"x must hold this rank's {graph.num_nodes} rows")
            fwd = graph.fwd_csr()
            ex = ops.HaloExchange(graph, x, False)
            if graph.halo_mode != "gather" and graph.world > 1 and w_r is not None:
                with _beside_collective(graph):
                    out = dense.linear_raw([x], [w_r], bias)        # while the blocks of x are on the wire
            ex.aggregate(agg, fwd.cnt if mean else None, None, ops.ACT_RELU)
        else:
            if x.size(0) != graph.num_nodes:
                raise ValueError("x must be [num_nodes, F]")
            fwd = graph.grouped_by(1)
            ops._aggregate_call(fwd, x.size(0), x, agg, row_cnt=fwd.cnt if mean else None, act=ops.ACT_RELU)
        if out is not None:
            out = dense.linear_raw([agg], [w_l], None, out=out, accumulate=True)
        elif w_r is not None:
            out = dense.linear_raw([agg, x], [w_l, w_r], bias)
        else:
            out = dense.linear_raw([agg], [w_l], bias)
        ctx.graph, ctx.mean, ctx.part = graph, mean, part
        ctx.has_r, ctx.has_bias = w_r is not None, bias is not None
        ctx.save_for_backward(x, agg, w_l, w_r)
        return out

    @staticmethod
    def backward(ctx, g):
        x, agg, w_l, w_r = ctx.saved_tensors
        graph, part = ctx.graph, ctx.part
        g = dense._rowmajor(ops._f32c(g, "grad"))
        need_x, need_wl, need_wr, need_b = (ctx.needs_input_grad[0], ctx.needs_input_grad[3],
                                            ctx.has_r and ctx.needs_input_grad[4],
                                            ctx.has_bias and ctx.needs_input_grad[5])
        g_x = ex = g_agg = None
        if need_x:
            cnt = (graph.fwd_csr() if part else graph.grouped_by(1)).cnt if ctx.mean else None
            if _ROW_DIV_IN_EPILOGUE:
                g_agg = dense.linear_raw([g], [w_l], None, w_transposed=True, row_div=cnt)  # g @ W_l / count[dst]
            else:
                g_agg = dense.linear_raw([g], [w_l], None, w_transposed=True)
                if cnt is not None:
                    g_agg = ops._row_scale(g_agg, cnt)
            if part:
                ex = ops.HaloExchange(graph, g_agg, True)       # on the wire while the products below run
        gws, gb = [None, None], None
        with _beside_collective(graph if ex is not None else None):
            if need_wl or need_wr or need_b:
                srcs = [agg, x] if ctx.has_r else [agg]
                gws, gb = dense.linear_wgrad_raw(g, srcs, want_bias=need_b)
                gws = list(gws) + [None] * (2 - len(gws))
            if need_x:
                g_x = dense.linear_raw([g], [w_r], None, w_transposed=True) if ctx.has_r else torch.empty_like(x)
        if need_x:
            add = g_x if ctx.has_r else None
            if part:
                if graph.halo_mode != "gather" and add is not None:
                    g_x = torch.empty_like(x)       # the chained block calls build their sum in `out`: no aliasing
                ex.aggregate(g_x, None, x, ops.ACT_RELU, row_add=add)
            else:
                ops._aggregate_call(graph.grouped_by(0), x.size(0), g_agg, g_x, row_x=x, act=ops.ACT_RELU,
                                    backward=True, row_add=add)
        return (g_x, None, None, gws[0] if need_wl else None, gws[1] if need_wr else None, gb)


def sage_layer(x, graph, reduce, lin_l, lin_r):
    """lin_l(aggregate(relu(x_j))) + lin_r(x) for a GraphStructure or PartitionedGraph, plain messages, widths that
    are multiples of 4 and at most dense.K_SINGLE columns in total (wider layers take the composed form, whose K
    slabs chain several GEMM calls)."""
    r = {"mean": ops.REDUCE_MEAN, "sum": ops.REDUCE_SUM, "add": ops.REDUCE_SUM}[reduce]
    if torch.is_grad_enabled() and not isinstance(graph, PartitionedGraph):
        graph.prepare_training()
    return _SageLayer.apply(x, graph, r, lin_l.weight, lin_r.weight if lin_r is not None else None, lin_l.bias)


def fusable(x, graph, edge_attr, out_channels, root_weight):
    f = x.size(1)
    k = 2 * f if root_weight else f
    return (edge_attr is None and x.is_cuda and x.dtype == torch.float32 and f % 4 == 0 and out_channels % 4 == 0
            and k <= dense.K_SINGLE and out_channels <= dense.K_SINGLE and not getattr(graph, "fused", False))
