"""Autograd operators over the C-ABI: aggregation (K1 forward / K2 backward) and the VQ
assignment / statistics / EMA calls (K3, K4)."""
import ctypes as C
import os

import torch

from . import _lib
from .graph import GraphStructure
from .partition import PartitionedGraph

REDUCE_SUM, REDUCE_MEAN = 0, 1
L2_BYTES = 126e6
ACT_NONE, ACT_RELU = 0, 1


def _f32c(t, name):
    if t is None:
        return None
    if not t.is_cuda:
        raise _lib.StemGnnError(f"{name} must be a CUDA tensor (no CPU fallback)")
    if t.dtype != torch.float32:
        raise TypeError(f"{name} must be float32 (got {t.dtype})")
    return t.contiguous()


def _aggregate_call(csr, n_gather_rows, gathered, out, *, row_cnt=None, gather_cnt=None,
                    edge_feat=None, type_table=None, edge_type=None, edge_weight=None, row_x=None,
                    act=ACT_NONE, backward=False, out_cols=None, gathered_rows=None, accumulate=False,
                    row_add=None):
    """One stemgnn_aggregate call.  `out_cols=(lo, hi)` writes columns [lo, hi) of the wider `out`
    (and reads the same columns of `row_x`): the feature-sliced halo pipeline."""
    per_edge = edge_feat is not None or edge_type is not None or edge_weight is not None
    a = _lib.AggregateArgs(
        rowptr=_lib.ptr(csr.rowptr), col=_lib.ptr(csr.col),
        perm=_lib.ptr(csr.perm) if per_edge else None,
        n_rows=csr.n, n_gather_rows=n_gather_rows, num_edges=csr.e, feat=gathered.size(1),
        gathered=_lib.ptr(gathered), row_cnt=_lib.ptr(row_cnt), gather_cnt=_lib.ptr(gather_cnt),
        edge_feat=_lib.ptr(edge_feat), type_table=_lib.ptr(type_table),
        edge_type=_lib.ptr(edge_type), edge_weight=_lib.ptr(edge_weight), row_x=_lib.ptr(row_x),
        act=act, backward=int(backward), out=_lib.ptr(out), accumulate=int(accumulate))
    if row_add is not None:                      # out[r] = finish(sum) + row_add[r]; may alias out
        a.row_add, a.row_add_ld = _lib.ptr(row_add), row_add.stride(0)
        if out_cols is not None:
            a.row_add = row_add.data_ptr() + 4 * out_cols[0]
    hub = getattr(csr, "hub", None)
    if hub is not None and os.environ.get("STEMGNN_AGG_NO_HUB") != "1":      # A/B switch for scripts/agg_bench.py
        a.hub_ws, a.hub_ws_bytes = hub.data_ptr(), hub.numel()
    if out_cols is not None:
        lo, hi = out_cols
        if hi - lo != gathered.size(1) or lo % 4:
            raise ValueError("column slice must match the gathered width and start on a 16-byte boundary")
        a.out = out.data_ptr() + 4 * lo
        a.out_ld = out.stride(0)
        if row_x is not None:
            a.row_x = row_x.data_ptr() + 4 * lo
            a.row_x_ld = row_x.stride(0)
    # algorithmic bytes: every row read once (X fits the 126 MB L2) or one source
    # row per edge (it does not); index, count and per-edge streams once
    n_g, F, E = n_gather_rows, gathered.size(1), csr.e
    idx = E * 4 + (csr.n + 1) * 4 + (csr.n * 4 if (row_cnt is not None or gather_cnt is not None) else 0)
    if per_edge:
        idx += E * 4 + (E * F * 4 if edge_feat is not None else 0) + (E * 4 if edge_type is not None or edge_weight is not None else 0)
    rows = csr.n * F * 4 * (2 if row_x is not None else 1)
    b_min, b_gather = n_g * F * 4 + rows + idx, E * F * 4 + rows + idx
    # the compulsory model while the gathered matrix fits the 126 MB L2, one
    # source row per edge once it does not
    b_alg = b_min if n_g * F * 4 <= L2_BYTES else b_gather
    name = "aggregate_bwd" if backward else "aggregate_fwd"
    with _lib.traced(name, (b_min, b_gather, b_alg), 0, 1):
        _lib.check(_lib.lib().stemgnn_aggregate(C.byref(a), _lib.stream()))


def _row_scale(g, cnt):
    """g[r, :] / cnt[r] (K2a): MeanAggregation's backward division, once per destination row."""
    out = torch.empty_like(g)
    n, f = g.shape
    with _lib.traced("row_scale", 2 * n * f * 4 + n * 4, 0, 1):
        _lib.check(_lib.lib().stemgnn_row_scale(n, f, _lib.ptr(g), _lib.ptr(cnt), _lib.ptr(out), _lib.stream()))
    return out


class _Aggregate(torch.autograd.Function):
    """out[i] = reduce_{e: to(e)=i} act(x[from(e)] + xe[e]) [* w[e]]  and its exact autograd.
    `to_row` is the edge_index row aggregated AT (1 for PyG source_to_target)."""

    @staticmethod
    def forward(ctx, x, graph, to_row, reduce, act, edge_feat, type_table, edge_type, edge_weight):
        x = _f32c(x, "x")
        if x.dim() != 2 or x.size(0) != graph.num_nodes:
            raise ValueError("x must be [num_nodes, F]")
        fwd = graph.grouped_by(to_row)
        out = torch.empty_like(x)
        _aggregate_call(fwd, x.size(0), x, out,
                        row_cnt=fwd.cnt if reduce == REDUCE_MEAN else None,
                        edge_feat=edge_feat, type_table=type_table, edge_type=edge_type,
                        edge_weight=edge_weight, act=act)
        ctx.graph, ctx.to_row, ctx.reduce, ctx.act = graph, to_row, reduce, act
        ctx.save_for_backward(x, edge_feat, type_table, edge_type, edge_weight)
        return out

    @staticmethod
    def backward(ctx, g_out):
        x, edge_feat, type_table, edge_type, edge_weight = ctx.saved_tensors
        g_out = _f32c(g_out, "grad")
        graph = ctx.graph
        bwd = graph.grouped_by(1 - ctx.to_row)
        g_x = torch.empty_like(x)
        if ctx.reduce == REDUCE_MEAN:
            g_out = _row_scale(g_out, graph.grouped_by(ctx.to_row).cnt)
        _aggregate_call(bwd, x.size(0), g_out, g_x,
                        edge_feat=edge_feat, type_table=type_table, edge_type=edge_type,
                        edge_weight=edge_weight, row_x=x if ctx.act == ACT_RELU else None,
                        act=ctx.act, backward=True)
        return g_x, None, None, None, None, None, None, None, None


class HaloExchange:
    """One halo exchange of a partitioned graph, cut in two so that a caller can put other work between the
    start of the transfers and the aggregation that consumes them (sage.py: the dense products of the layer
    that do not need the aggregate).  Constructing it issues every transfer on the collective's stream;
    `aggregate` waits (stream-ordered) and runs K1 / K2 on the owned rows."""

    def __init__(self, pg, local, backward):
        self.pg, self.backward, self.local = pg, backward, local
        if pg.halo_mode in ("blocks", "chunks"):
            self.blocks = pg.exchange_groups(local)
            self.first = next(self.blocks)                 # every transfer is on the wire once the generator yields
        else:
            self.slices = pg.feature_slices(local.size(1))
            if pg.world == 1 and pg._gather is None:
                self.pending = [lambda: local]
                self.slices = [(0, local.size(1))]
            else:
                self.pending = [pg.gather_cols_async(local, lo, hi) for lo, hi in self.slices]

    def aggregate(self, out, row_cnt, row_x, act, row_add=None):
        pg, backward = self.pg, self.backward
        if pg.halo_mode in ("blocks", "chunks"):
            # overlapped exchange: the own block first, every other owner block (or: row chunk after row chunk) as
            # it lands, accumulated into `out`; the count division / backward relu mask / row_add belong to the
            # finished sum: the last call
            csrs = pg.group_csrs(backward)
            last = pg.n_groups - 1
            items = [self.first]
            i = 0
            while True:
                b, blk = items.pop() if items else next(self.blocks)
                fin = i == last
                _aggregate_call(csrs[b], blk.size(0), blk, out, row_cnt=row_cnt if fin else None,
                                row_x=row_x if (fin and backward) else None,
                                act=act if (fin or not backward) else ACT_NONE, backward=backward, accumulate=i > 0,
                                row_add=row_add if fin else None)
                if fin:
                    return
                i += 1
        csr = pg.bwd_csr() if backward else pg.fwd_csr()
        whole = len(self.slices) == 1
        for (lo, hi), fin in zip(self.slices, self.pending):
            full = fin()                                                           # wait for this slice only
            _aggregate_call(csr, full.size(0), full, out, row_cnt=row_cnt, row_x=row_x, act=act, backward=backward,
                            out_cols=None if whole else (lo, hi), row_add=row_add)


class _PartAggregate(torch.autograd.Function):
    """The same aggregation on a 1-D destination-partitioned graph: x holds this rank's rows;
    the forward all-gathers the row blocks (halo exchange) and runs K1 on the owned destination
    rows, the backward all-gathers the (pre-scaled) g_out blocks and runs K2 on the owned source
    rows.  The exchange is pipelined by FEATURE slices: all slices are issued asynchronously, the
    aggregation of slice k runs while slices k+1.. are still on the wire.  A row's sum still walks
    the same edges in the same order, so the result stays bit-identical to one GPU."""

    @staticmethod
    def _run(pg, csr, local, out, row_cnt, row_x, act, backward):
        HaloExchange(pg, local, backward).aggregate(out, row_cnt, row_x, act)

    @staticmethod
    def forward(ctx, x, pg, reduce, act):
        x = _f32c(x, "x")
        if x.dim() != 2 or x.size(0) != pg.num_nodes:
            raise ValueError(f"x must hold this rank's {pg.num_nodes} rows")
        fwd = pg.fwd_csr()
        out = torch.empty_like(x)
        _PartAggregate._run(pg, fwd, x, out, fwd.cnt if reduce == REDUCE_MEAN else None, None, act, False)
        ctx.pg, ctx.reduce, ctx.act = pg, reduce, act
        ctx.save_for_backward(x)
        return out

    @staticmethod
    def backward(ctx, g_out):
        (x,) = ctx.saved_tensors
        pg = ctx.pg
        g_out = _f32c(g_out, "grad")
        if ctx.reduce == REDUCE_MEAN:       # own rows, own counts: no count exchange
            g_out = _row_scale(g_out, pg.fwd_csr().cnt)
        g_x = torch.empty_like(x)
        _PartAggregate._run(pg, pg.bwd_csr(), g_out, g_x, None, x if ctx.act == ACT_RELU else None, ctx.act, True)
        return g_x, None, None, None


def aggregate(x, graph, *, to_row=1, reduce="mean", act="none", edge_attr=None, edge_type=None,
              type_table=None, edge_weight=None):
    """Fused gather -> message -> reduce over a GraphStructure.

    edge_attr: dense [E, F] per-edge features in original edge order (what the reference passes,
    model/encoder.py:94-97), or (type_table [T,F], edge_type [E] int32/int64) for the table form.
    Gradients flow to `x` only: edge features / weights are data on this path."""
    if isinstance(graph, PartitionedGraph):
        if to_row != 1 or any(t is not None for t in (edge_attr, edge_type, type_table, edge_weight)):
            raise NotImplementedError("partitioned graphs: destination aggregation without per-edge "
                                      "data (the BASELINE products-shaped configuration)")
        r = {"mean": REDUCE_MEAN, "sum": REDUCE_SUM, "add": REDUCE_SUM}[reduce]
        a = {"none": ACT_NONE, "relu": ACT_RELU}[act]
        return _PartAggregate.apply(x, graph, r, a)
    if not isinstance(graph, GraphStructure):
        raise TypeError("graph must be a GraphStructure or a PartitionedGraph")
    for t, n in ((edge_attr, "edge_attr"), (type_table, "type_table"), (edge_weight, "edge_weight")):
        if t is not None and t.requires_grad:
            raise NotImplementedError(f"gradient w.r.t. {n} is not part of the hot path")
    E = graph.edge_index.size(1)
    edge_attr = _f32c(edge_attr, "edge_attr")
    type_table = _f32c(type_table, "type_table")
    edge_weight = _f32c(edge_weight, "edge_weight")
    if edge_attr is not None and tuple(edge_attr.shape) != (E, x.size(1)):
        raise ValueError(f"edge_attr must be [E, F] = [{E}, {x.size(1)}], got {tuple(edge_attr.shape)}")
    if edge_type is not None:
        if type_table is None or type_table.size(1) != x.size(1) or edge_type.numel() != E:
            raise ValueError("edge_type needs a [T, F] type_table and one id per edge")
        edge_type = edge_type.to(torch.int32).contiguous()
    if edge_weight is not None and edge_weight.numel() != E:
        raise ValueError("edge_weight must have one entry per edge")
    r = {"mean": REDUCE_MEAN, "sum": REDUCE_SUM, "add": REDUCE_SUM}[reduce]
    a = {"none": ACT_NONE, "relu": ACT_RELU}[act]
    if torch.is_grad_enabled():
        graph.prepare_training()            # a training step walks both groupings: one build for the pair
    return _Aggregate.apply(x, graph, to_row, r, a, edge_attr, type_table, edge_type, edge_weight)


# ------------------------------------------------------------------------------- l2norm (K11)
class _L2Norm(torch.autograd.Function):
    """F.normalize(x, p=2, dim=-1) over the rows of a contiguous [R, D] matrix, one kernel each way."""

    @staticmethod
    def forward(ctx, x, eps):
        x = _f32c(x, "x")
        r, d = x.shape
        y = torch.empty_like(x)
        norm = torch.empty(r, dtype=torch.float32, device=x.device)
        with _lib.traced("l2norm_forward", 2 * r * d * 4 + r * 4, 0, 1):
            _lib.check(_lib.lib().stemgnn_l2norm_forward(r, d, _lib.ptr(x), eps, _lib.ptr(y), _lib.ptr(norm),
                                                         _lib.stream()))
        ctx.save_for_backward(y, norm)
        ctx.eps = eps
        return y

    @staticmethod
    def backward(ctx, g):
        y, norm = ctx.saved_tensors
        g = _f32c(g, "grad")
        r, d = y.shape
        gx = torch.empty_like(y)
        with _lib.traced("l2norm_backward", 3 * r * d * 4 + r * 4, 0, 1):
            _lib.check(_lib.lib().stemgnn_l2norm_backward(r, d, _lib.ptr(g), _lib.ptr(y), _lib.ptr(norm), ctx.eps,
                                                          _lib.ptr(gx), _lib.stream()))
        return gx, None


def l2norm_rows(x, eps=1e-12):
    """x / max(||x||_2, eps) per row of a 2-D fp32 CUDA matrix (model/vq.py:48-49), differentiable."""
    if x.dim() != 2 or x.size(1) % 4:
        raise ValueError("l2norm_rows: a 2-D matrix whose width is a multiple of 4")
    return _L2Norm.apply(x, float(eps))


# ------------------------------------------------------------------------------- VQ (K3 / K4)
METRIC_EUCLID, METRIC_DOT = 0, 1
PATH_AUTO, PATH_EXACT, PATH_TENSOR = 0, 1, 2


def _head_layout(x, heads, embed):
    """x [M, heads*D] row-major; embed [Hc, K, D].  A codebook shared by several heads sees the
    heads as extra rows ('b n (h d) -> 1 (b h) n d', model/vq.py:885-887), i.e. x viewed as
    [M*heads, D] with one head."""
    Hc, K, D = embed.shape
    M = x.size(0)
    if x.size(1) != heads * D:
        raise ValueError(f"x has {x.size(1)} columns, expected heads*codebook_dim = {heads * D}")
    if Hc == heads:
        return dict(heads=heads, embed_heads=Hc, rows=M, codes=K, dim=D, row_stride=heads * D,
                    head_stride=D, ind_row_stride=heads, ind_head_stride=1)
    if Hc != 1:
        raise ValueError("embed must have 1 or `heads` codebooks")
    return dict(heads=1, embed_heads=1, rows=M * heads, codes=K, dim=D, row_stride=D,
                head_stride=0, ind_row_stride=1, ind_head_stride=0)


def vq_assign_raw(x, embed, heads, metric, path=PATH_AUTO, want_sq_err=True, want_stats=False):
    """One stemgnn_vq_assign call.  Returns (ind [M, heads] int64, quantize [M, heads*D] fp32,
    sq_err [heads'] fp64 or None, stats int64[4] or None)."""
    x, embed = _f32c(x, "x"), _f32c(embed, "embed")
    lay = _head_layout(x, heads, embed)
    dev = x.device
    ind = torch.empty((x.size(0), heads), dtype=torch.int64, device=dev)
    q = torch.empty_like(x)
    sq = torch.empty(lay["heads"], dtype=torch.float64, device=dev) if want_sq_err else None
    stats = torch.zeros(4, dtype=torch.int64, device=dev) if want_stats else None
    a = _lib.VqAssignArgs(
        x=_lib.ptr(x), x_row_stride=lay["row_stride"], x_head_stride=lay["head_stride"],
        embed=_lib.ptr(embed), heads=lay["heads"], embed_heads=lay["embed_heads"],
        rows=lay["rows"], codes=lay["codes"], dim=lay["dim"], metric=metric, path=path,
        ind=_lib.ptr(ind), ind_row_stride=lay["ind_row_stride"],
        ind_head_stride=lay["ind_head_stride"], quantize=_lib.ptr(q),
        q_row_stride=lay["row_stride"], q_head_stride=lay["head_stride"],
        sq_err=_lib.ptr(sq), stats=_lib.ptr(stats))
    L = _lib.lib()
    need = C.c_size_t(0)
    _lib.check(L.stemgnn_vq_assign_workspace(C.byref(a), C.byref(need)))
    ws = _lib.workspace(need.value, dev)
    M, K, D, H = lay["rows"], lay["codes"], lay["dim"], lay["heads"]
    nbytes = H * M * D * 4 * 2 + H * M * 8 + embed.numel() * 4
    with _lib.traced("vq_assign", nbytes, 2 * H * M * K * D):
        _lib.check(L.stemgnn_vq_assign(C.byref(a), _lib.ptr(ws), need.value, _lib.stream()))
    return ind, q, sq, stats


def vq_code_stats(x, ind, embed_shape, heads):
    """counts [Hc, K], sums [Hc, K, D] of the rows assigned to every code (model/vq.py:492,497)."""
    x = _f32c(x, "x")
    Hc, K, D = embed_shape
    lay = _head_layout(x, heads, torch.empty(embed_shape, device="meta"))
    dev = x.device
    ind = ind.contiguous()
    # one flat buffer [counts | sums] so a multi-rank caller all-reduces it in a single call
    flat = torch.empty(Hc * K * (1 + D), dtype=torch.float32, device=dev)
    counts, sums = flat[:Hc * K].view(Hc, K), flat[Hc * K:].view(Hc, K, D)
    a = _lib.VqCodeStatsArgs(
        x=_lib.ptr(x), x_row_stride=lay["row_stride"], x_head_stride=lay["head_stride"],
        ind=_lib.ptr(ind), ind_row_stride=lay["ind_row_stride"],
        ind_head_stride=lay["ind_head_stride"], heads=lay["heads"],
        embed_heads=lay["embed_heads"], rows=lay["rows"], codes=K, dim=D,
        counts=_lib.ptr(counts), sums=_lib.ptr(sums))
    L = _lib.lib()
    need = C.c_size_t(0)
    _lib.check(L.stemgnn_vq_code_stats_workspace(C.byref(a), C.byref(need)))
    ws = _lib.workspace(need.value, dev)
    T = lay["heads"] * lay["rows"]
    with _lib.traced("vq_code_stats", T * D * 4 + T * 8 + 2 * Hc * K * (D + 1) * 4, 0):
        _lib.check(L.stemgnn_vq_code_stats(C.byref(a), _lib.ptr(ws), need.value, _lib.stream()))
    return counts, sums, flat


def vq_ema_apply(counts, sums, cluster_size, embed_avg, embed, decay, eps, cosine):
    """In-place lerp_ + laplace smoothing + divide (+ double l2norm), model/vq.py:495-506."""
    Hc, K, D = embed.shape
    for t in (cluster_size, embed_avg, embed):
        if not t.is_contiguous() or t.dtype != torch.float32:
            raise ValueError("codebook buffers must be contiguous float32")
    scratch = torch.empty(Hc, dtype=torch.float32, device=embed.device)
    with _lib.traced("vq_ema_apply", 4 * Hc * K * (D + 1) * 4, 0, 2):
        _lib.check(_lib.lib().stemgnn_vq_ema_apply(
            _lib.ptr(counts), _lib.ptr(sums), Hc, K, D, float(decay), float(eps), int(cosine),
            _lib.ptr(cluster_size), _lib.ptr(embed_avg), _lib.ptr(embed), _lib.ptr(scratch),
            _lib.stream()))


class _VqQuantize(torch.autograd.Function):
    """(quantize, ind, mse) = nearest code + straight-through + mean((quantize - x)^2).

    backward (model/vq.py:937,1007-1009): g_x = g_quantize + g_mse * 2 (x - q) / numel.  When the
    codebook is learnable the commitment term also reaches embed (commit_quantize is not
    detached, vq.py:931-933): g_embed[k] = g_mse * 2/numel * (count_k * embed_k - sum_{n->k} x_n),
    obtained from the K4 statistics instead of an N x K one-hot GEMM."""

    @staticmethod
    def forward(ctx, x, embed, heads, metric, path, straight_through, embed_grad):
        ind, q, sq, _ = vq_assign_raw(x, embed, heads, metric, path)
        numel = x.numel()
        mse = (sq.sum() / max(numel, 1)).to(torch.float32)
        ctx.heads, ctx.st, ctx.embed_grad, ctx.numel = heads, straight_through, embed_grad, numel
        ctx.save_for_backward(x, q, ind, embed)
        ctx.mark_non_differentiable(ind)
        return q, ind, mse

    @staticmethod
    def backward(ctx, g_q, _g_ind, g_mse):
        x, q, ind, embed = ctx.saved_tensors
        g_x = g_embed = None
        if ctx.needs_input_grad[0]:
            gq = _f32c(g_q, "grad") if (g_q is not None and ctx.st) else None
            gl = None
            if g_mse is not None:
                gl = g_mse.reshape(1).to(torch.float32).contiguous()
            g_x = torch.empty_like(x)
            with _lib.traced("vq_backward", x.numel() * 4 * (4 if gq is not None else 3), 0, 1):
                _lib.check(_lib.lib().stemgnn_vq_backward(
                    _lib.ptr(gq), _lib.ptr(x), _lib.ptr(q), _lib.ptr(gl), 2.0 / max(ctx.numel, 1),
                    x.numel(), _lib.ptr(g_x), _lib.stream()))
        if ctx.embed_grad and ctx.needs_input_grad[1] and g_mse is not None:
            counts, sums, _ = vq_code_stats(x, ind, tuple(embed.shape), ctx.heads)
            g_embed = (counts.unsqueeze(-1) * embed - sums) * (g_mse * (2.0 / max(ctx.numel, 1)))
        return g_x, g_embed, None, None, None, None, None


def vq_quantize(x, embed, heads=1, metric=METRIC_EUCLID, path=PATH_AUTO, straight_through=True,
                embed_grad=False):
    return _VqQuantize.apply(x, embed, heads, metric, path, straight_through, embed_grad)
