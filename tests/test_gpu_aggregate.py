"""GPU: K0 (CSR build) and K1/K2 (aggregation fwd/bwd) through the C-ABI against the oracle.
Integer outputs bit-exact; fp32 aggregation bit-exact too, because the kernels add a row's
messages one by one in stable edge order like scatter_add_ (north_star bound: 1e-5 relative)."""
import numpy as np
import pytest
import torch

import stemgnn_b200 as S
from oracle import graph_ref
from oracle import propagate_ref as P

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _graph(n, e, seed, hub=0):
    g = torch.Generator().manual_seed(seed)
    ei = torch.randint(0, n, (2, e), generator=g)
    if hub:
        ei[1, :hub] = n // 3
        ei[0, hub:2 * hub] = n // 2
    return ei


@pytest.mark.parametrize("n,e,hub", [(1, 0, 0), (7, 0, 0), (50, 400, 0), (2708, 10556, 0),
                                     (5000, 70000, 9000), (169343, 1166243, 0)])
def test_csr_build_bit_exact(n, e, hub):
    ei = _graph(n, e, 1, hub)
    g = S.GraphStructure(ei.to(DEV), n)
    for by in (0, 1):
        rowptr, col, perm, deg = graph_ref.csr_build_ref(ei.numpy(), n, by)
        c = g.grouped_by(by)
        assert np.array_equal(c.rowptr.cpu().numpy(), rowptr)
        assert np.array_equal(c.col.cpu().numpy(), col)
        assert np.array_equal(c.perm.cpu().numpy(), perm)
        assert np.array_equal(c.cnt.cpu().numpy(), np.maximum(deg, 1).astype(np.float32))


@pytest.mark.parametrize("n,e,hub", [(1, 0, 0), (3, 1, 0), (300, 5000, 0), (512, 40000, 9000), (513, 3000, 0),
                                     (5000, 70000, 9000), (40000, 300000, 21000), (169343, 1166243, 0),
                                     (600000, 2000000, 30000), (2449029 // 4, 61859140 // 16, 0)])
def test_csr_single_and_pair_builds_bit_exact(n, e, hub):
    """Every plan of the round-2 build: no upper pass (N <= 512), one upper pass of 256 / 512 bins, two upper
    passes (N > 2^18), segments of several chunks (hubs), both through stemgnn_csr_build (no_grad: one grouping per
    call) and stemgnn_csr_build_pair (grad enabled: both groupings in one set of launches)."""
    ei = _graph(n, e, 3, hub)
    ei_d = ei.to(DEV)
    refs = [graph_ref.csr_build_ref(ei.numpy(), n, by) for by in (0, 1)]
    with torch.no_grad():
        g1 = S.GraphStructure(ei_d, n)
        single = [g1.grouped_by(0), g1.grouped_by(1)]
    g2 = S.GraphStructure(ei_d, n)
    g2.prepare_training()
    assert set(g2._csr) == {0, 1}                       # one call built both
    pair = [g2.grouped_by(0), g2.grouped_by(1)]
    if e:
        x = torch.randn(n, 4, device=DEV, requires_grad=True)
        g3 = S.GraphStructure(ei_d, n)
        c0 = S._lib.launch_count()
        S.aggregate(x, g3, reduce="sum")                # grad mode on: the op asks for the pair up front
        assert set(g3._csr) == {0, 1}
        if n == 169343:                                 # 18-bit ids: hist, scan, scatter, bucket + the aggregation's two
            assert S._lib.launch_count() - c0 == 6
    for by in (0, 1):
        rowptr, col, perm, deg = refs[by]
        for c in (single[by], pair[by]):
            assert np.array_equal(c.rowptr.cpu().numpy(), rowptr)
            assert np.array_equal(c.perm.cpu().numpy(), perm)
            assert np.array_equal(c.col.cpu().numpy(), col)
            assert np.array_equal(c.cnt.cpu().numpy(), np.maximum(deg, 1).astype(np.float32))


def _check(x, ei, n, *, reduce, act, xe=None, table=None, etype=None, exact=True):
    xc = x.clone().requires_grad_(True)
    xe_full = xe if xe is not None else (table[etype] if table is not None else None)
    if act == "relu":
        ref = P.sage_propagate_ref(xc, ei, xe_full, n, reduce)
    else:
        assert xe_full is None
        xj = xc.index_select(0, ei[0])
        ref = P.scatter_mean_ref(xj, ei[1], n) if reduce == "mean" else P._scatter_sum(xj, ei[1], n)
    go = torch.randn(ref.shape, generator=torch.Generator().manual_seed(5))
    ref.backward(go)
    xg = x.to(DEV).requires_grad_(True)
    g = S.GraphStructure(ei.to(DEV), n)
    kw = {}
    if xe is not None:
        kw["edge_attr"] = xe.to(DEV)
    if table is not None:
        kw.update(type_table=table.to(DEV), edge_type=etype.to(DEV))
    out = S.aggregate(xg, g, reduce=reduce, act=act, **kw)
    out.backward(go.to(DEV))
    cmp = np.testing.assert_array_equal if exact else \
        (lambda a, b: np.testing.assert_allclose(a, b, rtol=1e-5, atol=1e-6))
    cmp(out.detach().cpu().numpy(), ref.detach().numpy())
    cmp(xg.grad.cpu().numpy(), xc.grad.numpy())


@pytest.mark.parametrize("f", [1, 3, 20, 32, 37, 64, 100, 128, 256, 260, 768, 1433])
@pytest.mark.parametrize("reduce,act", [("mean", "relu"), ("sum", "none"), ("mean", "none")])
def test_aggregate_shapes(f, reduce, act):
    n, e = 300, 3000
    ei = _graph(n, e, 2, hub=400)
    ei[1][ei[1] == 5] = 6       # isolated destination -> 0 / clamp(count, 1)
    x = torch.randn(n, f, generator=torch.Generator().manual_seed(3))
    _check(x, ei, n, reduce=reduce, act=act)


@pytest.mark.parametrize("f", [24, 100, 768])
def test_aggregate_edge_features_dense_and_table(f):
    n, e = 200, 2500
    g = torch.Generator().manual_seed(4)
    ei = _graph(n, e, 4, hub=300)
    x = torch.randn(n, f, generator=g)
    table = torch.randn(5, f, generator=g)
    etype = torch.randint(0, 5, (e,), generator=g)
    _check(x, ei, n, reduce="mean", act="relu", xe=table[etype])
    _check(x, ei, n, reduce="mean", act="relu", table=table, etype=etype)


def _hub_graph(n, e, seed, hubs):
    """Uniform graph in which node h of `hubs` = {node: degree} receives (and node h + 1 sends) that many
    edges, interleaved with the others so that a hub's edges are spread over the whole edge list."""
    g = torch.Generator().manual_seed(seed)
    ei = torch.randint(0, n, (2, e), generator=g)
    pos = torch.randperm(e, generator=g)
    at = 0
    for node, deg in hubs.items():
        ei[1, pos[at:at + deg]] = node
        ei[0, pos[at + deg:at + 2 * deg]] = node + 1
        at += 2 * deg
    return ei


@pytest.mark.parametrize("f", [8, 36, 100, 128, 256, 768])
def test_hub_rows_every_mode_bit_exact(f):
    """Rows of 255 / 256 / 257 edges sit on both sides of the hub threshold, 4,095 / 4,096 on both sides of
    the big-hub one (8-column slices), 700 and 5,003 span several staged tiles with a ragged tail: same
    bits as the sequential scatter_add_ in every mode."""
    n, e = 600, 40000
    hubs = {7: 255, 100: 256, 200: 257, 300: 700, 400: 5003, 500: 4096, 550: 4095}
    g = torch.Generator().manual_seed(11)
    ei = _hub_graph(n, e, 11, hubs)
    deg = torch.bincount(ei[1], minlength=n)
    assert deg[400] >= 5003 and deg[100] >= 256
    x = torch.randn(n, f, generator=g)
    _check(x, ei, n, reduce="mean", act="relu")
    _check(x, ei, n, reduce="sum", act="none")
    table = torch.randn(7, f, generator=g)
    etype = torch.randint(0, 7, (e,), generator=g)
    _check(x, ei, n, reduce="mean", act="relu", xe=table[etype])
    _check(x, ei, n, reduce="mean", act="relu", table=table, etype=etype)


def test_hub_rows_weighted_gcn_sum():
    n, e, f = 500, 12000, 64
    g = torch.Generator().manual_seed(12)
    ei = _hub_graph(n, e, 12, {3: 2000, 250: 300})
    xw = torch.randn(n, f, generator=g)
    xc = xw.clone().requires_grad_(True)
    ref = P.gcn_propagate_ref(xc, ei, n)
    go = torch.randn(n, f, generator=g)
    ref.backward(go)
    g_sl, w = S.GraphStructure(ei.to(DEV), n).gcn()
    xg = xw.to(DEV).requires_grad_(True)
    out = S.aggregate(xg, g_sl, reduce="sum", edge_weight=w)
    out.backward(go.to(DEV))
    np.testing.assert_array_equal(out.detach().cpu().numpy(), ref.detach().numpy())
    np.testing.assert_array_equal(xg.grad.cpu().numpy(), xc.grad.numpy())


def test_hub_path_equals_row_per_group_path(monkeypatch):
    """A/B inside the library: the same skewed graph with the hub list switched off (every row walked by
    one group of lanes) gives identical bits, forward and backward."""
    n, e, f = 20000, 400000, 256
    g = torch.Generator().manual_seed(13)
    ei = _hub_graph(n, e, 13, {11: 17000, 5000: 13155, 9000: 1000, 15000: 256}).to(DEV)
    x = torch.randn(n, f, generator=g).to(DEV)
    go = torch.randn(n, f, generator=g).to(DEV)
    res = []
    for off in ("0", "1"):
        monkeypatch.setenv("STEMGNN_AGG_NO_HUB", off)
        xg = x.clone().requires_grad_(True)
        out = S.aggregate(xg, S.GraphStructure(ei, n), reduce="mean", act="relu")
        out.backward(go)
        res.append((out.detach(), xg.grad))
    assert torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][1], res[1][1])


def test_aggregate_empty_graph_and_empty_features():
    n = 10
    ei = torch.zeros(2, 0, dtype=torch.int64)
    x = torch.randn(n, 8)
    _check(x, ei, n, reduce="mean", act="relu")


@pytest.mark.parametrize("f", [19, 64])
def test_gcn_weighted_sum(f):
    n, e = 400, 3000
    g = torch.Generator().manual_seed(6)
    ei = _graph(n, e, 6, hub=200)
    ei[1, 500:540] = ei[0, 500:540]   # explicit self loops
    xw = torch.randn(n, f, generator=g)
    xc = xw.clone().requires_grad_(True)
    ref = P.gcn_propagate_ref(xc, ei, n)
    go = torch.randn(n, f, generator=g)
    ref.backward(go)
    gs = S.GraphStructure(ei.to(DEV), n)
    g_sl, w = gs.gcn()
    ei_ref, w_ref = P.gcn_norm_ref(ei, n)
    assert torch.equal(g_sl.edge_index.cpu(), ei_ref)
    np.testing.assert_array_equal(w.cpu().numpy(), w_ref.numpy())
    xg = xw.to(DEV).requires_grad_(True)
    out = S.aggregate(xg, g_sl, reduce="sum", edge_weight=w)
    out.backward(go.to(DEV))
    np.testing.assert_array_equal(out.detach().cpu().numpy(), ref.detach().numpy())
    np.testing.assert_array_equal(xg.grad.cpu().numpy(), xc.grad.numpy())


def test_reverse_direction_mixture_aggregate():
    n, e, f = 300, 2000, 48
    ei = _graph(n, e, 7, hub=100)
    x = torch.randn(n, f, generator=torch.Generator().manual_seed(7))
    xc = x.clone().requires_grad_(True)
    ref = P.mixture_aggregate_ref(xc, ei)
    go = torch.randn(n, f, generator=torch.Generator().manual_seed(8))
    ref.backward(go)
    xg = x.to(DEV).requires_grad_(True)
    out = S.aggregate(xg, S.GraphStructure(ei.to(DEV), n), to_row=0, reduce="mean")
    out.backward(go.to(DEV))
    np.testing.assert_array_equal(out.detach().cpu().numpy(), ref.detach().numpy())
    # the reference gathers with advanced indexing x[col] (encoder.py:124); its CPU backward is
    # index_put_(accumulate=True), whose summation order is not the edge order -> 1e-5 bound
    np.testing.assert_allclose(xg.grad.cpu().numpy(), xc.grad.numpy(), rtol=1e-5, atol=1e-6)


def test_arxiv_shaped_full_size_bit_exact():
    n, e, f = 169343, 1166243, 128
    ei = _graph(n, e, 9)
    x = torch.randn(n, f, generator=torch.Generator().manual_seed(9))
    _check(x, ei, n, reduce="mean", act="relu")


def test_linearity_property_large():
    """Size-independent property at a products-like degree: sum aggregation without activation
    is linear, agg(a x + y) == a agg(x) + agg(y) up to fp32 rounding."""
    n, e, f = 200000, 5000000, 100
    g = torch.Generator(device=DEV).manual_seed(10)
    ei = torch.randint(0, n, (2, e), generator=g, device=DEV)
    gs = S.GraphStructure(ei, n)
    x = torch.randn(n, f, generator=g, device=DEV)
    y = torch.randn(n, f, generator=g, device=DEV)
    lhs = S.aggregate(2.0 * x + y, gs, reduce="sum")
    rhs = 2.0 * S.aggregate(x, gs, reduce="sum") + S.aggregate(y, gs, reduce="sum")
    torch.testing.assert_close(lhs, rhs, rtol=1e-4, atol=1e-4)
    deg = torch.bincount(ei[1], minlength=n).float()
    ones = S.aggregate(torch.ones(n, 4, device=DEV), gs, reduce="sum")
    assert torch.equal(ones[:, 0], deg)


@pytest.mark.gpu
@pytest.mark.parametrize("n,e,f", [(300, 2000, 256), (1000, 3000, 100), (50, 400, 7)])
def test_row_scale_then_sum_equals_per_edge_division(n, e, f):
    """K2a + K2 without gather_cnt gives the same bits as K2 dividing every gathered row per edge
    (the C-ABI keeps both; the host op uses the first)."""
    from importlib import import_module
    ops = import_module("stem-gnn_b200.ops"); graph = import_module("stem-gnn_b200.graph")
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(n + e)
    ei = torch.randint(0, n, (2, e), generator=g).to(dev)
    gs = graph.GraphStructure(ei, n)
    go = torch.randn(n, f, generator=g).to(dev)
    cnt = gs.grouped_by(1).cnt
    bwd = gs.grouped_by(0)
    a = torch.empty_like(go); b = torch.empty_like(go)
    ops._aggregate_call(bwd, n, go, a, gather_cnt=cnt, backward=True)
    scaled = ops._row_scale(go, cnt)
    assert torch.equal(scaled, go / cnt[:, None])
    ops._aggregate_call(bwd, n, scaled, b, backward=True)
    assert torch.equal(a, b)


@pytest.mark.parametrize("f", [36, 100, 256])
@pytest.mark.parametrize("bwd_mask", [False, True])
def test_accumulate_chain_is_the_block_ordered_sum(f, bwd_mask):
    """ABI 5 `accumulate`: one aggregation cut into calls over disjoint edge subsets (here: by source block, as the
    partitioned path does while halo blocks land), each starting from the sums the previous call left, the mean
    division / backward relu mask only in the last call.  The result is the sequential sum over the edges taken
    block by block -- bit for bit, hub rows (>= 256 edges inside one block) included."""
    from stemgnn_b200 import ops
    from stemgnn_b200.graph import CSR
    n, e, blocks = 900, 30000, 3
    g = torch.Generator().manual_seed(41)
    ei = _hub_graph(n, e, 41, {17: 1500, 300: 290})
    x = torch.randn(n, f, generator=g)
    rowx = torch.randn(n, f, generator=g)
    per = n // blocks
    order = torch.cat([((ei[0] // per) == b).nonzero().view(-1) for b in range(blocks)])
    ei_blockwise = ei[:, order]
    xj = x.index_select(0, ei_blockwise[0])
    msg = xj if bwd_mask else torch.relu(xj)
    ref = P._scatter_sum(msg, ei_blockwise[1], n)
    cnt = torch.bincount(ei[1], minlength=n).clamp(min=1).to(torch.float32)
    ref = torch.where(rowx > 0, ref, torch.zeros(())) if bwd_mask else ref / cnt.view(-1, 1)

    xd, out = x.to(DEV), torch.full((n, f), 7.0, device=DEV)
    for b in range(blocks):
        sub = ei[:, (ei[0] // per) == b].to(DEV)
        csr = CSR(sub, n, 1)
        last = b == blocks - 1
        ops._aggregate_call(csr, n, xd, out, accumulate=b > 0,
                            row_cnt=cnt.to(DEV) if (last and not bwd_mask) else None,
                            row_x=rowx.to(DEV) if (last and bwd_mask) else None,
                            act=(ops.ACT_RELU if (not bwd_mask or last) else ops.ACT_NONE),
                            backward=bool(bwd_mask and last))
    np.testing.assert_array_equal(out.cpu().numpy(), ref.numpy())
