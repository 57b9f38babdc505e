"""GPU: K9 (dropout_adj, mask_feature, gather_rows) against the oracle restatement of the
torch_geometric.utils 2.3.0 functions the reference calls at pretrain.py:33-44.  Integer / copy
work: bit-exact."""
import numpy as np
import pytest
import torch

import stemgnn_b200 as S
from oracle.augment_ref import dropout_adj_ref, mask_feature_ref

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _graph(n, e, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, n, (2, e), generator=g)


@pytest.mark.parametrize("e", [0, 1, 255, 2047, 2048, 2049, 100_003, 1_166_243])
@pytest.mark.parametrize("undirected", [False, True])
def test_dropout_adj_bit_exact_with_torch_rand(e, undirected):
    n = max(2, e // 7)
    ei = _graph(n, e, e + 1)
    attr = torch.randn(e, 8, generator=torch.Generator().manual_seed(2))
    etype = torch.randint(0, 5, (e,), generator=torch.Generator().manual_seed(3))
    for p in (0.3, 0.85):
        torch.manual_seed(11)
        out, a, eid = S.dropout_adj(ei.to(DEV), attr.to(DEV), p=p, force_undirected=undirected,
                                    num_nodes=n, return_edge_id=True)
        torch.manual_seed(11)
        rand = torch.rand(e, device=DEV).cpu()            # the draw dropout_adj made
        ro, ra, rid = dropout_adj_ref(ei, attr, rand, p, undirected)
        assert out.dtype == torch.int64 and tuple(out.shape) == tuple(ro.shape)
        assert torch.equal(out.cpu(), ro) and torch.equal(eid.cpu(), rid)
        assert torch.equal(a.cpu(), ra)
        torch.manual_seed(11)
        _, t = S.dropout_adj(ei.to(DEV), etype.to(DEV), p=p, force_undirected=undirected)
        assert torch.equal(t.cpu(), etype[rid])


def test_dropout_adj_extremes_and_sorted_edges():
    torch.manual_seed(0)
    e = 5000
    ei = _graph(50, e, 0).to(DEV)
    out, _ = S.dropout_adj(ei, p=1.0)
    assert tuple(out.shape) == (2, 0)
    out, _ = S.dropout_adj(ei, p=1e-12)                   # keeps everything, in order
    assert torch.equal(out, ei)
    up = torch.stack([torch.zeros(e, dtype=torch.int64), torch.ones(e, dtype=torch.int64)]).to(DEV)
    out, _ = S.dropout_adj(up, p=1e-12, force_undirected=True)
    assert tuple(out.shape) == (2, 2 * e) and torch.equal(out[0, :e], up[0]) and torch.equal(out[0, e:], up[1])
    down = up.flip(0)                                      # row > col everywhere: nothing survives
    out, _ = S.dropout_adj(down.contiguous(), p=1e-12, force_undirected=True)
    assert out.size(1) == 0


def test_dropout_adj_philox_stream():
    n, e, p = 1000, 400_000, 0.25
    ei = _graph(n, e, 5).to(DEV)
    out, _, eid = S.dropout_adj(ei, p=p, force_undirected=True, generator="philox", return_edge_id=True)
    m = out.size(1) // 2
    assert torch.equal(out[0, :m], out[1, m:]) and torch.equal(out[1, :m], out[0, m:])
    assert bool((out[0, :m] <= out[1, :m]).all())
    assert bool((eid[1:m] > eid[:m - 1]).all()) and torch.equal(ei[:, eid[:m]], out[:, :m])
    eligible = int((ei[0] <= ei[1]).sum())
    assert abs(m / eligible - (1 - p)) < 0.01             # 4 sigma ~ 0.004
    out2, _ = S.dropout_adj(ei, p=p, generator="philox")
    assert abs(out2.size(1) / e - (1 - p)) < 0.01


@pytest.mark.parametrize("mode", ["col", "row", "all"])
def test_mask_feature_matches_masked_fill(mode):
    n, f = 3001, 768
    x = torch.randn(n, f, generator=torch.Generator().manual_seed(1))
    x[5, 7] = float("nan")
    xg = x.to(DEV).requires_grad_(True)
    torch.manual_seed(3)
    out, mask = S.mask_feature(xg, p=0.4, mode=mode, fill_value=-2.5)
    assert mask.dtype == torch.bool
    assert tuple(mask.shape) == {"col": (1, f), "row": (n, 1), "all": (n, f)}[mode]
    ref = mask_feature_ref(x, mask.cpu(), -2.5)
    assert np.array_equal(out.detach().cpu().numpy(), ref.numpy(), equal_nan=True)
    rate = float(mask.float().mean())
    tol = {"col": 0.08, "row": 0.04, "all": 0.003}[mode]
    assert abs(rate - 0.6) < tol
    if mode != "all":                                      # same draw as PyG on this device and seed
        torch.manual_seed(3)
        want = torch.rand(f if mode == "col" else n, device=DEV) >= 0.4
        assert torch.equal(mask.view(-1), want)
    g = torch.randn(n, f, generator=torch.Generator().manual_seed(2)).to(DEV)
    out.backward(g)
    assert torch.equal(xg.grad, g * mask)                  # masked_fill's backward


@pytest.mark.parametrize("f", [4, 32, 100, 768])
def test_gather_rows_bit_exact(f):
    src = torch.randn(5000, f, generator=torch.Generator().manual_seed(f)).to(DEV)
    ids = torch.randint(0, 5000, (20011,), generator=torch.Generator().manual_seed(1)).to(DEV)
    assert torch.equal(S.gather_rows(src, ids), src[ids])
    assert S.gather_rows(src, ids[:0]).shape == (0, f)
    bad = torch.tensor([1, 5000, -1], device=DEV)
    out = S.gather_rows(src, bad)
    assert torch.equal(out[0], src[1]) and bool(out[1:].isnan().all())


def test_negative_sampling_properties_and_determinism():
    from oracle.augment_ref import check_negative_edges
    n, e = 20000, 150000
    ei = _graph(n, e, 8)
    ei[1, :50] = ei[0, :50]                                 # self loops among the positives
    ei[:, 50:100] = ei[:, 100:150]                          # duplicated positives
    torch.manual_seed(5)
    neg = S.negative_sampling(ei.to(DEV), n)
    assert neg.dtype == torch.int64 and tuple(neg.shape) == (2, e)
    check_negative_edges(ei.numpy(), n, neg.cpu().numpy())
    torch.manual_seed(5)
    assert torch.equal(S.negative_sampling(ei.to(DEV), n), neg)      # same torch seed, same sample
    torch.manual_seed(6)
    assert not torch.equal(S.negative_sampling(ei.to(DEV), n), neg)
    few = S.negative_sampling(ei.to(DEV), n, num_neg_samples=1000)
    assert tuple(few.shape) == (2, 1000)
    assert S.negative_sampling(ei[:, :0].to(DEV), n).shape == (2, 0)


def test_negative_sampling_dense_graph_and_uniformity():
    from oracle.augment_ref import check_negative_edges
    n = 12
    full = torch.tensor([[i, j] for i in range(n) for j in range(n) if i != j]).t().contiguous()
    keep = torch.randperm(full.size(1), generator=torch.Generator().manual_seed(1))[:80]
    ei = full[:, keep]                                       # 80 of the 132 pairs are edges
    non_edges = 132 - 80
    hits = torch.zeros(n * n)
    torch.manual_seed(0)
    trials = 600
    for _ in range(trials):
        neg = S.negative_sampling(ei.to(DEV), n, num_neg_samples=20).cpu()
        check_negative_edges(ei.numpy(), n, neg.numpy())
        assert neg.size(1) == 20
        hits[neg[0] * n + neg[1]] += 1
    assert int((hits > 0).sum()) == non_edges               # every non-edge is reachable
    expect = trials * 20 / non_edges
    chi2 = float(((hits[hits > 0] - expect) ** 2 / expect).sum())
    assert chi2 < 110, chi2                                  # 51 dof: P(chi2 > 110) ~ 3e-6
    # more requested than exist: everything that is left comes back
    torch.manual_seed(1)
    allneg = S.negative_sampling(ei.to(DEV), n, num_neg_samples=1000)
    check_negative_edges(ei.numpy(), n, allneg.cpu().numpy())
    assert 40 <= allneg.size(1) <= non_edges            # like PyG, three rounds may leave a few undrawn


def test_edge_concat_linear_matches_cat_linear():
    n, e, d, o = 3000, 20011, 64, 32
    g = torch.Generator().manual_seed(2)
    z = torch.randn(n, d, generator=g).to(DEV).requires_grad_(True)
    ei = torch.randint(0, n, (2, e), generator=g).to(DEV)
    lin = torch.nn.Linear(2 * d, o).to(DEV)
    up = torch.randn(e, o, generator=g).to(DEV)
    out = S.edge_concat_linear(z, ei, lin)
    out.backward(up)
    gz, gw, gb = z.grad.clone(), lin.weight.grad.clone(), lin.bias.grad.clone()
    z.grad = None; lin.zero_grad()
    ref = lin(torch.cat([z[ei[0]], z[ei[1]]], dim=-1))      # pt_model.py:80-81
    ref.backward(up)
    scale = lambda t: t.abs().max().item()
    assert (out - ref).abs().max().item() < 1e-5 * scale(ref)
    assert (gz - z.grad).abs().max().item() < 2e-5 * scale(z.grad)
    assert (gw - lin.weight.grad).abs().max().item() < 2e-5 * scale(lin.weight.grad)
    assert (gb - lin.bias.grad).abs().max().item() < 2e-5 * scale(lin.bias.grad)
    z.grad = None
    S.edge_concat_linear(z, ei, lin).backward(up)            # the gather's backward is a pull in position order:
    assert torch.equal(z.grad, gz)                           # bit-identical from run to run (index_add_ is not)
