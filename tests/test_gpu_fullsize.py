"""GPU: size-independent properties at BASELINE.json's FULL sizes (where the CPU oracle would take
minutes): the ogbn-products-shaped graph (2,449,029 nodes, 61,859,140 edges, 100 features) and the
10M-row VQ sweep point.  Integer results are checked exactly, fp32 ones through identities."""
import pytest
import torch

import stemgnn_b200 as S

pytestmark = pytest.mark.gpu
DEV = "cuda"


def test_products_shaped_csr_and_aggregation_properties():
    n, e, f = 2449029, 61859140, 100
    g = torch.Generator(device=DEV).manual_seed(4)
    ei = torch.randint(0, n, (2, e), generator=g, device=DEV)
    gs = S.GraphStructure(ei, n)
    for row in (1, 0):
        c = gs.grouped_by(row)
        rowptr, perm, col = c.rowptr.long(), c.perm.long(), c.col.long()
        deg = torch.bincount(ei[row], minlength=n)
        assert int(rowptr[0]) == 0 and int(rowptr[-1]) == e
        assert torch.equal(rowptr[1:] - rowptr[:-1], deg)                       # degrees, exactly
        assert torch.equal(c.cnt, deg.clamp(min=1).float())                     # MeanAggregation's clamped count
        keys = ei[row][perm]
        assert bool((keys[1:] >= keys[:-1]).all())                              # grouped by the key ...
        same = keys[1:] == keys[:-1]
        assert bool((perm[1:][same] > perm[:-1][same]).all())                   # ... stably (edge order kept)
        assert torch.equal(torch.sort(perm).values, torch.arange(e, device=DEV))   # a permutation
        assert torch.equal(col, ei[1 - row][perm])
        del rowptr, perm, col, keys, same
    x = torch.randn(n, f, generator=g, device=DEV)
    ones = S.aggregate(torch.ones(n, 4, device=DEV), gs, reduce="sum")
    assert torch.equal(ones[:, 0], torch.bincount(ei[1], minlength=n).float())  # sum of ones = in-degree
    mean_const = S.aggregate(torch.full((n, 4), 3.0, device=DEV), gs, reduce="mean")
    has = torch.bincount(ei[1], minlength=n) > 0
    assert torch.equal(mean_const[:, 0], torch.where(has, 3.0, 0.0))            # mean of a constant; isolated -> 0
    y = torch.randn(n, f, generator=g, device=DEV)
    lhs = S.aggregate(2.0 * x + y, gs, reduce="sum")
    rhs = 2.0 * S.aggregate(x, gs, reduce="sum") + S.aggregate(y, gs, reduce="sum")
    torch.testing.assert_close(lhs, rhs, rtol=1e-4, atol=2e-4)                  # linearity
    # adjoint identity <A x, y> = <x, A^T y>: the backward kernel against the forward one
    xr = x.clone().requires_grad_(True)
    out = S.aggregate(xr, gs, reduce="mean")
    out.backward(y)
    a = (out.detach().double() * y.double()).sum()
    b = (x.double() * xr.grad.double()).sum()
    assert abs(float(a - b)) < 1e-6 * float(out.detach().double().norm() * y.double().norm())


def test_vq_sweep_point_full_size_properties():
    rows, K, D = 10_000_000, 1024, 128
    g = torch.Generator(device=DEV).manual_seed(5)
    embed = torch.randn(1, K, D, generator=g, device=DEV)
    x = torch.randn(rows, D, generator=g, device=DEV)
    pick = torch.randint(0, K, (rows // 10,), generator=g, device=DEV)
    x[::10] = embed[0][pick]                                # every 10th row IS a code row
    ind, q, sq, st = S.vq_assign_raw(x, embed, 1, S.METRIC_EUCLID, S.PATH_AUTO, want_stats=True)
    ind = ind.view(-1)
    assert int(ind.min()) >= 0 and int(ind.max()) < K
    assert torch.equal(q, embed[0][ind])                    # quantize is the gathered code row, bit for bit
    assert torch.equal(ind[::10], pick)                     # a code row is assigned to itself (distance 0)
    ind2, q2, _, _ = S.vq_assign_raw(q, embed, 1, S.METRIC_EUCLID, S.PATH_AUTO)
    assert torch.equal(ind2.view(-1), ind) and torch.equal(q2, q)               # idempotence
    err = (q.double() - x.double()).square().sum()
    assert abs(float(sq[0] - err)) < 1e-6 * float(err)      # the commitment partial is the squared error
    counts, sums, _ = S.vq_code_stats(x, ind.view(-1, 1), (1, K, D), 1)
    assert float(counts.sum()) == rows and torch.equal(counts.view(-1), torch.bincount(ind, minlength=K).float())
    total = x.double().sum(0)
    assert float((sums.double().sum(1).view(-1) - total).abs().max()) < 1e-5 * float(x.double().abs().sum(0).max())
