"""CPU: the oracle restatements of the augmentations (K9) and the neighbour sampler (K10) on
hand-checked cases, and the host-only entry points of the C-ABI (bounds / workspace / argument
errors -- no compute without a GPU)."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import graph_ref
from oracle.augment_ref import dropout_adj_ref, mask_feature_ref
from oracle.sampler_ref import check_sample, neighbor_sample_ref
from stemgnn_b200 import _lib


def test_dropout_adj_ref_hand_case():
    ei = torch.tensor([[0, 1, 2, 2, 3, 1], [1, 0, 2, 1, 0, 3]])
    rand = torch.tensor([0.9, 0.9, 0.1, 0.9, 0.9, 0.5])
    attr = torch.arange(6.).view(6, 1)
    out, a, eid = dropout_adj_ref(ei, attr, rand, 0.5, False)
    assert out.tolist() == [[0, 1, 2, 3, 1], [1, 0, 1, 0, 3]] and eid.tolist() == [0, 1, 3, 4, 5]
    assert a.view(-1).tolist() == [0., 1., 3., 4., 5.]
    out, a, eid = dropout_adj_ref(ei, attr, rand, 0.5, True)      # keeps row <= col only, then mirrors
    assert out.tolist() == [[0, 1, 1, 3], [1, 3, 0, 1]] and eid.tolist() == [0, 5, 0, 5]
    assert a.view(-1).tolist() == [0., 5., 0., 5.]
    out, _, _ = dropout_adj_ref(ei, None, rand, 1.0, False)
    assert out.shape == (2, 0)


def test_mask_feature_ref_fill_semantics():
    x = torch.tensor([[1., float("nan")], [3., 4.]])
    out = mask_feature_ref(x, torch.tensor([[True, False]]), 0.0)
    assert out.tolist() == [[1., 0.], [3., 0.]]                   # masked NaN becomes the fill value


def _toy():
    # in-edges: 0 <- {1, 2, 3}, 1 <- {2}, 2 <- {}, 3 <- {0, 4}, 4 <- {0}
    ei = np.array([[1, 2, 2, 3, 0, 4, 0], [0, 0, 1, 0, 3, 3, 4]])
    return ei, 5


def test_neighbor_sample_ref_hand_case():
    ei, n = _toy()
    rowptr, col, perm, _ = graph_ref.csr_build_ref(ei, n, 1)
    n_id, src, dst, e_id, hops = neighbor_sample_ref(rowptr, col, perm, [0, 3], [-1, -1], None)
    # hop 0: 0 <- 1, 2, 3(already local 1);  3 <- 0 (local 0), 4 -> new order [0, 3, 1, 2, 4]
    assert n_id.tolist() == [0, 3, 1, 2, 4]
    assert src[:5].tolist() == [2, 3, 1, 0, 4] and dst[:5].tolist() == [0, 0, 0, 1, 1]
    assert e_id[:5].tolist() == [0, 1, 3, 4, 5]
    # hop 1: frontier = locals 2 (node 1 <- 2), 3 (node 2, none), 4 (node 4 <- 0)
    assert src[5:].tolist() == [3, 0] and dst[5:].tolist() == [2, 4] and e_id[5:].tolist() == [2, 6]
    assert hops == [(5, 5), (5, 7)]
    check_sample(ei, n, rowptr, [0, 3], [-1, -1], n_id, src, dst, e_id)


def test_neighbor_sample_ref_random_subsets_are_valid():
    rng = np.random.default_rng(0)
    n, e = 300, 6000
    ei = rng.integers(0, n, (2, e))
    rowptr, col, perm, _ = graph_ref.csr_build_ref(ei, n, 1)
    seeds = rng.permutation(n)[:32]
    choose = lambda h, i, deg, k: rng.permutation(deg)[:k]
    out = neighbor_sample_ref(rowptr, col, perm, seeds, [10, 5], choose)
    check_sample(ei, n, rowptr, seeds, [10, 5], *out[:4])
    with pytest.raises(AssertionError):
        bad = out[3].copy()
        bad[1] = bad[0]
        check_sample(ei, n, rowptr, seeds, [10, 5], out[0], out[1], out[2], bad)


def test_sampler_bounds_and_argument_errors():
    h = _lib.lib()
    nc, ec, ws = C.c_int64(0), C.c_int64(0), C.c_size_t(0)
    fan = (C.c_int32 * 2)(10, 10)
    assert h.stemgnn_neighbor_sample_bounds(1024, 195095, 1274551, 2, fan, C.byref(nc), C.byref(ec), C.byref(ws)) == 0
    assert ec.value == 1024 * 10 + 1024 * 100 and nc.value == 1024 + 10240 + 102400 and ws.value > 0
    fan = (C.c_int32 * 2)(-1, -1)          # all neighbours: bounded by E per hop and by N per frontier
    assert h.stemgnn_neighbor_sample_bounds(512, 1000, 5000, 2, fan, C.byref(nc), C.byref(ec), C.byref(ws)) == 0
    assert ec.value == 10000 and nc.value == 512 + 1000 + 1000
    fan = (C.c_int32 * 1)(65)
    assert h.stemgnn_neighbor_sample_bounds(8, 100, 500, 1, fan, C.byref(nc), C.byref(ec), C.byref(ws)) == -4
    assert b"fan-out" in h.stemgnn_last_error()
    assert h.stemgnn_neighbor_sample_bounds(-1, 100, 500, 1, fan, None, None, None) == -1


def test_dropout_adj_argument_errors():
    import stemgnn_b200 as S
    h = _lib.lib()
    need = C.c_size_t(0)
    assert h.stemgnn_dropout_adj_workspace(10 ** 6, C.byref(need)) == 0 and need.value >= 4 * (10 ** 6 // 2048)
    assert h.stemgnn_dropout_adj_workspace(-3, C.byref(need)) == -1
    ei = torch.zeros(2, 3, dtype=torch.int64)
    with pytest.raises(ValueError):
        S.dropout_adj(ei, p=1.5)
    with pytest.raises(ValueError):
        S.mask_feature(torch.zeros(2, 4), p=-0.1)
    out, attr = S.dropout_adj(ei, None, p=0.0)                    # PyG: returned unchanged
    assert out is ei and attr is None
    x = torch.zeros(2, 4)
    y, m = S.mask_feature(x, p=0.3, training=False)
    assert y is x and m.dtype == torch.bool and bool(m.all())
    with pytest.raises(S.StemGnnError):
        S.mask_feature(x, p=0.3)                                   # CPU tensor: no fallback


def test_oracle_properties_on_random_cases():
    """Randomised (hypothesis) invariants of the restatements, whatever the draw: dropout_adj keeps a
    sub-multiset of the edges in order (and, undirected, a symmetric list built from row <= col edges only);
    the CSR restatement is a stable grouping; a full-fan-out sample contains every in-edge of its frontier."""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=40, deadline=None)
    @given(st.integers(1, 30), st.integers(0, 120), st.floats(0.0, 1.0), st.integers(0, 2 ** 31 - 1))
    def run(n, e, p, seed):
        g = torch.Generator().manual_seed(seed)
        ei = torch.randint(0, n, (2, e), generator=g)
        rand = torch.rand(e, generator=g)
        out, _, eid = dropout_adj_ref(ei, None, rand, p, False)
        assert torch.equal(out, ei[:, eid]) and (eid[1:] > eid[:-1]).all() and torch.equal(eid, (rand >= p).nonzero().view(-1))
        und, _, eid_u = dropout_adj_ref(ei, None, rand, p, True)
        half = und.size(1) // 2
        assert und.size(1) == 2 * half and torch.equal(und[:, :half], und[:, half:].flip(0))
        assert (und[0, :half] <= und[1, :half]).all() and torch.equal(und[:, :half], ei[:, eid_u[:half]])
        # stable grouping
        rowptr, col, perm, deg = graph_ref.csr_build_ref(ei.numpy(), n, 1)
        assert np.array_equal(perm, np.argsort(ei[1].numpy(), kind="stable")) and np.array_equal(col, ei[0].numpy()[perm])
        assert np.array_equal(deg, np.bincount(ei[1].numpy(), minlength=n)) and rowptr[-1] == e
        # full fan-out, one hop: every in-edge of every seed, seeds first in n_id
        seeds = torch.randperm(n, generator=g)[: max(1, n // 3)]
        n_id, src_l, dst_l, e_id, _ = neighbor_sample_ref(rowptr, col, perm, seeds.tolist(), [-1], None)
        assert n_id[: seeds.numel()].tolist() == seeds.tolist()
        want = torch.isin(ei[1], seeds).nonzero().view(-1)
        assert sorted(e_id.tolist()) == want.tolist()
        assert np.array_equal(n_id[src_l], ei[0].numpy()[e_id]) and np.array_equal(n_id[dst_l], ei[1].numpy()[e_id])
        check_sample(ei.numpy(), n, rowptr, seeds.tolist(), [-1], n_id, src_l, dst_l, e_id)

    run()
