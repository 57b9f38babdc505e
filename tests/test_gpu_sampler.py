"""GPU: K10 neighbour sampler against the sequential restatement of the NeighborLoader sampling
algorithm (oracle/sampler_ref.py).  Everything that is not a random draw is compared bit for bit:
with full fan-outs directly; with random fan-outs by replaying the subsets the GPU drew through the
sequential algorithm.  The draws themselves are checked for validity and uniformity."""
import numpy as np
import pytest
import torch

import stemgnn_b200 as S
from oracle import graph_ref
from oracle.sampler_ref import check_sample, neighbor_sample_ref

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _graph(n, e, seed, skew=False):
    rng = np.random.default_rng(seed)
    src = rng.integers(0, n, e)
    dst = (n * rng.random(e) ** 3).astype(np.int64) if skew else rng.integers(0, n, e)
    ei = np.stack([src, dst])
    ei[1][ei[1] == 3] = 4                                   # node 3 has no in-edges
    return ei


def _replay(ei, n, seeds, fanouts, n_id, edge_index, e_id):
    rowptr, col, perm, _ = graph_ref.csr_build_ref(ei, n, 1)
    slot_of = np.empty(ei.shape[1], np.int64)
    slot_of[perm] = np.arange(ei.shape[1])
    n_id, src, dst, e_id = n_id.cpu().numpy(), edge_index[0].cpu().numpy(), edge_index[1].cpu().numpy(), e_id.cpu().numpy()
    check_sample(ei, n, rowptr, seeds, fanouts, n_id, src, dst, e_id)
    picks = {}
    for t, e in zip(dst, e_id):                              # the subsets the GPU drew, per target
        picks.setdefault(int(t), []).append(int(slot_of[e] - rowptr[n_id[t]]))
    choose = lambda h, i, deg, k: picks[i]
    r_id, r_src, r_dst, r_e, hops = neighbor_sample_ref(rowptr, col, perm, seeds, fanouts, choose)
    assert np.array_equal(n_id, r_id) and np.array_equal(src, r_src)
    assert np.array_equal(dst, r_dst) and np.array_equal(e_id, r_e)
    return hops


@pytest.mark.parametrize("fanouts", [[-1], [-1, -1], [64, 64], [-1, -1, -1]])
def test_full_fanout_is_bit_exact(fanouts):
    n, e = 2000, 12000
    ei = _graph(n, e, 1)
    seeds = np.random.default_rng(2).permutation(n)[:97]
    out = S.neighbor_sample(torch.from_numpy(ei).to(DEV), n, torch.from_numpy(seeds).to(DEV), fanouts, seed=1)
    rowptr, col, perm, _ = graph_ref.csr_build_ref(ei, n, 1)
    r_id, r_src, r_dst, r_e, _ = neighbor_sample_ref(rowptr, col, perm, seeds, fanouts, None)
    assert np.array_equal(out[0].cpu().numpy(), r_id)
    assert np.array_equal(out[1].cpu().numpy(), np.stack([r_src, r_dst]))
    assert np.array_equal(out[2].cpu().numpy(), r_e)


@pytest.mark.parametrize("n,e,b,fanouts,skew", [
    (500, 20000, 64, [10, 10], False),
    (20000, 300000, 1024, [10, 10], True),          # the reference's pretrain loader shape
    (20000, 300000, 512, [30, 30], True),           # link-task loader (utils/loader.py:28)
    (3000, 40000, 100, [5, -1, 3], False),
])
def test_random_fanout_replays_through_sequential_algorithm(n, e, b, fanouts, skew):
    ei = _graph(n, e, 7, skew)
    seeds = np.random.default_rng(3).permutation(n)[:b]
    smp = S.NeighborSampler(torch.from_numpy(ei).to(DEV), n, fanouts, seed=5)
    assert bool((smp.node_map == -2 ** 31).all())
    for _ in range(2):
        n_id, edge_index, e_id = smp.sample(torch.from_numpy(seeds))
        _replay(ei, n, seeds, fanouts, n_id, edge_index, e_id)
        assert bool((smp.node_map == -2 ** 31).all()), "the node map is left initialised"


def test_duplicate_seeds_isolated_seeds_and_empty_inputs():
    n, e = 300, 4000
    ei = _graph(n, e, 9)
    seeds = np.array([5, 3, 5, 7, 3, 3, 11])
    smp = S.NeighborSampler(torch.from_numpy(ei).to(DEV), n, [4, 4], seed=0)
    n_id, edge_index, e_id = smp.sample(torch.from_numpy(seeds))
    _replay(ei, n, seeds, [4, 4], n_id, edge_index, e_id)
    n_id, edge_index, e_id = smp.sample(torch.zeros(0, dtype=torch.int64))
    assert n_id.numel() == 0 and edge_index.shape == (2, 0) and e_id.numel() == 0
    empty = S.NeighborSampler(torch.zeros(2, 0, dtype=torch.int64, device=DEV), 10, [3], seed=0)
    n_id, edge_index, e_id = empty.sample(torch.tensor([1, 2]))
    assert n_id.tolist() == [1, 2] and edge_index.shape == (2, 0)
    with pytest.raises(S.StemGnnError):
        smp.sample(torch.tensor([n + 5]))
    assert bool((smp.node_map == -2 ** 31).all())


def test_same_seed_same_sample_and_subsets_are_uniform():
    n = 64
    deg, k = 20, 5
    src = np.arange(1, deg + 1)
    ei = np.stack([src, np.zeros(deg, np.int64)])            # node 0 has 20 in-edges
    t = torch.from_numpy(ei).to(DEV)
    a = S.neighbor_sample(t, n, torch.tensor([0]), [k], seed=123)
    b = S.neighbor_sample(t, n, torch.tensor([0]), [k], seed=123)
    assert all(torch.equal(x, y) for x, y in zip(a, b))
    smp = S.NeighborSampler(t, n, [k], seed=1)
    trials = 4000
    seeds = torch.zeros(1, dtype=torch.int64)
    hits = np.zeros(deg)
    for _ in range(trials):                                   # the device seed advances per call
        _, _, e_id = smp.sample(seeds)
        assert e_id.numel() == k and len(set(e_id.tolist())) == k
        hits[e_id.cpu().numpy()] += 1
    expect = trials * k / deg
    chi2 = float(((hits - expect) ** 2 / expect).sum())
    assert chi2 < 60, (chi2, hits)                            # 19 dof: P(chi2 > 60) ~ 4e-6


def test_neighbor_loader_slices_attributes_like_pyg():
    n, e, f = 5000, 60000, 32
    ei = _graph(n, e, 4)
    data = dict(edge_index=torch.from_numpy(ei), x=torch.arange(n), xe=torch.arange(e) % 7,
                feat=torch.randn(n, f), node_text_feat=torch.randn(11, f), num_nodes=n)
    idx = torch.arange(0, n, 3)
    loader = S.NeighborLoader(data, num_neighbors=[10, 10], input_nodes=idx, batch_size=512, shuffle=True, seed=2)
    assert len(loader) == -(-idx.numel() // 512)
    seen = []
    for batch in loader:
        bs = batch.batch_size
        seen.append(batch.n_id[:bs].cpu())
        assert torch.equal(batch.x.cpu(), batch.n_id.cpu())                  # node attribute sliced by n_id
        assert torch.equal(batch.xe.cpu(), batch.e_id.cpu() % 7)             # edge attribute sliced by e_id
        assert torch.equal(batch.feat.cpu(), data["feat"][batch.n_id.cpu()])
        assert batch.node_text_feat.shape == (11, f)                         # neither: passed through
        g = torch.from_numpy(ei)[:, batch.e_id.cpu()]
        assert torch.equal(batch.n_id.cpu()[batch.edge_index.cpu()], g)
    assert torch.equal(torch.cat(seen).sort().values, idx)                   # every seed exactly once
