"""TEST INFRASTRUCTURE ONLY -- the torch_geometric.utils names the reference's model/ package imports
(encoder.py:14 spmm, pt_model.py:6 negative_sampling)."""
import torch

# tests swap this to feed both arms of a comparison the same negative edges (a callable with
# negative_sampling's signature); None = the restatement below
NEGATIVE_SAMPLING_HOOK = None


def spmm(*a, **k):
    raise NotImplementedError("pyg_shim: spmm path (encoder.py:99-102) needs a SparseTensor; unused")


def negative_sampling(edge_index, num_nodes=None, num_neg_samples=None, method="sparse", force_undirected=False):
    """torch_geometric.utils.negative_sampling (2.3.0), structured mode off: up to `num_neg_samples`
    (default E) distinct pairs (i, j), i != j, that are not edges.  PyG draws with python's
    random.sample over the N*N-N pair ids and filters with numpy.isin, up to three times; the draws
    cannot be reproduced, the output properties can (oracle/augment_ref.check_negative_edges)."""
    if NEGATIVE_SAMPLING_HOOK is not None:
        return NEGATIVE_SAMPLING_HOOK(edge_index, num_nodes, num_neg_samples)
    assert not force_undirected
    n = int(num_nodes) if num_nodes is not None else int(edge_index.max()) + 1
    e = edge_index.size(1)
    want = e if num_neg_samples is None else int(num_neg_samples)
    pop = n * n - n
    want = min(want, pop - e)
    if want <= 0:
        return edge_index.new_empty((2, 0))
    i, j = edge_index[0], edge_index[1]
    pos = torch.unique(i * (n - 1) + j - (i < j).to(i.dtype))[: e]          # pair ids without the diagonal
    got = edge_index.new_empty(0)
    for _ in range(3):
        cand = torch.randint(0, pop, (int(1.2 * want / max(1.0 - e / pop, 1e-3)) + 16,))
        cand = cand[~torch.isin(cand, pos)]
        got = torch.unique(torch.cat([got, cand]))
        if got.numel() >= want:
            break
    got = got[torch.randperm(got.numel())[:want]]
    r, c = got // (n - 1), got % (n - 1)
    c = c + (c >= r).to(c.dtype)
    return torch.stack([r, c]).to(edge_index.device)
