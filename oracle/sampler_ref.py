"""TEST INFRASTRUCTURE ONLY -- sequential restatement of the neighbour sampler underneath
torch_geometric.loader.NeighborLoader 2.3.0 (torch_sparse 0.6.15 `neighbor_sample`, replace=False,
directed=True; environment.yml:292,315), which the reference drives at STEM-GNN/pretrain.py:151-153
and utils/loader.py:11-24.  Neither package is under /root/reference or installable here: the
algorithm below is the published one (a node list grown hop by hop with a node -> local-id hash map,
sampled edges appended in frontier order) -- parity unpinned by the reference.  The random subset of
a node with more in-edges than the fan-out comes from the C++ RNG there; here it is a callback, so a
test can replay the subsets another implementation drew and compare everything that is NOT random:
node order, relabelling, edge order, edge ids."""
import numpy as np


def neighbor_sample_ref(rowptr, col, perm, seeds, fanouts, choose):
    """rowptr/col/perm: CSR grouped by destination (oracle.graph_ref.csr_build_ref(..., 1)).
    choose(hop, local_target, deg, k) -> k distinct slot offsets in [0, deg), in emission order.
    Returns (n_id, src_local, dst_local, e_id, cumulative (nodes, edges) per hop)."""
    n_id = [int(s) for s in seeds]
    to_local = {}
    for i, v in enumerate(n_id):
        to_local.setdefault(v, i)
    src, dst, e_id, per_hop = [], [], [], []
    begin, end = 0, len(n_id)
    for h, k in enumerate(fanouts):
        for i in range(begin, end):
            v = n_id[i]
            p0, deg = int(rowptr[v]), int(rowptr[v + 1] - rowptr[v])
            if deg == 0:
                continue
            offs = range(deg) if (k < 0 or deg <= k) else choose(h, i, deg, k)
            for t in offs:
                u = int(col[p0 + t])
                if u not in to_local:
                    to_local[u] = len(n_id)
                    n_id.append(u)
                src.append(to_local[u])
                dst.append(i)
                e_id.append(int(perm[p0 + t]))
        begin, end = end, len(n_id)
        per_hop.append((len(n_id), len(src)))
    return (np.array(n_id, np.int64), np.array(src, np.int64), np.array(dst, np.int64),
            np.array(e_id, np.int64), per_hop)


def check_sample(edge_index, num_nodes, rowptr, seeds, fanouts, n_id, src, dst, e_id):
    """Properties every valid sample has, whatever the random subsets were."""
    ei = np.asarray(edge_index)
    n_id, src, dst, e_id = map(np.asarray, (n_id, src, dst, e_id))
    B = len(seeds)
    assert np.array_equal(n_id[:B], np.asarray(seeds)), "seeds first, in order"
    if len(np.unique(seeds)) == B:
        assert len(np.unique(n_id)) == len(n_id), "nodes are unique"
    # every sampled edge is the original edge e_id, expressed in local ids
    assert np.array_equal(ei[0][e_id], n_id[src]) and np.array_equal(ei[1][e_id], n_id[dst])
    # per target: min(deg, k) distinct edges, targets visited in increasing local order
    assert np.all(np.diff(dst) >= 0), "edges are grouped by target in frontier order"
    deg = np.diff(np.asarray(rowptr)).astype(np.int64)
    begin, end, epos = 0, B, 0
    for k in fanouts:
        for i in range(begin, end):
            want = deg[n_id[i]] if (k < 0 or deg[n_id[i]] <= k) else k
            got = e_id[epos:epos + want]
            assert np.all(dst[epos:epos + want] == i)
            assert len(np.unique(got)) == want, "edges of a target are distinct"
            epos += want
        new_end = int(src[:epos].max()) + 1 if epos else end
        begin, end = end, max(end, new_end)
    assert epos == len(e_id) and end == len(n_id)
