"""TEST INFRASTRUCTURE ONLY -- integer oracle for the graph-structure work (K0).

The reference never builds a CSR explicitly: PyG's `propagate` scatters in COO edge order
(encoder.py:82 -> MessagePassing.aggregate -> scatter_add_).  The order in which messages are
added to a destination row is therefore "original edge order".  A CSR that groups edges by
destination with a STABLE sort reproduces that order, which is what these functions define
(SURVEY.md section 8c "Integer oracle").
"""
import numpy as np


def csr_build_ref(edge_index, num_nodes, group_by_row):
    """Group the E edges by `edge_index[group_by_row]`, keeping original edge order inside a
    group (stable).  Returns (rowptr[N+1] int32, col[E] int32, perm[E] int32, deg[N] int32)
    where perm[p] is the original edge id stored at CSR slot p and col[p] is the OTHER
    endpoint of that edge.

    group_by_row=1: rows are destinations edge_index[1], col = sources  (encoder.py:82, PyG
                    source_to_target flow: j = edge_index[0], i = edge_index[1]).
    group_by_row=0: rows are edge_index[0], col = edge_index[1] (the transposed structure
                    used by the backward pass, and the forward of MixtureSageLayer,
                    encoder.py:123-124).
    """
    ei = np.asarray(edge_index, dtype=np.int64)
    assert ei.ndim == 2 and ei.shape[0] == 2
    key = ei[group_by_row]
    other = ei[1 - group_by_row]
    perm = np.argsort(key, kind="stable")
    deg = np.bincount(key, minlength=num_nodes).astype(np.int64)
    rowptr = np.zeros(num_nodes + 1, dtype=np.int64)
    np.cumsum(deg, out=rowptr[1:])
    col = other[perm]
    return (rowptr.astype(np.int32), col.astype(np.int32), perm.astype(np.int32),
            deg.astype(np.int32))


def add_remaining_self_loops_ref(edge_index, num_nodes):
    """PyG 2.3.0 `add_remaining_self_loops` with edge_attr=None as used by GCNConv.gcn_norm
    (encoder.py:197): keep the non-loop edges in order, then append one self loop per node
    (arange(N)).  Existing self loops are dropped from the front part and re-created in the
    appended block (with fill weight 1 when no weights are given)."""
    ei = np.asarray(edge_index, dtype=np.int64)
    mask = ei[0] != ei[1]
    loops = np.arange(num_nodes, dtype=np.int64)
    return np.concatenate([ei[:, mask], np.stack([loops, loops])], axis=1)


def partition_1d_ref(num_nodes, world_size):
    """1-D contiguous partition by destination node: rank r owns
    [r*ceil(N/P), min(N, (r+1)*ceil(N/P)))."""
    per = -(-num_nodes // world_size)
    return [(min(num_nodes, r * per), min(num_nodes, (r + 1) * per)) for r in range(world_size)]
