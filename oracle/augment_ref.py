"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the two torch_geometric.utils (2.3.0,
environment.yml:292) augmentations the reference calls at STEM-GNN/pretrain.py:41-44.  The package
is absent from /root/reference and not installable here, so these follow the published functions
(`torch_geometric/utils/dropout.py::dropout_adj`, `utils/mask.py`/`augmentation.py::mask_feature`)
with the random draw passed in: parity unpinned by the reference (it has no tests or fixtures)."""
import torch


def dropout_adj_ref(edge_index, edge_attr, rand, p, force_undirected):
    """rand plays `torch.rand(row.size(0))`.  Returns (edge_index, edge_attr, edge_id)."""
    row, col = edge_index
    mask = rand >= p
    if force_undirected:
        mask = mask.clone()
        mask[row > col] = False
    eid = mask.nonzero().view(-1)
    row, col = row[mask], col[mask]                       # filter_adj
    if edge_attr is not None:
        edge_attr = edge_attr[mask]
    if force_undirected:
        out = torch.stack([torch.cat([row, col]), torch.cat([col, row])])
        if edge_attr is not None:
            edge_attr = torch.cat([edge_attr, edge_attr])
        eid = torch.cat([eid, eid])
    else:
        out = torch.stack([row, col])
    return out, edge_attr, eid


def mask_feature_ref(x, mask, fill_value=0.0):
    """`x.masked_fill(~mask, fill_value)` with mask broadcast as [1,F] (col), [N,1] (row) or [N,F]."""
    return x.masked_fill(~mask, fill_value)


def check_negative_edges(edge_index, num_nodes, neg):
    """Properties every output of torch_geometric.utils.negative_sampling (2.3.0) has, whatever its
    random draws were: no self loops, no positive edge, no pair twice, ids in range."""
    import numpy as np
    ei, neg = np.asarray(edge_index), np.asarray(neg)
    assert neg.shape[0] == 2
    assert ((neg >= 0) & (neg < num_nodes)).all()
    assert (neg[0] != neg[1]).all(), "self loop drawn"
    pos = set((ei[0] * num_nodes + ei[1]).tolist())
    ids = neg[0] * num_nodes + neg[1]
    assert not pos.intersection(ids.tolist()), "a positive edge was returned"
    assert len(np.unique(ids)) == ids.size, "a pair was returned twice"
