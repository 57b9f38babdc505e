"""TEST INFRASTRUCTURE ONLY -- op-for-op CPU restatement of the aggregation the reference
runs through PyG (not vendored under /root/reference; torch_geometric 2.3.0 and
torch_scatter 2.1.0, environment.yml:292,314).  Everything is plain differentiable torch so
autograd provides the backward oracle as well.

Summation order: `zeros.scatter_add_(0, idx, m)` on CPU adds the messages of a destination
one by one in original edge order (verified bit-for-bit against a sequential C loop in
tests/test_oracle_golden.py), which is the order the CUDA kernels reproduce.
"""
import torch
import torch.nn.functional as F


def _scatter_sum(src, index, dim_size):
    # torch_scatter.scatter_sum -> out.scatter_add_(0, index.expand_as(src), src)
    out = torch.zeros((dim_size,) + tuple(src.shape[1:]), dtype=src.dtype)
    idx = index.view(-1, *([1] * (src.dim() - 1))).expand_as(src)
    return out.scatter_add_(0, idx, src)


def scatter_mean_ref(src, index, dim_size):
    """torch_scatter.scatter_mean (encoder.py:124) and PyG MeanAggregation (encoder.py:43 via
    aggr='mean' at encoder.py:193): sum / clamp(count, 1), count accumulated in the feature
    dtype."""
    out = _scatter_sum(src, index, dim_size)
    ones = torch.ones(index.numel(), dtype=src.dtype)
    count = torch.zeros(dim_size, dtype=src.dtype).scatter_add_(0, index, ones)
    count = count.clamp(min=1)
    return out / count.view(-1, *([1] * (src.dim() - 1)))


def sage_propagate_ref(x, edge_index, xe=None, num_nodes=None, aggr="mean"):
    """MySAGEConv.propagate (encoder.py:82) with message = relu(x_j [+ xe]) (encoder.py:94-97):
    x_j = x[edge_index[0]], aggregated at edge_index[1]."""
    n = x.size(0) if num_nodes is None else num_nodes
    x_j = x.index_select(0, edge_index[0])
    if xe is not None:
        x_j = x_j + xe
    m = F.relu(x_j)
    if aggr == "mean":
        return scatter_mean_ref(m, edge_index[1], n)
    if aggr in ("sum", "add"):
        return _scatter_sum(m, edge_index[1], n)
    raise ValueError(aggr)


def mixture_aggregate_ref(x, edge_index):
    """MixtureSageLayer aggregation (encoder.py:123-124): scatter_mean(x[col], row) -- the
    REVERSE direction of MySAGEConv, no relu, no edge features."""
    row, col = edge_index[0], edge_index[1]
    return scatter_mean_ref(x[col], row, x.size(0))


def gin_aggregate_ref(x, edge_index):
    """GINConv.propagate (encoder.py:199): plain sum of x_j at the destination."""
    return _scatter_sum(x.index_select(0, edge_index[0]), edge_index[1], x.size(0))


def gcn_norm_ref(edge_index, num_nodes, dtype=torch.float32):
    """PyG 2.3.0 gcn_norm(edge_index, None, N, improved=False, add_self_loops=True,
    flow='source_to_target') as called from GCNConv.forward (encoder.py:197): remaining self
    loops with weight 1, deg = scatter_add(w, dst), w_e = deg^-1/2[src] * w * deg^-1/2[dst],
    inf -> 0."""
    mask = edge_index[0] != edge_index[1]
    loops = torch.arange(num_nodes, dtype=edge_index.dtype)
    ei = torch.cat([edge_index[:, mask], torch.stack([loops, loops])], dim=1)
    w = torch.ones(ei.size(1), dtype=dtype)
    row, col = ei[0], ei[1]
    deg = torch.zeros(num_nodes, dtype=dtype).scatter_add_(0, col, w)
    dis = deg.pow(-0.5)
    dis.masked_fill_(dis == float("inf"), 0)
    w = dis[row] * w * dis[col]
    return ei, w


def gcn_propagate_ref(xw, edge_index, num_nodes):
    """GCNConv.propagate: out[i] = sum_e w_e * xw[src_e] (message = edge_weight.view(-1,1) * x_j,
    aggr='add'); the multiply and the add are separate roundings."""
    ei, w = gcn_norm_ref(edge_index, num_nodes, xw.dtype)
    m = w.view(-1, 1) * xw.index_select(0, ei[0])
    return _scatter_sum(m, ei[1], num_nodes)
