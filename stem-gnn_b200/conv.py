"""Drop-in conv layers of the STEM-GNN encoder (reference model/encoder.py:17-129 and the PyG
GCNConv / GINConv it instantiates at :197/:199), with the propagate step replaced by the fused
sm_100a aggregation kernels.  Constructor / forward signatures and state_dict keys follow the
reference so checkpoints and the unchanged pretrain.py / finetune.py keep working."""
from typing import Optional, Tuple, Union

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch import Tensor

from .graph import GraphStructure, structure_of
from .partition import PartitionedGraph
import os

from . import dense, sage
from .ops import aggregate

_SAGE_FUSED = os.environ.get("STEMGNN_SAGE_FUSED", "1") != "0"      # A/B switch: 0 = the composed form


def _sparse_coo(adj_t):
    """(row, col) int64 of a sparse adjacency in row-major order, or None if `adj_t` is not one.  Accepted: torch
    sparse CSR / COO tensors and anything with torch_sparse.SparseTensor's `.coo()` (row, col, value sorted by
    (row, col)).  Values are dropped: the reference's fused path removes them (`adj_t.set_value(None)`,
    model/encoder.py:99-102)."""
    if isinstance(adj_t, Tensor):
        if adj_t.layout == torch.sparse_csr:
            crow, col = adj_t.crow_indices(), adj_t.col_indices()
            row = torch.repeat_interleave(torch.arange(crow.numel() - 1, device=col.device), crow[1:] - crow[:-1])
            return row.to(torch.int64), col.to(torch.int64)
        if adj_t.layout == torch.sparse_coo:
            idx = adj_t.coalesce().indices()
            return idx[0], idx[1]
        return None
    coo = getattr(adj_t, "coo", None)
    if callable(coo):
        row, col = coo()[:2]
        return row.to(torch.int64), col.to(torch.int64)
    return None


def _graph(edge_index, num_nodes):
    """The aggregation structure of whatever the caller passed as `edge_index` (PyG's `Adj`): an int64 [2, E]
    tensor, a GraphStructure / PartitionedGraph, or a sparse TRANSPOSED adjacency `adj_t` (rows = targets, columns
    = sources: what PyG's ToSparseTensor produces and `message_and_aggregate` consumes, model/encoder.py:99-102).
    A sparse adjacency becomes the edge list (col -> row) in its row-major order, so the K0 structure sums a row's
    neighbours in column order like spmm does, and carries `fused = True`: PyG then skips `message` (no relu, no
    edge features) and computes spmm(adj_t, x, reduce)."""
    if isinstance(edge_index, (GraphStructure, PartitionedGraph)):
        return edge_index
    if isinstance(edge_index, Tensor) and edge_index.layout == torch.strided:
        return structure_of(edge_index, num_nodes)
    cached = getattr(edge_index, "_stemgnn_structure", None)
    if cached is not None and cached.num_nodes == int(num_nodes):
        return cached
    rc = _sparse_coo(edge_index)
    if rc is None:
        raise TypeError("edge_index must be an int64 [2, E] tensor, a GraphStructure / PartitionedGraph, or a "
                        f"sparse adjacency (torch sparse CSR / COO, torch_sparse.SparseTensor); got {type(edge_index)}")
    g = GraphStructure(torch.stack([rc[1], rc[0]]).contiguous(), num_nodes)
    g.fused = True
    try:
        edge_index._stemgnn_structure = g          # torch_sparse.SparseTensor is a plain Python object
    except (AttributeError, RuntimeError):
        pass
    return g


class MySAGEConv(nn.Module):
    """out = lin_l(mean_{j->i} relu(x_j [+ xe])) + lin_r(x_i)   (model/encoder.py:72-97)."""

    def __init__(self, in_channels: Union[int, Tuple[int, int]], out_channels: int,
                 aggr: Optional[str] = "mean", normalize: bool = False, root_weight: bool = True,
                 project: bool = False, bias: bool = True, **kwargs):
        super().__init__()
        if kwargs:
            raise NotImplementedError(f"MessagePassing options {sorted(kwargs)} are not supported")
        if aggr not in ("mean", "sum", "add"):
            raise NotImplementedError(f"aggr={aggr!r}: only 'mean' and 'sum'/'add' are on the hot path")
        self.in_channels, self.out_channels = in_channels, out_channels
        self.aggr, self.normalize, self.root_weight, self.project = aggr, normalize, root_weight, project
        if isinstance(in_channels, int):
            in_channels = (in_channels, in_channels)
        if in_channels[0] != in_channels[1]:
            raise NotImplementedError("bipartite (x_src, x_dst) inputs are not supported")
        if project:
            self.lin = nn.Linear(in_channels[0], in_channels[0], bias=True)
        self.lin_l = nn.Linear(in_channels[0], out_channels, bias=bias)
        if root_weight:
            self.lin_r = nn.Linear(in_channels[1], out_channels, bias=False)
        self.reset_parameters()

    def reset_parameters(self):
        if self.project:
            self.lin.reset_parameters()
        self.lin_l.reset_parameters()
        if self.root_weight:
            self.lin_r.reset_parameters()

    def forward(self, x: Tensor, edge_index, edge_attr: Optional[Tensor] = None, size=None) -> Tensor:
        if isinstance(x, (tuple, list)):
            if x[1] is not None and x[1] is not x[0]:
                raise NotImplementedError("bipartite (x_src, x_dst) inputs are not supported")
            x = x[0]
        if size is not None and tuple(size) != (x.size(0), x.size(0)):
            raise NotImplementedError("size must be None or (N, N)")
        x_r = x
        if self.project:
            x = dense.linear_sum([x], [self.lin.weight], self.lin.bias).relu()
        g = _graph(edge_index, x.size(0))
        if x.size(1) % 4:
            # a width like Cora's 1,433: pad the features (and per-edge features) once, so the aggregation runs the
            # 128-bit kernels and the padded columns -- zeros all the way -- meet zero weight columns in the GEMM
            x_r = dense.pad4(x_r)
            x = dense.pad4(x) if self.project else x_r
            if isinstance(edge_attr, tuple):
                edge_attr = (dense.pad4(edge_attr[0]), edge_attr[1])
            elif edge_attr is not None:
                edge_attr = dense.pad4(edge_attr)
        if (_SAGE_FUSED and not self.project and x_r is x and x.size(1) == self.lin_l.weight.size(1)
                and sage.fusable(x, g, edge_attr, self.out_channels, self.root_weight)):
            # the whole layer as one autograd node (sage.py): count division and gradient accumulation ride on
            # kernel epilogues, a partitioned graph overlaps its halo exchange with the dense products
            out = sage.sage_layer(x, g, self.aggr, self.lin_l, self.lin_r if self.root_weight else None)
            return F.normalize(out, p=2.0, dim=-1) if self.normalize else out
        if getattr(g, "fused", False):
            # sparse adj_t: PyG's propagate takes message_and_aggregate = spmm(adj_t, x, reduce) and never calls
            # message(): no relu, xe unused (model/encoder.py:99-102)
            out = aggregate(x, g, reduce=self.aggr, act="none")
        elif isinstance(edge_attr, tuple):          # (type_table [T,F], edge_type [E]) fast form
            out = aggregate(x, g, reduce=self.aggr, act="relu", type_table=edge_attr[0],
                            edge_type=edge_attr[1])
        else:
            out = aggregate(x, g, reduce=self.aggr, act="relu", edge_attr=edge_attr)
        # lin_l(out) + lin_r(x_r) as ONE tensor-core GEMM over the concatenated K (K5/K6)
        if self.root_weight:
            out = dense.linear_sum([out, x_r], [self.lin_l.weight, self.lin_r.weight], self.lin_l.bias)
        else:
            out = dense.linear_sum([out], [self.lin_l.weight], self.lin_l.bias)
        if self.normalize:
            out = F.normalize(out, p=2.0, dim=-1)
        return out

    def __repr__(self):
        return f"{self.__class__.__name__}({self.in_channels}, {self.out_channels}, aggr={self.aggr})"


class GCNConv(nn.Module):
    """PyG GCNConv(in, out) as built at model/encoder.py:197: x W, symmetric-normalised sum over
    the edges plus remaining self loops, + bias."""

    def __init__(self, in_channels: int, out_channels: int, improved: bool = False,
                 cached: bool = False, add_self_loops: bool = True, normalize: bool = True,
                 bias: bool = True, **kwargs):
        super().__init__()
        if improved or not add_self_loops or not normalize or kwargs:
            raise NotImplementedError("only the default GCNConv(in, out) configuration is supported")
        self.in_channels, self.out_channels = in_channels, out_channels
        self.lin = nn.Linear(in_channels, out_channels, bias=False)
        if bias:
            self.bias = nn.Parameter(torch.empty(out_channels))
        else:
            self.register_parameter("bias", None)
        self.reset_parameters()

    def reset_parameters(self):
        nn.init.xavier_uniform_(self.lin.weight)
        if self.bias is not None:
            nn.init.zeros_(self.bias)

    def forward(self, x: Tensor, edge_index, edge_weight: Optional[Tensor] = None) -> Tensor:
        if edge_weight is not None:
            raise NotImplementedError("GCNConv with explicit edge_weight is not on the hot path")
        g = _graph(edge_index, x.size(0))
        if isinstance(g, PartitionedGraph):
            raise NotImplementedError("GCNConv on a partitioned graph (symmetric normalisation needs "
                                      "the degrees of remote sources) is not built yet")
        g_sl, w = g.gcn()
        h = dense.linear_sum([x], [self.lin.weight], None)            # K5/K6 (tcgen05, fp32 accuracy)
        fo = h.size(1)
        out = aggregate(dense.pad4(h), g_sl, reduce="sum", edge_weight=w)
        out = out[:, :fo] if out.size(1) != fo else out
        return out + self.bias if self.bias is not None else out


class GINConv(nn.Module):
    """PyG GINConv(nn) as built at model/encoder.py:199: nn((1 + eps) x_i + sum_j x_j)."""

    def __init__(self, nn_module: nn.Module, eps: float = 0.0, train_eps: bool = False, **kwargs):
        super().__init__()
        if kwargs:
            raise NotImplementedError(f"GINConv options {sorted(kwargs)} are not supported")
        self.nn = nn_module
        self.initial_eps = eps
        if train_eps:
            self.eps = nn.Parameter(torch.empty(1))
        else:
            self.register_buffer("eps", torch.empty(1))
        self.reset_parameters()

    def reset_parameters(self):
        for m in self.nn.modules():
            if m is not self.nn and hasattr(m, "reset_parameters"):
                m.reset_parameters()
        if hasattr(self.nn, "reset_parameters"):
            self.nn.reset_parameters()
        self.eps.data.fill_(self.initial_eps)

    def forward(self, x: Tensor, edge_index, size=None) -> Tensor:
        if size is not None and not isinstance(size, (tuple, list)):
            raise NotImplementedError("GINConv takes no edge features (model/encoder.py:311 passes "
                                      "edge_attr as `size`; use edge_attr=None with backbone='gin')")
        f = x.size(1)
        out = aggregate(dense.pad4(x), _graph(edge_index, x.size(0)), reduce="sum")
        out = out[:, :f] if out.size(1) != f else out
        return self.nn(out + (1 + self.eps) * x)


class MixtureSageLayer(nn.Module):
    """model/encoder.py:109-129.  Aggregates x[edge_index[1]] AT edge_index[0] (the reverse of
    MySAGEConv), no relu, no edge features."""

    def __init__(self, in_dim: int, out_dim: int, num_experts: int, residual: bool = True):
        super().__init__()
        self.in_dim, self.out_dim, self.num_experts = in_dim, out_dim, num_experts
        self.residual = residual and (in_dim == out_dim)
        self.weights = nn.Parameter(torch.empty(num_experts, in_dim * 2, out_dim))
        self.reset_parameters()

    def reset_parameters(self):
        nn.init.xavier_uniform_(self.weights)

    def forward(self, x: Tensor, edge_index, edge_attr: Optional[Tensor] = None) -> Tensor:
        xp = dense.pad4(x)
        agg = aggregate(xp, _graph(edge_index, x.size(0)), to_row=0, reduce="mean")
        # einsum('nd,kdo->nko', [agg|x], W) for all experts as ONE two-source GEMM (K5/K6): the
        # experts are stacked along the output dimension, [agg|x] is never concatenated
        w_all = self.weights.permute(0, 2, 1).reshape(self.num_experts * self.out_dim, 2 * self.in_dim)
        outputs = dense.linear_sum([agg, xp], [w_all[:, :self.in_dim], w_all[:, self.in_dim:]], None)
        outputs = outputs.reshape(x.size(0), self.num_experts, self.out_dim)
        if self.residual:
            outputs = outputs + x.unsqueeze(1)
        return outputs
