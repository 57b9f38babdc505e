"""TEST INFRASTRUCTURE ONLY -- minimal stand-in for the torch_geometric 2.3.0 symbols that
/root/reference/STEM-GNN/model/encoder.py:9-14 imports, so the UNMODIFIED reference
encoder.py can execute on CPU in the build container (torch_geometric is not installed and
cannot be: no network).  Written from the published PyG 2.3.0 semantics ("PyG-knowledge",
SURVEY.md section 8c); used only by tests/golden/gen_golden.py to generate golden vectors.
"""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from .aggr import Aggregation, MeanAggregation, MultiAggregation, SumAggregation
from .dense.linear import Linear


def global_mean_pool(x, batch, size=None):
    size = int(batch.max()) + 1 if size is None else size
    out = torch.zeros(size, x.size(1), dtype=x.dtype, device=x.device).index_add_(0, batch, x)
    cnt = torch.zeros(size, dtype=x.dtype, device=x.device).index_add_(0, batch, torch.ones_like(batch, dtype=x.dtype))
    return out / cnt.clamp(min=1).view(-1, 1)


def global_add_pool(x, batch, size=None):
    size = int(batch.max()) + 1 if size is None else size
    return torch.zeros(size, x.size(1), dtype=x.dtype, device=x.device).index_add_(0, batch, x)


def global_max_pool(x, batch, size=None):
    raise NotImplementedError


class MessagePassing(nn.Module):
    """propagate = _collect (x_j = x[0].index_select(0, edge_index[0])) -> message ->
    aggregate at edge_index[1] with dim_size = N (flow source_to_target)."""

    def __init__(self, aggr="add", **kwargs):
        super().__init__()
        self.aggr = aggr
        if aggr in ("add", "sum"):
            self.aggr_module = SumAggregation()
        elif aggr == "mean":
            self.aggr_module = MeanAggregation()
        else:
            raise NotImplementedError(f"pyg_shim: aggr={aggr!r}")
        self.fuse = True

    def reset_parameters(self):
        pass

    def propagate(self, edge_index, size=None, **kwargs):
        x = kwargs.pop("x")
        if isinstance(x, torch.Tensor):
            x = (x, x)
        x_j = x[0].index_select(0, edge_index[0])
        n = x[1].size(0) if x[1] is not None else x[0].size(0)
        msg_kwargs = dict(kwargs)
        import inspect
        params = inspect.signature(self.message).parameters
        call = {}
        for name in params:
            if name == "x_j":
                call[name] = x_j
            elif name in msg_kwargs:
                call[name] = msg_kwargs[name]
        m = self.message(**call)
        return self.aggr_module(m, edge_index[1], dim_size=n)

    def message(self, x_j):
        return x_j


class GCNConv(MessagePassing):
    """PyG 2.3.0 GCNConv: lin (no bias) -> gcn_norm -> weighted sum -> + bias."""

    def __init__(self, in_channels, out_channels, bias=True, **kwargs):
        super().__init__(aggr="add")
        self.lin = Linear(in_channels, out_channels, bias=False, weight_initializer="glorot")
        self.bias = nn.Parameter(torch.zeros(out_channels)) if bias else None
        self.reset_parameters()

    def reset_parameters(self):
        self.lin.reset_parameters()
        if self.bias is not None:
            nn.init.zeros_(self.bias)

    def forward(self, x, edge_index, edge_weight=None):
        assert edge_weight is None, "pyg_shim GCNConv: edge_weight unsupported"
        from oracle.propagate_ref import gcn_norm_ref
        n = x.size(0)
        x = self.lin(x)
        ei, w = gcn_norm_ref(edge_index, n, x.dtype)
        out = self.propagate(ei, x=x, edge_weight=w)
        if self.bias is not None:
            out = out + self.bias
        return out

    def message(self, x_j, edge_weight):
        return edge_weight.view(-1, 1) * x_j


class GINConv(MessagePassing):
    """PyG 2.3.0 GINConv(nn, eps=0, train_eps=False): nn((1+eps)*x_r + sum_j x_j)."""

    def __init__(self, nn_module, eps=0.0, train_eps=False, **kwargs):
        super().__init__(aggr="add")
        self.nn = nn_module
        self.initial_eps = eps
        self.register_buffer("eps", torch.Tensor([eps]))

    def reset_parameters(self):
        if hasattr(self.nn, "reset_parameters"):
            self.nn.reset_parameters()
        self.eps.data.fill_(self.initial_eps)

    def forward(self, x, edge_index, size=None):
        out = self.propagate(edge_index, x=(x, x))
        out = out + (1 + self.eps) * x
        return self.nn(out)


class GATConv(MessagePassing):
    def __init__(self, *a, **k):
        raise NotImplementedError("pyg_shim: GATConv is outside the north_star path")


class SAGEConv(MessagePassing):
    def __init__(self, *a, **k):
        raise NotImplementedError("pyg_shim: SAGEConv is imported by encoder.py:9 but never used")
