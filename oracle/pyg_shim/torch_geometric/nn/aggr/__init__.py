"""TEST INFRASTRUCTURE ONLY -- PyG 2.3.0 nn.aggr.basic semantics: scatter(..., reduce) =
torch_scatter-style scatter_add_ followed (for 'mean') by / count.clamp(min=1)."""
import torch
import torch.nn as nn


class Aggregation(nn.Module):
    def forward(self, x, index, dim_size=None):
        raise NotImplementedError


class SumAggregation(Aggregation):
    def forward(self, x, index, dim_size=None):
        out = torch.zeros(dim_size, x.size(1), dtype=x.dtype)
        return out.scatter_add_(0, index.view(-1, 1).expand_as(x), x)


class MeanAggregation(Aggregation):
    def forward(self, x, index, dim_size=None):
        out = torch.zeros(dim_size, x.size(1), dtype=x.dtype)
        out = out.scatter_add_(0, index.view(-1, 1).expand_as(x), x)
        count = torch.zeros(dim_size, dtype=x.dtype)
        count.scatter_add_(0, index, torch.ones(index.numel(), dtype=x.dtype))
        return out / count.clamp(min=1).view(-1, 1)


class MultiAggregation(Aggregation):
    pass
