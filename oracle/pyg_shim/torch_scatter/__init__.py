"""TEST INFRASTRUCTURE ONLY -- torch_scatter 2.1.0 scatter_mean: scatter_add_ then divide by
count.clamp_(1) (count accumulated in the source dtype)."""
import torch


def scatter_mean(src, index, dim=0, out=None, dim_size=None):
    assert dim == 0 and out is None
    res = torch.zeros((dim_size,) + tuple(src.shape[1:]), dtype=src.dtype)
    res.scatter_add_(0, index.view(-1, *([1] * (src.dim() - 1))).expand_as(src), src)
    count = torch.zeros(dim_size, dtype=src.dtype)
    count.scatter_add_(0, index, torch.ones(index.numel(), dtype=src.dtype))
    count.clamp_(1)
    return res / count.view(-1, *([1] * (src.dim() - 1)))
