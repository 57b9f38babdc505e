"""GPU: a sparse transposed adjacency `adj_t` in place of edge_index -- PyG's fused path
`message_and_aggregate(adj_t, x) = spmm(adj_t.set_value(None), x[0], reduce=self.aggr)` (reference
model/encoder.py:99-102): no relu, no edge features, values dropped."""
import pytest
import torch

import stemgnn_b200 as S

pytestmark = pytest.mark.gpu
DEV = "cuda"


class _DuckSparseTensor:
    """What the layers use of torch_sparse.SparseTensor (absent from this image): `.coo()` -> (row, col, value),
    sorted by (row, col)."""

    def __init__(self, row, col, value):
        self.row, self.col, self.value = row, col, value

    def coo(self):
        return self.row, self.col, self.value


def _adj(n, e, seed):
    g = torch.Generator().manual_seed(seed)
    src, dst = torch.randint(0, n, (e,), generator=g), torch.randint(0, n, (e,), generator=g)
    dst[:300] = 3                                               # a hub target
    val = torch.rand(e, generator=g) + 0.5                      # values the fused path must ignore
    return torch.sparse_coo_tensor(torch.stack([dst, src]), val, (n, n)).coalesce()     # rows = targets


@pytest.mark.parametrize("aggr", ["mean", "sum"])
@pytest.mark.parametrize("fmt", ["coo", "csr", "duck"])
def test_sage_conv_on_sparse_adj_t_is_unweighted_spmm(aggr, fmt):
    n, e, f, o = 400, 3000, 32, 48
    adj = _adj(n, e, 1)
    row, col = adj.indices()
    A = torch.zeros(n, n)
    A[row, col] = 1.0                                           # set_value(None): the pattern only
    torch.manual_seed(0)
    conv = S.MySAGEConv(f, o, aggr=aggr).to(DEV)
    x = torch.randn(n, f)
    go = torch.randn(n, o)
    adj_in = {"coo": lambda: adj.to(DEV), "csr": lambda: adj.to_sparse_csr().to(DEV),
              "duck": lambda: _DuckSparseTensor(row.to(DEV), col.to(DEV), adj.values().to(DEV))}[fmt]()
    xg = x.to(DEV).requires_grad_(True)
    out = conv(xg, adj_in, edge_attr=torch.randn(adj._nnz(), f, device=DEV))     # xe never reaches the fused path
    out.backward(go.to(DEV))
    xr = x.double().requires_grad_(True)
    agg = A.double() @ xr
    if aggr == "mean":
        agg = agg / A.sum(1).clamp(min=1).double().view(-1, 1)
    ref = agg @ conv.lin_l.weight.detach().cpu().double().t() + conv.lin_l.bias.detach().cpu().double() \
        + xr @ conv.lin_r.weight.detach().cpu().double().t()
    ref.backward(go.double())
    assert ((out.detach().cpu().double() - ref.detach()).abs().max() / ref.detach().abs().max()).item() < 1e-5
    assert ((xg.grad.cpu().double() - xr.grad).abs().max() / xr.grad.abs().max()).item() < 1e-5
    # a second call reuses the structure cached on the adjacency object
    out2 = conv(xg.detach(), adj_in)
    assert torch.equal(out2, out.detach())


def test_sparse_adj_t_equals_the_edge_list_form_without_relu():
    n, e, f = 500, 4000, 64
    adj = _adj(n, e, 2)
    row, col = adj.indices()
    x = torch.randn(n, f).to(DEV)
    gs = S.GraphStructure(torch.stack([col, row]).to(DEV), n)           # edge j -> i for adj_t[i, j]
    want = S.aggregate(x, gs, reduce="mean", act="none")
    from stemgnn_b200.conv import _graph
    g = _graph(adj.to_sparse_csr().to(DEV), n)
    assert g.fused
    assert torch.equal(S.aggregate(x, g, reduce="mean", act="none"), want)


def test_dense_matrix_is_not_an_adjacency():
    conv = S.MySAGEConv(8, 8).to(DEV)
    with pytest.raises((TypeError, ValueError)):
        conv(torch.randn(10, 8, device=DEV), object())
