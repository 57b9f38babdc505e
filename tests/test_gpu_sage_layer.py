"""GPU: the SAGE layer as one autograd node (sage.py: count division in the epilogue of g @ W_l, the lin_r gradient
added by the aggregation's epilogue) against the composed form aggregate -> linear_sum it replaces: bit-identical on
one GPU, forward and backward; on a partition (ranks run one after another, collectives replaced by the known
matrices) bit-identical in the all-gather mode and within 1e-5 in the overlapped owner-block mode."""
import numpy as np
import pytest
import torch
import torch.nn as nn

import stemgnn_b200 as S
from stemgnn_b200 import dense, sage

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _layer(f, o, root, seed):
    torch.manual_seed(seed)
    lin_l = nn.Linear(f, o).to(DEV)
    lin_r = nn.Linear(f, o, bias=False).to(DEV) if root else None
    return lin_l, lin_r


def _composed(x, g, reduce, lin_l, lin_r):
    agg = S.aggregate(x, g, reduce=reduce, act="relu")
    if lin_r is not None:
        return dense.linear_sum([agg, x], [lin_l.weight, lin_r.weight], lin_l.bias)
    return dense.linear_sum([agg], [lin_l.weight], lin_l.bias)


def _grads(mods):
    return [p.grad.clone() for m in mods if m is not None for p in m.parameters()]


@pytest.mark.parametrize("n,e,f,o,hub", [(1003, 9000, 36, 48, 600), (5000, 60000, 128, 256, 0), (3000, 20000, 256, 256, 5000)])
@pytest.mark.parametrize("reduce", ["mean", "sum"])
@pytest.mark.parametrize("root", [True, False])
def test_fused_layer_is_bit_identical_to_the_composed_form(n, e, f, o, hub, reduce, root):
    g = torch.Generator().manual_seed(n + f)
    ei = torch.randint(0, n, (2, e), generator=g)
    if hub:
        ei[1, :hub] = 11
        ei[0, hub:2 * hub] = 13
    ei = ei.to(DEV)
    x = torch.randn(n, f, generator=g).to(DEV)
    go = torch.randn(n, o, generator=g).to(DEV)
    lin_l, lin_r = _layer(f, o, root, 3)
    res = []
    for fn in (_composed, sage.sage_layer):
        for m in (lin_l, lin_r):
            if m is not None:
                m.zero_grad()
        xs = x.clone().requires_grad_(True)
        out = fn(xs, S.GraphStructure(ei, n), reduce, lin_l, lin_r)
        out.backward(go)
        res.append([out.detach(), xs.grad] + _grads((lin_l, lin_r)))
    for a, b in zip(*res):
        assert torch.equal(a, b)
    # an input that needs no gradient (the first layer of an encoder): parameters only
    out = sage.sage_layer(x, S.GraphStructure(ei, n), reduce, lin_l, lin_r)
    assert torch.equal(out, res[0][0])
    out.backward(go)


class _Queue:
    def __init__(self, items):
        self.items = list(items)

    def __call__(self, blk):
        return self.items.pop(0)


class _Blocks:
    def __init__(self, fulls, per, world):
        self.fulls, self.per, self.world = list(fulls), per, world

    def __call__(self, blk):
        full = self.fulls.pop(0)
        return [full[b * self.per:(b + 1) * self.per] for b in range(self.world)]


@pytest.mark.parametrize("mode", ["gather", "blocks", "chunks"])
@pytest.mark.parametrize("world", [2, 4])
def test_fused_layer_on_a_partition(mode, world):
    n, e, f, o = 1003, 9000, 64, 96
    g = torch.Generator().manual_seed(world)
    ei = torch.randint(0, n, (2, e), generator=g)
    ei[1, :500] = 17
    ei[0, 500:900] = 40
    ei = ei.to(DEV)
    x = torch.randn(n, f, generator=g).to(DEV)
    go = torch.randn(n, o, generator=g).to(DEV)
    lin_l, lin_r = _layer(f, o, True, 5)
    xs = x.clone().requires_grad_(True)
    ref = sage.sage_layer(xs, S.GraphStructure(ei, n), "mean", lin_l, lin_r)
    ref.backward(go)
    ref_w = _grads((lin_l, lin_r))
    per = S.block_size(n, world)
    pad = lambda t: torch.cat([t, t.new_zeros(world * per - n, t.size(1))])
    deg = torch.bincount(ei[1], minlength=n).clamp(min=1).float().view(-1, 1)
    g_agg_full = pad((go @ lin_l.weight.detach()) / deg)        # what the ranks exchange in the backward (1e-6 of K5's)
    sum_w = [torch.zeros_like(w) for w in ref_w]
    for rank, (lo, hi) in enumerate(S.partition_1d(n, world)):
        for m in (lin_l, lin_r):
            m.zero_grad()
        if mode == "gather":
            pg = S.PartitionedGraph(ei, n, rank=rank, world=world, gather=_Queue([pad(x), g_agg_full]))
        else:
            pg = S.PartitionedGraph(ei, n, rank=rank, world=world, halo_mode=mode, halo_chunks=3,
                                    blocks=_Blocks([pad(x), g_agg_full], per, world))
        xl = x[lo:hi].clone().requires_grad_(True)
        out = sage.sage_layer(xl, pg, "mean", lin_l, lin_r)
        out.backward(go[lo:hi])
        so, sg = ref.detach().abs().max().item(), xs.grad.abs().max().item()
        if mode == "gather":
            np.testing.assert_array_equal(out.detach().cpu().numpy(), ref.detach()[lo:hi].cpu().numpy())
        else:
            assert (out.detach() - ref.detach()[lo:hi]).abs().max().item() <= 1e-5 * so
        assert (xl.grad - xs.grad[lo:hi]).abs().max().item() <= 2e-5 * sg      # the exchanged g_agg is torch's product
        for s_, w in zip(sum_w, _grads((lin_l, lin_r))):
            s_ += w
    for s_, w in zip(sum_w, ref_w):                              # the ranks' parameter gradients add up to the graph's
        assert ((s_ - w).abs().max() / w.abs().max()).item() <= 2e-5
