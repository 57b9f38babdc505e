"""GPU: the 1-D destination-partitioned aggregation through the C-ABI.  The ranks of a P-way
partition are run one after another on one GPU with the halo all-gather replaced by the known
full matrices (no spinning multi-process kernels on one GPU); every rank's rows must equal the
single-graph result bit for bit, forward and backward."""
import numpy as np
import pytest
import torch

import stemgnn_b200 as S

pytestmark = pytest.mark.gpu
DEV = "cuda"


class _Queue:
    """Stand-in for the collective: returns the pre-computed gathered matrices in call order."""

    def __init__(self, items):
        self.items = list(items)

    def __call__(self, blk):
        full = self.items.pop(0)
        assert blk.size(1) == full.size(1)
        return full


@pytest.mark.parametrize("world", [1, 2, 3, 8])
@pytest.mark.parametrize("reduce,act", [("mean", "relu"), ("sum", "none")])
def test_partitioned_rows_equal_single_graph(world, reduce, act):
    n, e, f = 1003, 9000, 36
    g = torch.Generator().manual_seed(world)
    ei = torch.randint(0, n, (2, e), generator=g)
    ei[1, :500] = 17                                     # a hub destination
    x = torch.randn(n, f, generator=g).to(DEV)
    go = torch.randn(n, f, generator=g).to(DEV)
    ei = ei.to(DEV)
    xs = x.clone().requires_grad_(True)
    ref = S.aggregate(xs, S.GraphStructure(ei, n), reduce=reduce, act=act)
    ref.backward(go)
    per = S.block_size(n, world)
    pad = lambda t: torch.cat([t, t.new_zeros(world * per - n, t.size(1))])
    deg = torch.bincount(ei[1], minlength=n).clamp(min=1).float().view(-1, 1)
    for rank, (lo, hi) in enumerate(S.partition_1d(n, world)):
        q = [pad(x), pad(go / deg if reduce == "mean" else go)]   # each rank scales its own g_out rows
        pg = S.PartitionedGraph(ei, n, rank=rank, world=world, gather=_Queue(q))
        assert pg.num_nodes == hi - lo
        xl = x[lo:hi].clone().requires_grad_(True)
        out = S.aggregate(xl, pg, reduce=reduce, act=act)
        out.backward(go[lo:hi])
        np.testing.assert_array_equal(out.detach().cpu().numpy(), ref.detach()[lo:hi].cpu().numpy())
        np.testing.assert_array_equal(xl.grad.cpu().numpy(), xs.grad[lo:hi].cpu().numpy())


def test_encoder_on_a_partition_matches_full_graph_rows():
    """A 2-layer SAGE encoder without BatchNorm on rank 1 of 2: its rows of the output equal the
    full-graph encoder's (the dense layers act row-wise)."""
    import torch.nn as nn
    n, e, f, hid = 400, 3000, 24, 32
    g = torch.Generator().manual_seed(5)
    ei = torch.randint(0, n, (2, e), generator=g).to(DEV)
    x = torch.randn(n, f, generator=g).to(DEV)
    torch.manual_seed(0)
    enc = S.Encoder(f, hid, nn.ReLU, 2, backbone="sage", normalize="none").to(DEV)
    with torch.no_grad():
        z_full = enc(x, ei)
        h1_full = enc.layers[0](x, ei).relu()            # layer-1 input of every node
        world, rank = 2, 1
        per = S.block_size(n, world)
        lo, hi = S.partition_1d(n, world)[rank]
        pad = lambda t: torch.cat([t, t.new_zeros(world * per - n, t.size(1))])
        pg = S.PartitionedGraph(ei, n, rank=rank, world=world, gather=_Queue([pad(x), pad(h1_full)]))
        z_loc = enc(x[lo:hi], pg)
    torch.testing.assert_close(z_loc, z_full[lo:hi], rtol=1e-6, atol=1e-6)


def test_encoder_with_batchnorm_on_a_partition_uses_full_batch_statistics():
    """normalize='batch': a rank's BatchNorm must use the statistics of ALL nodes, so its
    rows equal the single-GPU encoder's.  Collectives are replaced by their known results."""
    import torch.nn as nn
    n, e, f, hid, world, rank = 600, 5000, 24, 32, 3, 1
    g = torch.Generator().manual_seed(6)
    ei = torch.randint(0, n, (2, e), generator=g).to(DEV)
    x = torch.randn(n, f, generator=g).to(DEV)
    torch.manual_seed(0)
    enc = S.Encoder(f, hid, nn.ReLU, 2, backbone="sage", normalize="batch", dropout=0.0).to(DEV).train()
    state = {k: v.clone() for k, v in enc.state_dict().items()}
    pre_bn = []
    hooks = [l.register_forward_hook(lambda m, i, o: pre_bn.append(o.detach())) for l in enc.layers]
    with torch.no_grad():
        z_full = enc(x, ei)
        for h in hooks:
            h.remove()
        h1_full = S.bn_act_drop(pre_bn[0], nn.BatchNorm1d(hid).to(DEV).train(), True, 0.0, True)   # default affine
        per = S.block_size(n, world)
        parts = S.partition_1d(n, world)
        lo, hi = parts[rank]
        pad = lambda t: torch.cat([t, t.new_zeros(world * per - n, t.size(1))])
        packs = [torch.stack([torch.cat([*S.batch_stats(z[a:b]), torch.tensor([float(b - a)], device=DEV)])
                              for a, b in parts]) for z in pre_bn]
        enc.load_state_dict(state)                           # running statistics back to their start
        pg = S.PartitionedGraph(ei, n, rank=rank, world=world, gather=_Queue([pad(x), pad(h1_full)]),
                                stat_gather=lambda v, it=iter(packs): next(it), stat_reduce=lambda v: v)
        z_loc = enc(x[lo:hi], pg)
    torch.testing.assert_close(z_loc, z_full[lo:hi], rtol=1e-5, atol=1e-5)


def test_feature_sliced_halo_pipeline_is_bit_identical():
    """The exchange pipelined over 3 feature slices (columns 0-87, 88-175, 176-255 gathered and
    aggregated one after the other into column ranges of the output) gives the same bits."""
    n, e, f, world, rank = 1003, 9000, 256, 4, 2
    g = torch.Generator().manual_seed(9)
    ei = torch.randint(0, n, (2, e), generator=g).to(DEV)
    x = torch.randn(n, f, generator=g).to(DEV)
    go = torch.randn(n, f, generator=g).to(DEV)
    xs = x.clone().requires_grad_(True)
    ref = S.aggregate(xs, S.GraphStructure(ei, n), reduce="mean", act="relu")
    ref.backward(go)
    per = S.block_size(n, world)
    lo, hi = S.partition_1d(n, world)[rank]
    pad = lambda t: torch.cat([t, t.new_zeros(world * per - n, t.size(1))])
    deg = torch.bincount(ei[1], minlength=n).clamp(min=1).float().view(-1, 1)
    pg = S.PartitionedGraph(ei, n, rank=rank, world=world, gather=lambda blk: None, halo_slices=3)
    sl = pg.feature_slices(f)
    assert sl == [(0, 88), (88, 176), (176, 256)]
    q = [pad(x)[:, a:b].contiguous() for a, b in sl] + [pad(go / deg)[:, a:b].contiguous() for a, b in sl]
    pg._gather = _Queue(q)
    xl = x[lo:hi].clone().requires_grad_(True)
    out = S.aggregate(xl, pg, reduce="mean", act="relu")
    out.backward(go[lo:hi])
    assert not q or len(pg._gather.items) == 0
    np.testing.assert_array_equal(out.detach().cpu().numpy(), ref.detach()[lo:hi].cpu().numpy())
    np.testing.assert_array_equal(xl.grad.cpu().numpy(), xs.grad[lo:hi].cpu().numpy())


class _Blocks:
    """Stand-in for the overlapped exchange: hands out the known padded row blocks of the other ranks, one matrix
    per exchange in call order."""

    def __init__(self, fulls, per, world):
        self.fulls, self.per, self.world = list(fulls), per, world

    def __call__(self, blk):
        full = self.fulls.pop(0)
        assert blk.shape == (self.per, full.size(1))
        return [full[b * self.per:(b + 1) * self.per] for b in range(self.world)]


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("transport", ["p2p", "bcast", "chunks3", "chunks4"])
@pytest.mark.parametrize("reduce,act,f", [("mean", "relu", 36), ("sum", "none", 256), ("mean", "relu", 256)])
def test_overlapped_block_exchange_within_1e5_of_single_graph(world, transport, reduce, act, f):
    """halo_mode='blocks': the rank's rows are summed owner block by owner block (accumulate = 1) in the order the
    blocks arrive.  Same terms as the single-graph sum in another order: 1e-5 of the output scale (north_star's
    bar); the bit-identical mode stays the default."""
    n, e = 1003, 9000
    g = torch.Generator().manual_seed(world)
    ei = torch.randint(0, n, (2, e), generator=g)
    ei[1, :500] = 17                                     # a hub destination: the hub launch accumulates too
    ei[0, 500:900] = 40                                  # and a hub source for the backward
    x = torch.randn(n, f, generator=g).to(DEV)
    go = torch.randn(n, f, generator=g).to(DEV)
    ei = ei.to(DEV)
    xs = x.clone().requires_grad_(True)
    ref = S.aggregate(xs, S.GraphStructure(ei, n), reduce=reduce, act=act)
    ref.backward(go)
    per = S.block_size(n, world)
    pad = lambda t: torch.cat([t, t.new_zeros(world * per - n, t.size(1))])
    deg = torch.bincount(ei[1], minlength=n).clamp(min=1).float().view(-1, 1)
    so, sg = ref.detach().abs().max().item(), xs.grad.abs().max().item()
    for rank, (lo, hi) in enumerate(S.partition_1d(n, world)):
        q = _Blocks([pad(x), pad(go / deg if reduce == "mean" else go)], per, world)
        kw = dict(halo_mode="chunks", halo_chunks=int(transport[-1])) if transport.startswith("chunks") else \
            dict(halo_mode="blocks", halo_transport=transport)
        pg = S.PartitionedGraph(ei, n, rank=rank, world=world, blocks=q, **kw)
        xl = x[lo:hi].clone().requires_grad_(True)
        out = S.aggregate(xl, pg, reduce=reduce, act=act)
        out.backward(go[lo:hi])
        assert not q.fulls
        assert (out.detach() - ref.detach()[lo:hi]).abs().max().item() <= 1e-5 * so
        assert (xl.grad - xs.grad[lo:hi]).abs().max().item() <= 1e-5 * sg
        # deterministic: the same call again gives the same bits
        q2 = _Blocks([pad(x)], per, world)
        pg2 = S.PartitionedGraph(ei, n, rank=rank, world=world, blocks=q2, **kw)
        with torch.no_grad():
            out2 = S.aggregate(x[lo:hi].clone(), pg2, reduce=reduce, act=act)
        assert torch.equal(out2, out.detach())


def test_overlapped_block_exchange_one_rank_is_the_plain_aggregation():
    n, e, f = 500, 4000, 64
    g = torch.Generator().manual_seed(3)
    ei = torch.randint(0, n, (2, e), generator=g).to(DEV)
    x = torch.randn(n, f, generator=g).to(DEV)
    ref = S.aggregate(x, S.GraphStructure(ei, n), reduce="mean", act="relu")
    pg = S.PartitionedGraph(ei, n, rank=0, world=1, halo_mode="blocks")
    out = S.aggregate(x, pg, reduce="mean", act="relu")
    assert torch.equal(out, ref)                         # one block: the edge order is the original one
