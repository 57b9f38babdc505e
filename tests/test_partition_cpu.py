"""CPU (gloo, world_size 2): the integer bookkeeping of the 1-D destination partition and the
halo all-gather layout.  The partitioned aggregation is emulated with the oracle's CPU propagate
on each rank's local edge list and must equal the single-graph oracle bit for bit."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def test_partition_ranges_cover_the_nodes():
    from stemgnn_b200.partition import block_size, partition_1d
    from oracle.graph_ref import partition_1d_ref
    for n, p in [(10, 4), (0, 3), (7, 8), (169343, 8), (2449029, 8), (16, 2)]:
        parts = partition_1d(n, p)
        assert parts == partition_1d_ref(n, p) if n > 0 else all(a == b == 0 for a, b in parts)
        assert sum(hi - lo for lo, hi in parts) == n
        assert all(hi - lo <= block_size(n, p) for lo, hi in parts)
        assert all(lo == r * block_size(n, p) or lo == n for r, (lo, hi) in enumerate(parts))


def test_local_edges_keep_original_order():
    from stemgnn_b200.partition import local_edges, partition_1d
    g = torch.Generator().manual_seed(0)
    n, e = 53, 700
    ei = torch.randint(0, n, (2, e), generator=g)
    seen = torch.zeros(e, dtype=torch.int64)
    for owner_row in (0, 1):
        seen.zero_()
        for lo, hi in partition_1d(n, 4):
            loc, eid = local_edges(ei, lo, hi, owner_row)
            assert torch.all(eid[1:] > eid[:-1])                       # original order kept
            assert torch.equal(loc[owner_row] + lo, ei[owner_row, eid])
            assert torch.equal(loc[1 - owner_row], ei[1 - owner_row, eid])
            assert loc.numel() == 0 or (0 <= int(loc[owner_row].min()) and int(loc[owner_row].max()) < hi - lo)
            seen[eid] += 1
        assert torch.all(seen == 1)                                    # every edge on exactly one rank


def _worker(rank, world, port, n, e, f, ret):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import propagate_ref as P
        from stemgnn_b200.partition import block_size, local_edges, partition_1d
        g = torch.Generator().manual_seed(3)
        ei = torch.randint(0, n, (2, e), generator=g)
        x = torch.randn(n, f, generator=g)
        go = torch.randn(n, f, generator=g)
        lo, hi = partition_1d(n, world)[rank]
        per = block_size(n, world)
        # halo exchange of the padded row blocks: row g of the result is global node g
        blk = torch.zeros(per, f)
        blk[:hi - lo] = x[lo:hi]
        full = torch.empty(world * per, f)
        dist.all_gather_into_tensor(full, blk)
        assert torch.equal(full[:n], x)
        # forward on the owned destinations == the rows of the single-graph result
        ref_x = x.clone().requires_grad_(True)
        ref = P.sage_propagate_ref(ref_x, ei, None, n, "mean")
        ref.backward(go)
        fwd, _ = local_edges(ei, lo, hi, 1)
        xf = full.clone()
        x_j = torch.relu(xf.index_select(0, fwd[0]))
        out = P.scatter_mean_ref(x_j, fwd[1], hi - lo)
        assert torch.equal(out, ref.detach()[lo:hi])
        # backward on the owned sources: gathered g_out / count of the destination, relu mask of x
        deg = torch.bincount(ei[1], minlength=n).clamp(min=1).float()
        gs = go / deg.view(-1, 1)
        bwd, _ = local_edges(ei, lo, hi, 0)
        contrib = gs.index_select(0, bwd[1]) * (x[lo:hi].index_select(0, bwd[0]) > 0)
        gx = torch.zeros(hi - lo, f).index_add_(0, bwd[0], contrib)
        assert torch.equal(gx, ref_x.grad[lo:hi])
        ret[rank] = 1
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_partitioned_aggregation_two_ranks_gloo():
    world = 2
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, world, port, 101, 900, 12, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(100)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    assert dict(ret) == {0: 1, 1: 1}


def _bn_worker(rank, world, port, n, f, ret):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from stemgnn_b200.norm_act import BatchSync, merge_batch_stats
        from stemgnn_b200.partition import partition_1d
        g = torch.Generator().manual_seed(5)
        x = (torch.randn(n, f, generator=g) * 3 + 7).double()
        gy = torch.randn(n, f, generator=g).double()
        lo, hi = partition_1d(n, world)[rank]
        xl, gl = x[lo:hi], gy[lo:hi]
        sync = BatchSync(n, world)                                   # torch.distributed defaults
        # forward: (mean, biased var, rows) of the uneven shares -> the full batch's
        mean, var = merge_batch_stats(xl.mean(0), xl.var(0, unbiased=False), hi - lo, sync)
        assert torch.allclose(mean, x.mean(0), rtol=1e-12, atol=1e-12)
        assert torch.allclose(var, x.var(0, unbiased=False), rtol=1e-11, atol=1e-12)
        # backward: the two column sums are the only thing the ranks exchange; with the batch's sums and
        # row count the local rows of the input gradient equal those of the full-batch formula
        invstd = (var + 1e-5).rsqrt()
        xhat_l = (xl - mean) * invstd
        sums = sync.reduce(torch.cat([gl.sum(0), (gl * xhat_l).sum(0)]))
        sum_g, sum_gx = sums[:f], sums[f:]
        gx_l = invstd * (gl - sum_g / n - xhat_l * sum_gx / n)       # gamma = 1
        xr = x.clone().requires_grad_(True)
        torch.nn.functional.batch_norm(xr, None, None, training=True, eps=1e-5).backward(gy)
        assert torch.allclose(gx_l, xr.grad[lo:hi], rtol=1e-9, atol=1e-11)
        # VectorQuantize follows the reference's rule for its statistics all-reduce (model/vq.py:771-772): on by
        # default exactly when a process group with more than one rank exists; several codebooks with a
        # distributed k-means init are refused like the reference's assert (vq.py:285)
        import stemgnn_b200 as S
        assert S.VectorQuantize(dim=8, codebook_size=4)._codebook.use_ddp is True
        assert S.VectorQuantize(dim=8, codebook_size=4, sync_codebook=False)._codebook.use_ddp is False
        with pytest.raises(NotImplementedError):
            S.VectorQuantize(dim=8, codebook_size=4, heads=2, separate_codebook_per_head=True, kmeans_init=True)
        # k-means init (model/vq.py:287-288): sync_kmeans on -> the initial means come from EVERY rank's rows
        # (multinomial split on rank 0, local draws, exchange) and are the same tensor on all ranks; off -> the
        # local sampler and no collective
        from stemgnn_b200 import vq as V
        cb = S.VectorQuantize(dim=8, codebook_size=4, kmeans_init=True)._codebook
        assert cb.sync_kmeans is True and cb.sample_fn is V._sample_vectors_distributed
        cb = S.VectorQuantize(dim=8, codebook_size=4, kmeans_init=True, sync_kmeans=False)._codebook
        assert cb.sync_kmeans is False and cb.sample_fn is V._batched_sample_vectors
        torch.manual_seed(100 + rank)
        rows = torch.arange((rank + 1) * 300, dtype=torch.float32).view(-1, 1).repeat(1, 3) + 10000 * rank
        means = V._sample_vectors_distributed(rows.unsqueeze(0), 64)     # rank 0 owns 300 rows, rank 1 600
        assert means.shape == (1, 64, 3)
        both = [torch.empty_like(means) for _ in range(world)]
        dist.all_gather(both, means)
        assert torch.equal(both[0], both[1])                              # identical on every rank
        owner = (means[0, :, 0] >= 10000).long()
        assert 0 < int(owner.sum()) < 64 and torch.all(owner[1:] >= owner[:-1])      # both ranks contribute, rank order
        assert means[0, :, 0].unique().numel() == 64                      # without replacement
        ret[rank] = 1
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_batchnorm_statistics_over_two_ranks_gloo():
    """norm_act.BatchSync over real collectives (all_gather_into_tensor / all_reduce on gloo): the merge of
    per-rank statistics and the split of the backward at its reduction (what K7 does on each side of
    stemgnn_bn_act_backward / _backward_apply) against torch's full-batch batch_norm in float64."""
    world = 2
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = 31500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_bn_worker, args=(r, world, port, 1001, 20, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(100)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    assert dict(ret) == {0: 1, 1: 1}


# ------------------------------------------------------------------ the overlapped exchange (halo_mode="blocks")
def test_block_edges_cut_by_owner_block_and_keep_order():
    from stemgnn_b200.partition import block_edges, block_size, exchange_order, local_edges, partition_1d
    g = torch.Generator().manual_seed(1)
    n, e, world = 53, 900, 4
    ei = torch.randint(0, n, (2, e), generator=g)
    per = block_size(n, world)
    for rank, (lo, hi) in enumerate(partition_1d(n, world)):
        for own_row in (1, 0):                             # forward: owned destinations; backward: owned sources
            loc, _ = local_edges(ei, lo, hi, own_row)
            seen = torch.zeros(loc.size(1), dtype=torch.int64)
            for b, (sub, pos) in enumerate(block_edges(loc, per, world, 1 - own_row)):
                assert torch.all(pos[1:] > pos[:-1])                                # relative order kept
                assert torch.equal(sub[own_row], loc[own_row, pos])
                assert torch.equal(sub[1 - own_row] + b * per, loc[1 - own_row, pos])
                assert sub.numel() == 0 or (0 <= int(sub[1 - own_row].min()) and int(sub[1 - own_row].max()) < per)
                seen[pos] += 1
            assert torch.all(seen == 1)
        for tr in ("p2p", "bcast"):
            order = exchange_order(rank, world, tr)
            assert order[0] == rank and sorted(order) == list(range(world))


def test_block_by_block_sum_equals_the_edge_order_sum_within_rounding():
    """What stemgnn_aggregate(accumulate=1) computes over the owner-block CSRs, restated on the CPU: the row sums
    run block by block (own block first) instead of in edge order -- 1e-5 of the output scale, not bit-identical."""
    from oracle import propagate_ref as P
    from stemgnn_b200.partition import block_edges, block_size, exchange_order, local_edges, partition_1d
    g = torch.Generator().manual_seed(2)
    n, e, f, world = 211, 6000, 16, 4
    ei = torch.randint(0, n, (2, e), generator=g)
    ei[1, :700] = 5
    x = torch.randn(n, f, generator=g)
    ref = P.sage_propagate_ref(x, ei, None, n, "mean")
    per = block_size(n, world)
    xp = torch.cat([x, x.new_zeros(world * per - n, f)])
    deg = torch.bincount(ei[1], minlength=n).clamp(min=1).float()
    for rank, (lo, hi) in enumerate(partition_1d(n, world)):
        loc, _ = local_edges(ei, lo, hi, 1)
        subs = block_edges(loc, per, world, 0)
        acc = torch.zeros(hi - lo, f)
        for b in exchange_order(rank, world, "p2p"):
            sub = subs[b][0]
            blk = xp[b * per:(b + 1) * per]
            acc = acc.index_add(0, sub[1], torch.relu(blk.index_select(0, sub[0])))
        out = acc / deg[lo:hi].view(-1, 1)
        err = (out - ref[lo:hi]).abs().max() / ref.abs().max()
        assert float(err) < 1e-5


def _blocks_worker(rank, world, port, n, f, transport, ret):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from stemgnn_b200.partition import PartitionedGraph, block_size, exchange_order, partition_1d
        g = torch.Generator().manual_seed(7)
        x = torch.randn(n, f, generator=g)
        per = block_size(n, world)
        lo, hi = partition_1d(n, world)[rank]
        empty = torch.zeros(2, 0, dtype=torch.int64)
        for wire in (None, torch.bfloat16):
            pg = PartitionedGraph(None, n, rank=rank, world=world, pre_split=(empty, empty), halo_mode="blocks",
                                  halo_transport=transport, halo_dtype=wire)
            got = list(pg.exchange_blocks(x[lo:hi]))
            assert [b for b, _ in got] == exchange_order(rank, world, transport)
            for b, blk in got:
                blo, bhi = partition_1d(n, world)[b]
                want = torch.zeros(per, f)
                want[:bhi - blo] = x[blo:bhi]
                assert blk.shape == (per, f) and blk.dtype == torch.float32
                if wire is None or b == rank:
                    assert torch.equal(blk, want)                  # own block never crosses the wire
                else:
                    assert torch.equal(blk, want.to(torch.bfloat16).float())
        ret[rank] = 1
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
@pytest.mark.parametrize("world,transport", [(2, "p2p"), (2, "bcast"), (3, "p2p"), (3, "bcast")])
def test_block_exchange_delivers_every_owner_block_gloo(world, transport):
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = 33500 + (os.getpid() % 2000) + 7 * world + (1 if transport == "p2p" else 0)
    procs = [ctx.Process(target=_blocks_worker, args=(r, world, port, 103, 8, transport, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(150)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    assert dict(ret) == {r: 1 for r in range(world)}


def test_chunk_edges_cut_by_row_chunk_and_renumber_into_the_chunk_all_gather():
    from stemgnn_b200.partition import block_size, chunk_edges, local_edges, partition_1d
    g = torch.Generator().manual_seed(4)
    n, e, world, chunks = 53, 900, 4, 3
    ei = torch.randint(0, n, (2, e), generator=g)
    per = block_size(n, world)
    cr = -(-per // chunks)
    for lo, hi in partition_1d(n, world):
        for own_row in (1, 0):
            loc, _ = local_edges(ei, lo, hi, own_row)
            seen = torch.zeros(loc.size(1), dtype=torch.int64)
            for c, (sub, pos) in enumerate(chunk_edges(loc, per, world, chunks, 1 - own_row)):
                assert torch.all(pos[1:] > pos[:-1])
                assert torch.equal(sub[own_row], loc[own_row, pos])
                gl = loc[1 - own_row, pos]
                b, l = gl // per, gl % per
                assert torch.all(l // cr == c)
                assert torch.equal(sub[1 - own_row], b * cr + (l - c * cr))        # row in the gather of the c-th chunks
                assert sub.numel() == 0 or int(sub[1 - own_row].max()) < world * cr
                seen[pos] += 1
            assert torch.all(seen == 1)


def _chunks_worker(rank, world, port, n, f, chunks, ret):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from stemgnn_b200.partition import PartitionedGraph, block_size, partition_1d
        g = torch.Generator().manual_seed(7)
        x = torch.randn(n, f, generator=g)
        per = block_size(n, world)
        cr = -(-per // chunks)
        lo, hi = partition_1d(n, world)[rank]
        empty = torch.zeros(2, 0, dtype=torch.int64)
        pg = PartitionedGraph(None, n, rank=rank, world=world, pre_split=(empty, empty), halo_mode="chunks",
                              halo_chunks=chunks)
        got = list(pg.exchange_groups(x[lo:hi]))
        assert [c for c, _ in got] == list(range(chunks)) and pg.n_groups == chunks
        for c, buf in got:
            assert buf.shape == (world * cr, f)
            for b, (blo, bhi) in enumerate(partition_1d(n, world)):
                want = torch.zeros(chunks * cr, f)
                want[:bhi - blo] = x[blo:bhi]
                assert torch.equal(buf[b * cr:(b + 1) * cr], want[c * cr:(c + 1) * cr])
        ret[rank] = 1
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
@pytest.mark.parametrize("world,chunks", [(2, 3), (3, 2)])
def test_chunked_all_gather_delivers_every_row_chunk_gloo(world, chunks):
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = 35500 + (os.getpid() % 2000) + 11 * world + chunks
    procs = [ctx.Process(target=_chunks_worker, args=(r, world, port, 103, 8, chunks, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(150)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    assert dict(ret) == {r: 1 for r in range(world)}


def test_chunk_by_chunk_sum_equals_the_edge_order_sum_within_rounding():
    """halo_mode='chunks' restated on the CPU: chunk c of every rank's block is gathered into one [world * cr, F]
    matrix, the edges whose source lies in chunk c are summed into the owned rows chunk after chunk -- 1e-5 of the
    output scale against the oracle's edge-order sum, forward (mean of relu messages) and backward (masked sums)."""
    from oracle import propagate_ref as P
    from stemgnn_b200.partition import block_size, chunk_edges, local_edges, partition_1d
    g = torch.Generator().manual_seed(6)
    n, e, f, world, chunks = 211, 6000, 16, 4, 3
    ei = torch.randint(0, n, (2, e), generator=g)
    ei[1, :700] = 5
    x = torch.randn(n, f, generator=g)
    go = torch.randn(n, f, generator=g)
    xr = x.clone().requires_grad_(True)
    ref = P.sage_propagate_ref(xr, ei, None, n, "mean")
    ref.backward(go)
    per = block_size(n, world)
    cr = -(-per // chunks)
    deg = torch.bincount(ei[1], minlength=n).clamp(min=1).float()

    def chunk_matrix(full, c):          # what all_gather_into_tensor of the ranks' c-th row chunks returns
        rows = []
        for lo, hi in partition_1d(n, world):
            blk = torch.zeros(chunks * cr, f)
            blk[:hi - lo] = full[lo:hi]
            rows.append(blk[c * cr:(c + 1) * cr])
        return torch.cat(rows)

    gs = go / deg.view(-1, 1)
    for lo, hi in partition_1d(n, world):
        fwd, _ = local_edges(ei, lo, hi, 1)
        acc = torch.zeros(hi - lo, f)
        for c, (sub, _) in enumerate(chunk_edges(fwd, per, world, chunks, 0)):
            acc = acc.index_add(0, sub[1], torch.relu(chunk_matrix(x, c).index_select(0, sub[0])))
        out = acc / deg[lo:hi].view(-1, 1)
        assert float((out - ref.detach()[lo:hi]).abs().max() / ref.detach().abs().max()) < 1e-5
        bwd, _ = local_edges(ei, lo, hi, 0)
        gacc = torch.zeros(hi - lo, f)
        for c, (sub, _) in enumerate(chunk_edges(bwd, per, world, chunks, 1)):
            gacc = gacc.index_add(0, sub[0], chunk_matrix(gs, c).index_select(0, sub[1]))
        gx = gacc * (x[lo:hi] > 0)
        assert float((gx - xr.grad[lo:hi]).abs().max() / xr.grad.abs().max()) < 1e-5
