"""torchrun --nproc-per-node P scripts/check_partition_multi.py : every rank builds its share of a
seeded graph, runs the partitioned aggregation with the NCCL halo all-gather and compares its rows
with the single-graph result computed locally (bit-exact), then times both."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import stemgnn_b200 as S

def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    opts = None
    if os.environ.get("STEMGNN_NCCL_HIGH_PRIO", "0") == "1":       # collectives on a high-priority stream: their copy
        opts = dist.ProcessGroupNCCL.Options(is_high_priority_stream=True)   # kernels get SMs while an aggregation runs
    dist.init_process_group("nccl", device_id=torch.device("cuda", local), pg_options=opts)
    n, e, f = int(os.environ.get("N", 300000)), int(os.environ.get("E", 6000000)), int(os.environ.get("F", 100))
    g = torch.Generator().manual_seed(7)
    ei = torch.randint(0, n, (2, e), generator=g).cuda()
    x = torch.randn(n, f, generator=g).cuda()
    go = torch.randn(n, f, generator=g).cuda()
    xs = x.clone().requires_grad_(True)
    ref = S.aggregate(xs, S.GraphStructure(ei, n), reduce="mean", act="relu")
    ref.backward(go)
    pg = S.PartitionedGraph(ei, n)
    lo, hi = pg.lo, pg.hi
    xl = x[lo:hi].clone().requires_grad_(True)
    out = S.aggregate(xl, pg, reduce="mean", act="relu")
    out.backward(go[lo:hi])
    ok = torch.equal(out.detach(), ref.detach()[lo:hi]) and torch.equal(xl.grad, xs.grad[lo:hi])
    t = torch.tensor([int(ok)], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MIN)
    # timing: fwd+bwd partitioned (with halo exchange)
    def step():
        xl.grad = None
        o = S.aggregate(xl, pg, reduce="mean", act="relu"); o.backward(go[lo:hi])
    for _ in range(3): step()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): step()
    e1.record(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / 10], device="cuda"); dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    slices = len(pg.feature_slices(f))
    # the same exchange unsliced (one blocking all-gather per direction) and on a bf16 wire
    def timed(pgx):
        def st():
            xl.grad = None
            o = S.aggregate(xl, pgx, reduce="mean", act="relu"); o.backward(go[lo:hi])
        for _ in range(3): st()
        torch.cuda.synchronize(); dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(10): st()
        b.record(); torch.cuda.synchronize()
        m = torch.tensor([a.elapsed_time(b) / 10], device="cuda"); dist.all_reduce(m, op=dist.ReduceOp.MAX)
        return m.item()
    pg1 = S.PartitionedGraph(ei, n, halo_slices=1)
    ms1 = timed(pg1)
    pgb = S.PartitionedGraph(ei, n, halo_dtype=torch.bfloat16)
    xl.grad = None
    ob = S.aggregate(xl, pgb, reduce="mean", act="relu"); ob.backward(go[lo:hi])
    errb = ((ob.detach() - ref.detach()[lo:hi]).abs().max() / ref.detach().abs().max()).item()
    errg = ((xl.grad - xs.grad[lo:hi]).abs().max() / xs.grad.abs().max()).item()
    msb = timed(pgb)
    if rank == 0:
        print(f"partition check world={world} N={n} E={e} F={f}: bit_exact={bool(t.item())} "
              f"fwd+bwd {ms.item():.3f} ms ({slices} feature slices)  unsliced {ms1:.3f} ms  "
              f"bf16 wire {msb:.3f} ms (max err fwd {errb:.1e} bwd {errg:.1e})  "
              f"{2 * e / ms.item() / 1e6:.1f} Medges/s (both directions)", flush=True)
    if os.environ.get("STEMGNN_CHECK_GATHER_ONLY") == "1":
        dist.destroy_process_group()
        sys.exit(0 if t.item() else 1)
    # the overlapped exchange (halo_mode="blocks"): owner blocks aggregated as they land, 1e-5 of the output scale
    ok_blocks = True
    for transport in ("p2p", "bcast"):
        for wire in (None, torch.bfloat16):
            pgo = S.PartitionedGraph(ei, n, halo_mode="blocks", halo_transport=transport, halo_dtype=wire)
            xl.grad = None
            oo = S.aggregate(xl, pgo, reduce="mean", act="relu"); oo.backward(go[lo:hi])
            ef = ((oo.detach() - ref.detach()[lo:hi]).abs().max() / ref.detach().abs().max()).item()
            eg = ((xl.grad - xs.grad[lo:hi]).abs().max() / xs.grad.abs().max()).item()
            err = torch.tensor([ef, eg], device="cuda"); dist.all_reduce(err, op=dist.ReduceOp.MAX)
            mso = timed(pgo)
            if wire is None and float(err.max()) > 1e-5:
                ok_blocks = False
            if rank == 0:
                print(f"  blocks/{transport}/{'bf16' if wire is not None else 'f32'}: fwd+bwd {mso:.3f} ms  "
                      f"max err fwd {float(err[0]):.1e} bwd {float(err[1]):.1e}", flush=True)
    for chunks in (2, 4, 8):
        for wire in (None, torch.bfloat16):
            pgo = S.PartitionedGraph(ei, n, halo_mode="chunks", halo_chunks=chunks, halo_dtype=wire)
            xl.grad = None
            oo = S.aggregate(xl, pgo, reduce="mean", act="relu"); oo.backward(go[lo:hi])
            ef = ((oo.detach() - ref.detach()[lo:hi]).abs().max() / ref.detach().abs().max()).item()
            eg = ((xl.grad - xs.grad[lo:hi]).abs().max() / xs.grad.abs().max()).item()
            err = torch.tensor([ef, eg], device="cuda"); dist.all_reduce(err, op=dist.ReduceOp.MAX)
            mso = timed(pgo)
            if wire is None and float(err.max()) > 1e-5:
                ok_blocks = False
            if rank == 0:
                print(f"  chunks/{chunks}/{'bf16' if wire is not None else 'f32'}: fwd+bwd {mso:.3f} ms  "
                      f"max err fwd {float(err[0]):.1e} bwd {float(err[1]):.1e}", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if (t.item() and ok_blocks) else 1)

if __name__ == "__main__":
    main()
