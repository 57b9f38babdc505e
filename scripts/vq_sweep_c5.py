"""BASELINE configs[4]: the VQ bottleneck sweep -- 10M node embeddings x codebook
sizes 256..16,384 x dims 128..512, one training-mode pass = nearest-code search (K3: tcgen05
argmin-GEMM + gather + commitment partial) + per-code statistics (K4) + EMA update, Euclidean, 1 head.
Rows are processed in chunks of 2.5M (the statistics of the chunks are summed, one EMA update per
pass -- what a replica does with its shard before the all-reduce).  Under torchrun the rows are
sharded over the ranks and the packed [counts | sums] buffer is all-reduced over NCCL
(model/vq.py:494,499).  One JSON line per (K, D) point plus a summary line.
    python scripts/vq_sweep_c5.py [--rows 10000000] [--cpu-rows 100000]"""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

KS, DS = (256, 1024, 4096, 16384), (128, 256, 512)
CHUNK = 2_500_000


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--cpu-rows", type=int, default=100_000)
    ap.add_argument("--iters", type=int, default=2)
    args = ap.parse_args()
    import torch.distributed as dist
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    import stemgnn_b200 as S
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    tc_peak, hbm_peak = peaks.get("bf16_tflops", 1590.0), peaks.get("hbm_gbs", 6650.0)
    rows = args.rows // world                                   # this rank's shard
    results = []
    for D in DS:
        g = torch.Generator(device=dev).manual_seed(1 + rank)
        x = torch.randn(rows, D, generator=g, device=dev)
        for K in KS:
            torch.manual_seed(0)
            embed = torch.empty(1, K, D)
            torch.nn.init.kaiming_uniform_(embed)
            embed = embed.to(dev)
            embed_avg, cluster = embed.clone(), torch.zeros(1, K, device=dev)

            def one_pass():
                flat_total = None
                for lo in range(0, rows, CHUNK):
                    xc = x[lo:lo + CHUNK]
                    ind, q, sq, _ = S.vq_assign_raw(xc, embed, 1, S.METRIC_EUCLID, S.PATH_AUTO)
                    counts, sums, flat = S.vq_code_stats(xc, ind, (1, K, D), 1)
                    flat_total = flat if flat_total is None else flat_total.add_(flat)
                if world > 1:
                    dist.all_reduce(flat_total)
                S.vq_ema_apply(flat_total[:K].view(1, K), flat_total[K:].view(1, K, D), cluster, embed_avg, embed,
                               0.8, 1e-5, False)

            one_pass()
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record()
            for _ in range(args.iters):
                one_pass()
            t1.record()
            torch.cuda.synchronize()
            ms = torch.tensor([t0.elapsed_time(t1) / args.iters], device=dev)
            if world > 1:
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            ms = float(ms)
            total = rows * world
            tflops = 2.0 * total * K * D / ms / 1e9
            gbs = total * D * 4 * 3 / ms / 1e6                   # x read by K3 and K4, quantize written
            # K = 256 with fp32 inputs sits left of the ridge -> reported against HBM
            bound = "hbm" if K <= 256 else "tensor"
            frac = gbs / (hbm_peak * world) if bound == "hbm" else tflops / (tc_peak * world)
            rec = dict(K=K, D=D, rows=total, n_gpus=world, ms=round(ms, 3), nodes_per_s=total / ms * 1e3,
                       tflops=round(tflops, 1), gbs=round(gbs, 1), bound=bound, frac=round(frac, 4))
            results.append(rec)
            if rank == 0:
                print(json.dumps(rec), flush=True)
        del x
        torch.cuda.empty_cache()
    if rank == 0 and args.cpu_rows > 0:
        from oracle.vq_ref import vq_forward_ref
        torch.set_num_threads(os.cpu_count() or 1)
        for K, D in ((1024, 256), (16384, 256)):
            xc = torch.randn(args.cpu_rows, D)
            emb = torch.empty(1, K, D)
            torch.nn.init.kaiming_uniform_(emb)
            st = dict(embed=emb, embed_avg=emb.clone(), cluster_size=torch.zeros(1, K))
            t = time.perf_counter()
            vq_forward_ref(xc, st, training=True, ema_update=True, decay=0.8)
            dt = time.perf_counter() - t
            print(json.dumps(dict(cpu_baseline=dict(K=K, D=D, rows=args.cpu_rows, nodes_per_s=args.cpu_rows / dt,
                                                    cores=os.cpu_count(), kind="port",
                                                    sample=f"{args.cpu_rows} rows, oracle/vq_ref.py (the reference "
                                                    "materialises the N x K distance matrix)"))), flush=True)
    if rank == 0:
        best = max(r["frac"] for r in results if r["bound"] == "tensor")
        print(json.dumps(dict(summary="C5 VQ sweep", n_gpus=world, rows=args.rows,
                              best_tensor_frac=best, peak_tflops=tc_peak, peak_gbs=hbm_peak)), flush=True)
    if world > 1:
        dist.barrier()
        os._exit(0)


if __name__ == "__main__":
    main()
