"""Print the tcgen05 VQ path's statistics and timing for a few (rows, K, D) points (GPU only)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import stemgnn_b200 as S

def run(rows, K, D, metric, scale=None, iters=5, exact=True):
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.randn(rows, D, generator=g, device="cuda")
    if scale is None:
        e = (torch.rand(1, K, D, generator=g, device="cuda") * 2 - 1) * (6.0 / D) ** 0.5
    else:
        e = torch.randn(1, K, D, generator=g, device="cuda") * scale
    if metric == S.METRIC_DOT:
        x = torch.nn.functional.normalize(x, dim=-1); e = torch.nn.functional.normalize(e, dim=-1)
    for path in (S.PATH_TENSOR, S.PATH_EXACT) if exact and rows * K * D < 1e12 else (S.PATH_TENSOR,):
        S.vq_assign_raw(x, e, 1, metric, path, want_stats=True)
        torch.cuda.synchronize()
        t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
        t0.record()
        for _ in range(iters):
            ind, q, sq, st = S.vq_assign_raw(x, e, 1, metric, path, want_stats=True)
        t1.record(); torch.cuda.synchronize()
        ms = t0.elapsed_time(t1) / iters
        if os.environ.get("VQ_KERNELS") and path == S.PATH_TENSOR:      # per-kernel device times (CUPTI)
            from torch.profiler import profile, ProfilerActivity
            with profile(activities=[ProfilerActivity.CUDA]) as prof:
                for _ in range(iters):
                    S.vq_assign_raw(x, e, 1, metric, path, want_stats=True)
                torch.cuda.synchronize()
            for ev in sorted(prof.key_averages(), key=lambda a: -a.device_time_total)[:6]:
                print(f"    {ev.device_time_total / ev.count:10.1f} us x{ev.count:<3d} {ev.key[:90]}")
        print(f"rows={rows} K={K} D={D} metric={metric} path={path}: {ms:.3f} ms  "
              f"{2*rows*K*D/ms/1e9:.1f} TFLOP/s  rescored={st[0].item()} rescanned={st[1].item()}", flush=True)

if __name__ == "__main__":
    if len(sys.argv) > 1:       # python scripts/vq_stats.py rows K D [iters]: the tensor path at one point
        r, k, d = (int(a) for a in sys.argv[1:4])
        run(r, k, d, S.METRIC_EUCLID, scale=1.0, iters=int(sys.argv[4]) if len(sys.argv) > 4 else 3, exact=False)
        sys.exit(0)
    run(169343, 512, 256, S.METRIC_EUCLID)
    run(169343, 512, 256, S.METRIC_EUCLID, scale=1.0)
    run(169343, 128, 768, S.METRIC_DOT, scale=1.0)
    run(1000000, 4096, 128, S.METRIC_EUCLID, scale=1.0, iters=2)
    run(1000000, 1024, 256, S.METRIC_EUCLID, scale=1.0, iters=3)
    run(1000000, 16384, 128, S.METRIC_EUCLID, scale=1.0, iters=2)
    run(1000000, 16384, 256, S.METRIC_EUCLID, scale=1.0, iters=2)
    run(1000000, 16384, 512, S.METRIC_EUCLID, scale=1.0, iters=2)
