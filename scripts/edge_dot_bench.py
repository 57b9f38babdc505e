"""K8 timing at the bench shapes next to the torch expression it replaces (GPU only)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import stemgnn_b200 as S

dev = "cuda"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, reps=5):
    ts = []
    for r in range(reps + 2):
        flush.zero_()
        a, b = torch.cuda.Event(True), torch.cuda.Event(True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        if r >= 2:
            ts.append(a.elapsed_time(b) * 1e3)
    return min(ts)


g = torch.Generator().manual_seed(0)
for n, e, f in [(169343, 1166243, 256), (169343, 1166243, 128), (2449029, 61859140 // 4, 256)]:
    z = torch.randn(n, f, generator=g).to(dev)
    ei = torch.randint(0, n, (2, e), generator=g).to(dev)
    go = torch.randn(e, generator=g).to(dev)
    t_f = timeit(lambda: S.edge_dot(z, ei, True))
    t_ref = timeit(lambda: torch.sigmoid((z[ei[0]] * z[ei[1]]).sum(1)))
    zr = z.clone().requires_grad_(True)
    S.edge_dot(zr, ei, True).backward(go)          # builds and caches the two CSR groupings
    def fb():
        zr.grad = None
        S.edge_dot(zr, ei, True).backward(go)
    t_fb = timeit(fb)
    def fb_ref():
        zr.grad = None
        torch.sigmoid((zr[ei[0]] * zr[ei[1]]).sum(1)).backward(go)
    t_fb_ref = timeit(fb_ref)
    by = 2 * e * f * 4 + e * 20
    print(json.dumps(dict(nodes=n, edges=e, feat=f, fwd_us=round(t_f, 1), torch_fwd_us=round(t_ref, 1),
                          fwd_GBps_gather_model=round(by / t_f / 1e3, 1), edges_per_s=round(e / t_f * 1e6),
                          fwd_bwd_us=round(t_fb, 1), torch_fwd_bwd_us=round(t_fb_ref, 1))), flush=True)
