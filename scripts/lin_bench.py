"""Timing of the dense kernel (K5) on the bench shapes, next to torch's fp32 path."""
import sys, os, json, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import stemgnn_b200 as S
dev = "cuda"
g = torch.Generator().manual_seed(0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timeit(fn, reps=5):
    ts = []
    for r in range(reps + 2):
        flush.zero_()
        s, t = torch.cuda.Event(True), torch.cuda.Event(True)
        s.record(); fn(); t.record(); torch.cuda.synchronize()
        if r >= 2: ts.append(s.elapsed_time(t) * 1e3)
    return min(ts)
MIX = len(sys.argv) > 1 and sys.argv[1] == "mix"      # the 768-wide shapes of the pretrain mix (C3) mini-batch
FWD = [(169343, 256, (128, 128)), (169343, 256, (256, 256)), (169343, 256, (256,)), (169343, 128, (256,)),
       (2449029 // 4, 256, (256, 256)), (169343, 512, (512, 512))]
WG = [(169343, 256, (128, 128)), (169343, 256, (256, 256)), (169343, 256, (256,))]
if MIX:
    FWD = [(43008, 768, (768, 768)), (43008, 768, (768,)), (43008, 3072, (768,)), (43008, 768, (3072,)), (43008, 768, (512,)), (43008, 768, (256,))]
    WG = [(43008, 768, (768, 768)), (43008, 3072, (768,)), (43008, 768, (3072,))]
for m, n, ks in FWD:
    srcs = [torch.randn(m, k, generator=g).to(dev) for k in ks]
    ws = [(torch.randn(n, k, generator=g) / k ** 0.5).to(dev) for k in ks]
    bias = torch.randn(n, generator=g).to(dev)
    out = torch.empty(m, n, device=dev)
    t_ours = timeit(lambda: S.linear_raw(srcs, ws, bias, out=out))
    def ref():
        o = torch.nn.functional.linear(srcs[0], ws[0], bias)
        for a, w in zip(srcs[1:], ws[1:]): o = o + torch.nn.functional.linear(a, w)
        return o
    t_ref = timeit(ref)
    kt = sum(ks); fl = 2 * m * n * kt; by = m * kt * 4 + m * n * 4
    print(json.dumps(dict(m=m, n=n, k=ks, ours_us=round(t_ours, 1), torch_us=round(t_ref, 1),
                          tflops_fp32_equiv=round(fl / t_ours / 1e6, 1), tf32_mma_tflops=round(3 * fl / t_ours / 1e6, 1),
                          hbm_GBps=round(by / t_ours / 1e3))), flush=True)
print("wgrad", flush=True)
for m, n, ks in WG:
    gr = torch.randn(m, n, generator=g).to(dev)
    srcs = [torch.randn(m, k, generator=g).to(dev) for k in ks]
    t_ours = timeit(lambda: S.linear_wgrad_raw(gr, srcs, want_bias=False))
    t_bias = timeit(lambda: S.linear_wgrad_raw(gr, srcs, want_bias=True))
    t_ref = timeit(lambda: [gr.t() @ a for a in srcs])
    kt = sum(ks); fl = 2 * m * n * kt
    print(json.dumps(dict(m=m, n=n, k=ks, ours_us=round(t_ours, 1), with_bias_us=round(t_bias, 1), torch_us=round(t_ref, 1),
                          tf32_mma_tflops=round(3 * fl / t_ours / 1e6, 1))), flush=True)
