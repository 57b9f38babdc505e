"""Error of the 3xTF32 dense kernel vs K (fp64 reference), next to torch's fp32 SGEMM."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import stemgnn_b200 as S
dev = "cuda"
g = torch.Generator().manual_seed(0)
for k in (16, 64, 256, 512, 1024, 4096):
    a = torch.randn(2048, k, generator=g).to(dev); w = (torch.randn(256, k, generator=g) / k ** 0.5).to(dev)
    ref = a.double() @ w.double().t()
    out = S.linear_raw([a], [w]); t32 = a @ w.t()
    sc = ref.abs().max()
    e = (out.double() - ref); e32 = (t32.double() - ref)
    print(f"K={k:5d} ours max {e.abs().max()/sc:.2e} rms {e.pow(2).mean().sqrt()/sc:.2e} mean {e.mean()/sc:+.2e} "
          f"bias_vs_sign {(e*ref.sign()).mean()/sc:+.2e} | torch max {e32.abs().max()/sc:.2e} rms {e32.pow(2).mean().sqrt()/sc:.2e}", flush=True)
# positive-only operands: a truncating accumulator shows as a negative bias growing with K
for k in (256, 1024):
    a = torch.rand(2048, k, generator=g).to(dev); w = (torch.rand(256, k, generator=g)).to(dev)
    ref = a.double() @ w.double().t(); out = S.linear_raw([a], [w]); t32 = a @ w.t()
    print(f"positive K={k}: ours mean rel {((out.double()-ref)/ref).mean():+.2e} max {((out.double()-ref)/ref).abs().max():.2e} | torch {((t32.double()-ref)/ref).mean():+.2e} {((t32.double()-ref)/ref).abs().max():.2e}")
