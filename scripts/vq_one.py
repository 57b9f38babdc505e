"""One (rows, K, D) point of the tcgen05 VQ search, for profiling: python scripts/vq_one.py rows K D [iters]."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import stemgnn_b200 as S

rows, K, D = (int(a) for a in sys.argv[1:4])
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 2
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.randn(rows, D, generator=g, device="cuda")
e = torch.randn(1, K, D, generator=g, device="cuda")
for _ in range(iters):
    ind, q, sq, st = S.vq_assign_raw(x, e, 1, S.METRIC_EUCLID, S.PATH_TENSOR, want_stats=True)
torch.cuda.synchronize()
print("ok", rows, K, D, st.tolist())
