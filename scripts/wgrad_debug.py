import sys, os, ctypes as C, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import stemgnn_b200 as S
from stemgnn_b200 import _lib
dev = "cuda"
torch.manual_seed(0)
for (m, n, k) in [(16, 256, 256), (512, 256, 256), (1024, 256, 256)]:
    g = torch.randn(m, n, device=dev); a_ = torch.randn(m, k, device=dev)
    gw = torch.full((n, k), 7.0, device=dev)
    a = _lib.LinearWgradArgs()
    a.m, a.n, a.n_src = m, n, 1
    a.g, a.ldg = g.data_ptr(), n
    a.a[0], a.lda[0], a.k[0] = a_.data_ptr(), k, k
    a.gw[0], a.ldgw[0] = gw.data_ptr(), k
    L = _lib.lib(); need = C.c_size_t(0)
    _lib.check(L.stemgnn_linear_wgrad_workspace(C.byref(a), C.byref(need)))
    ws = torch.full((need.value,), 0x11, dtype=torch.uint8, device=dev)
    _lib.check(L.stemgnn_linear_wgrad(C.byref(a), ws.data_ptr(), ws.numel(), _lib.stream()))
    torch.cuda.synchronize()
    base = (ws.data_ptr() + 255) // 256 * 256 - ws.data_ptr()
    wsf = ws[base:base + (need.value - 256) // 4 * 4].view(torch.float32)
    slabs = (m + 511) // 512; grid = min(slabs, 148)
    part = wsf[:grid * 65536].view(grid, 256, 256)
    seen = ws[base + grid * 262144: base + grid * 262144 + grid * 4].view(torch.int32)
    ref = g.double().t() @ a_.double()
    print(f"m={m}: gw[0,:4]={gw[0,:4].tolist()} ref={ref[0,:4].tolist()} seen={seen.tolist()[:4]} part0[0,:4]={part[0,0,:4].tolist()}")
    print("   err vs ref:", ((part.double().sum(0) - ref).abs().max() / ref.abs().max()).item(), "gw err", ((gw.double() - ref).abs().max() / ref.abs().max()).item())
    # where does part differ: check transposition / block structure
    p0 = part.double().sum(0)
    print("   err vs ref.T:", ((p0 - ref.t()).abs().max() / ref.abs().max()).item())
