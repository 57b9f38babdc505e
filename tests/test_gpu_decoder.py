"""K8 (edge-wise inner product of InnerProductDecoder) against the oracle restatement and the
golden vectors of the reference's own class (tests/golden/dec_*.npz); fp32 bar 1e-5."""
import os

import numpy as np
import pytest
import torch
import torch.nn as nn

import stemgnn_b200 as S
from oracle.decoder_ref import edge_dot_grad_ref, edge_dot_ref
from tests.golden.cases import DEC_CASES

pytestmark = pytest.mark.gpu
DEV = "cuda"
TOL = 1e-5          # north star: fp32 numerics within 1e-5 of the reference


def _close(a, b, tol=TOL):
    a, b = torch.as_tensor(a).double().cpu(), torch.as_tensor(b).double().cpu()
    scale = b.abs().max().clamp_min(1e-30)
    err = ((a - b).abs().max() / scale).item() if b.numel() else 0.0
    assert err < tol, err


@pytest.mark.parametrize("name", sorted(DEC_CASES))
def test_decoder_module_matches_reference_golden(golden_dir, name):
    c = DEC_CASES[name]
    g = dict(np.load(os.path.join(golden_dir, name + ".npz")))
    dec = S.InnerProductDecoder(c["dim"], c["proj"]) if c["proj"] else S.InnerProductDecoder()
    dec.load_state_dict({k[6:]: torch.from_numpy(v) for k, v in g.items() if k.startswith("state.")})
    dec = dec.to(DEV)
    z = torch.from_numpy(g["z"]).to(DEV).requires_grad_(True)
    ei = torch.from_numpy(g["edge_index"]).to(DEV)
    out = dec(z, ei, sigmoid=c["sigmoid"])
    assert out.shape == (c["e"],) and out.dtype == torch.float32
    _close(out.detach(), g["out"])
    (out * torch.from_numpy(g["g_out"]).to(DEV)).sum().backward()
    _close(z.grad, g["g_z"])
    for k, p in dec.named_parameters():
        _close(p.grad, g["grad." + k])


@pytest.mark.parametrize("n,e,f", [(500, 4000, 256), (500, 4000, 128), (300, 2500, 64), (300, 2500, 32),
                                   (100, 999, 8), (100, 999, 4), (200, 1500, 100), (200, 1500, 30),
                                   (64, 77, 520), (50, 1, 16), (3, 40, 12)])
@pytest.mark.parametrize("sigmoid", [False, True])
def test_edge_dot_forward_backward_vs_oracle(n, e, f, sigmoid):
    g = torch.Generator().manual_seed(n * 7 + e + f)
    z = torch.randn(n, f, generator=g) * (1.0 if not sigmoid else 0.3)
    ei = torch.randint(0, n, (2, e), generator=g)
    if e >= 20:
        ei[:, :5] = ei[:, 5:10]              # duplicate edges
        ei[1, 10:15] = ei[0, 10:15]          # self loops
    go = torch.randn(e, generator=g)
    zr = z.clone().requires_grad_(True)
    ref = edge_dot_ref(zr, ei, sigmoid)
    (ref * go).sum().backward()
    zd = z.to(DEV).requires_grad_(True)
    out = S.edge_dot(zd, ei.to(DEV), sigmoid)
    (out * go.to(DEV)).sum().backward()
    # the dot of two N(0,1) rows has scale sqrt(F): compare at the scale of the values
    _close(out.detach(), ref.detach())
    _close(zd.grad, zr.grad)
    g64 = edge_dot_grad_ref(z.numpy(), ei.numpy(), go.numpy(), sigmoid)
    _close(zd.grad, g64)


def test_edge_dot_is_deterministic_and_handles_strided_rows():
    g = torch.Generator().manual_seed(11)
    n, e, f = 4000, 60000, 128
    wide = torch.randn(n, f + 32, generator=g).to(DEV)
    z = wide[:, 16:16 + f]                       # row stride 160 floats, 64-byte offset
    ei = torch.randint(0, n, (2, e), generator=g).to(DEV)
    a = S.edge_dot(z, ei, False)
    b = S.edge_dot(z.contiguous(), ei, False)
    assert torch.equal(a, b)
    zr = z.clone().requires_grad_(True)
    go = torch.randn(e, generator=g).to(DEV)
    g1 = torch.autograd.grad((S.edge_dot(zr, ei, True) * go).sum(), zr)[0]
    g2 = torch.autograd.grad((S.edge_dot(zr, ei, True) * go).sum(), zr)[0]
    assert torch.equal(g1, g2)                   # no atomics anywhere on the path


def test_edge_dot_empty_and_errors():
    z = torch.randn(10, 8, device=DEV)
    out = S.edge_dot(z, torch.zeros((2, 0), dtype=torch.int64, device=DEV), True)
    assert out.shape == (0,)
    with pytest.raises(S.StemGnnError):
        S.edge_dot(z.cpu(), torch.zeros((2, 3), dtype=torch.int64), True)          # no CPU fallback
    with pytest.raises(ValueError):
        S.edge_dot(z, torch.zeros((3, 3), dtype=torch.int64, device=DEV), True)
    with pytest.raises(TypeError):
        S.edge_dot(z.double(), torch.zeros((2, 3), dtype=torch.int64, device=DEV), True)
    bad = torch.tensor([[0, 11], [1, 2]], dtype=torch.int64, device=DEV)           # 11 >= N
    v = S.edge_dot(z, bad, False)
    assert torch.isfinite(v[0]) and torch.isnan(v[1])


def test_edge_dot_arxiv_scale_properties():
    """Full bench size (169,343 nodes, 1,166,243 edges, F=256): size-independent properties --
    symmetry in the two end points, linearity in z, and agreement with a chunked torch evaluation."""
    g = torch.Generator().manual_seed(3)
    n, e, f = 169343, 1166243, 256
    z = (torch.randn(n, f, generator=g) / 16).to(DEV)
    ei = torch.randint(0, n, (2, e), generator=g).to(DEV)
    v = S.edge_dot(z, ei, False)
    assert torch.equal(v, S.edge_dot(z, ei.flip(0).contiguous(), False))            # <a,b> == <b,a>, same bits
    assert torch.equal(S.edge_dot(z * 2, ei, False), v * 4)                         # exact scaling by powers of two
    ref = torch.cat([(z[ei[0, i:i + 200000]].double() * z[ei[1, i:i + 200000]].double()).sum(1)
                     for i in range(0, e, 200000)])
    _close(v, ref)
