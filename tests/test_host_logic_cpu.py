"""CPU: host-side helpers that need no kernel -- zero-padding of odd widths, the rule that decides when a SAGE layer
runs as one autograd node, the bench's one-pass loss, the K slab plan of wide dense products."""
import importlib.util
import os

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_pad4_pads_columns_and_rows_with_zeros_and_is_differentiable():
    from stemgnn_b200 import dense
    x = torch.randn(5, 1433, requires_grad=True)
    p = dense.pad4(x)
    assert p.shape == (5, 1436) and torch.equal(p[:, :1433], x) and float(p[:, 1433:].abs().sum()) == 0.0
    p.sum().backward()
    assert torch.equal(x.grad, torch.ones_like(x))                     # the padding's gradient is dropped
    w = torch.randn(7, 1433)
    pw = dense.pad4(w, rows=True)
    assert pw.shape == (8, 1436) and torch.equal(pw[:7, :1433], w) and float(pw[7].abs().sum()) == 0.0
    assert dense.pad4(torch.randn(4, 8)) .shape == (4, 8)
    b = dense.pad4(torch.randn(7))
    assert b.shape == (8,) and float(b[7]) == 0.0
    assert dense.supported(1433, 768) and not dense.supported(0, 8)


def test_sage_layer_fusion_rule():
    from stemgnn_b200 import dense, sage

    class _G:                                   # stands in for a GraphStructure
        pass
    g = _G()
    x = torch.empty(10, 256, device="meta")
    x_cuda_like = type("T", (), {"size": lambda self, i: (10, 256)[i], "is_cuda": True, "dtype": torch.float32})()
    assert sage.fusable(x_cuda_like, g, None, 256, True)
    assert not sage.fusable(x_cuda_like, g, torch.empty(1), 256, True)            # edge features: composed form
    wide = type("T", (), {"size": lambda self, i: (10, 768)[i], "is_cuda": True, "dtype": torch.float32})()
    assert not sage.fusable(wide, g, None, 768, True)                              # 2 x 768 columns > one K5 call
    assert sage.fusable(wide, g, None, 768, False)
    odd = type("T", (), {"size": lambda self, i: (10, 130)[i], "is_cuda": True, "dtype": torch.float32})()
    assert not sage.fusable(odd, g, None, 256, True)
    g.fused = True                                                                  # sparse adj_t: spmm path
    assert not sage.fusable(x_cuda_like, g, None, 256, True)
    assert dense.K_SLAB_WIDE < dense.K_SLAB <= dense.K_SINGLE < dense.K_WIDE


def test_bench_mean_square_equals_square_mean_with_its_gradient():
    spec = importlib.util.spec_from_file_location("bench_ms", os.path.join(ROOT, "bench.py"))
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    q = torch.randn(300, 64, dtype=torch.float64, requires_grad=True)
    q2 = q.detach().clone().requires_grad_(True)
    (b.mean_square(q) * 3).backward()
    (q2.square().mean() * 3).backward()
    assert torch.allclose(q.grad, q2.grad, rtol=1e-12, atol=0)
    assert abs(float(b.mean_square(q.detach())) - float(q2.detach().square().mean())) < 1e-12
    assert abs(float(b.mean_square(q.detach(), 10)) - float(q2.detach().square().sum() / 10)) < 1e-9
