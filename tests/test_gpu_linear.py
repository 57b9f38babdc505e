"""K5 (tcgen05 3xTF32 dense layers) against torch: the fp64 product is the reference, torch's own
fp32 SGEMM gives the error scale an fp32 implementation has on the same inputs; north_star's
bound for fp32 numerics is 1e-5."""
import pytest
import torch

import stemgnn_b200 as S

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rel_err(got, ref64):
    """max |got - ref| over (|A| |W|^T) row scale: the natural unit of a dot-product error."""
    return ((got.double() - ref64).abs().max() / ref64.abs().max().clamp_min(1e-30)).item()


@pytest.mark.parametrize("m,n,ks", [
    (1000, 256, (128,)), (1000, 256, (256, 256)), (169, 64, (32, 64)), (256, 256, (16,)),
    (257, 272, (260,)), (5000, 512, (128, 128)), (77, 8, (4,)), (3, 300, (520, 36)),
    (70000, 256, (128, 128)),
])
def test_linear_matches_fp64_product(m, n, ks):
    g = torch.Generator().manual_seed(m + n)
    srcs = [torch.randn(m, k, generator=g).to(DEV) for k in ks]
    ws = [(torch.randn(n, k, generator=g) / k ** 0.5).to(DEV) for k in ks]
    bias = torch.randn(n, generator=g).to(DEV)
    out = S.linear_raw(srcs, ws, bias)
    ref = sum(a.double() @ w.double().t() for a, w in zip(srcs, ws)) + bias.double()
    t32 = sum(a @ w.t() for a, w in zip(srcs, ws)) + bias
    err, err32 = _rel_err(out, ref), _rel_err(t32, ref)
    assert err < 1e-5, (err, err32)
    # the accuracy model of the kernel: the tensor-core accumulator truncates, about 7.5e-9 of the
    # result scale per accumulated column (measured, scripts/lin_err.py), on top of fp32 rounding
    assert err < 1.5e-6 + 1.0e-8 * sum(ks), (err, err32)
    # elementwise against the |a|.|w| scale of each output
    scale = sum(a.double().abs() @ w.double().abs().t() for a, w in zip(srcs, ws)) + bias.double().abs()
    assert ((out.double() - ref).abs() / scale).max().item() < 2e-6


def test_linear_transposed_weights_and_strided_rows():
    g = torch.Generator().manual_seed(3)
    m, n, k = 1500, 128, 256
    big = torch.randn(m, k + 64, generator=g).to(DEV)
    a = big[:, 32:32 + k]                       # row stride k+64, 16-byte aligned offset
    w = torch.randn(k, n, generator=g).to(DEV)  # [k, n]: a @ w
    out = S.linear_raw([a], [w], None, w_transposed=True)
    ref = a.double() @ w.double()
    assert _rel_err(out, ref) < 1e-5
    buf = torch.zeros(m, n + 16, device=DEV)
    S.linear_raw([a], [w], None, w_transposed=True, out=buf[:, :n])
    assert torch.equal(buf[:, :n], out) and buf[:, n:].abs().max().item() == 0


def test_linear_extreme_magnitudes():
    """fp32 range is preserved: tiny gradients and large activations need no scaling."""
    g = torch.Generator().manual_seed(4)
    a = (torch.randn(600, 128, generator=g) * 1e-20).to(DEV)
    w = (torch.randn(256, 128, generator=g) * 1e12).to(DEV)
    out = S.linear_raw([a], [w])
    ref = a.double() @ w.double().t()
    assert _rel_err(out, ref) < 1e-5
    a2 = (torch.randn(600, 128, generator=g) * 1e15).to(DEV)
    out2 = S.linear_raw([a2], [w * 1e-10])
    assert _rel_err(out2, a2.double() @ (w * 1e-10).double().t()) < 1e-5


def test_linear_rejects_bad_arguments():
    a = torch.randn(8, 6, device=DEV)
    w = torch.randn(4, 6, device=DEV)
    with pytest.raises(ValueError):
        S.linear_raw([a], [w])                   # k % 4 != 0
    with pytest.raises(S.StemGnnError):
        S.linear_raw([a.cpu()], [w])


def _sum_scale(g, a):
    return (g.double().abs().t() @ a.double().abs())


@pytest.mark.parametrize("m,n,ks", [
    (1000, 256, (128,)), (5000, 256, (256, 256)), (513, 64, (32, 64)), (100, 8, (4,)),
    (3000, 272, (260, 36)), (40000, 256, (128, 128)), (512, 512, (512,)), (1, 16, (16,)),
])
def test_wgrad_matches_fp64_product(m, n, ks):
    g_ = torch.Generator().manual_seed(m * 3 + n)
    g = torch.randn(m, n, generator=g_).to(DEV)
    srcs = [torch.randn(m, k, generator=g_).to(DEV) for k in ks]
    gws, gb = S.linear_wgrad_raw(g, srcs, want_bias=True)
    for a, gw in zip(srcs, gws):
        ref = g.double().t() @ a.double()
        t32 = g.t() @ a
        err, err32 = _rel_err(gw, ref), _rel_err(t32, ref)
        assert err < 1e-5, (err, err32)
        # against the |g|^T |a| scale of each entry: a few 1e-6 (truncating accumulator, 512-row slabs)
        assert ((gw.double() - ref).abs() / _sum_scale(g, a).clamp_min(1e-30)).max().item() < 5e-6
    refb = g.double().sum(0)
    assert ((gb.double() - refb).abs().max() / refb.abs().max()).item() < 1e-5


def test_wgrad_is_deterministic_and_handles_positive_data():
    g_ = torch.Generator().manual_seed(11)
    m, n, k = 30000, 256, 256
    g = torch.rand(m, n, generator=g_).to(DEV)          # all positive: the truncation bias adds up
    a = torch.rand(m, k, generator=g_).to(DEV)
    (gw1,), _ = S.linear_wgrad_raw(g, [a])
    (gw2,), _ = S.linear_wgrad_raw(g, [a])
    assert torch.equal(gw1, gw2)
    ref = g.double().t() @ a.double()
    assert ((gw1.double() - ref) / ref).abs().max().item() < 1e-5


def test_linear_sum_autograd_matches_torch():
    g_ = torch.Generator().manual_seed(5)
    m, n, k0, k1 = 4000, 256, 128, 256
    x0 = torch.randn(m, k0, generator=g_).to(DEV).requires_grad_(True)
    x1 = torch.randn(m, k1, generator=g_).to(DEV).requires_grad_(True)
    w0 = (torch.randn(n, k0, generator=g_) / 11).to(DEV).requires_grad_(True)
    w1 = (torch.randn(n, k1, generator=g_) / 16).to(DEV).requires_grad_(True)
    b = torch.randn(n, generator=g_).to(DEV).requires_grad_(True)
    go = torch.randn(m, n, generator=g_).to(DEV)
    out = S.linear_sum([x0, x1], [w0, w1], b)
    out.backward(go)
    got = [out.detach()] + [t.grad.clone() for t in (x0, x1, w0, w1, b)]
    ts = [t.detach().double().requires_grad_(True) for t in (x0, x1, w0, w1, b)]
    ref = ts[0] @ ts[2].t() + ts[1] @ ts[3].t() + ts[4]
    ref.backward(go.double())
    refs = [ref.detach()] + [t.grad for t in ts]
    for name, a, r in zip(("out", "gx0", "gx1", "gw0", "gw1", "gb"), got, refs):
        assert _rel_err(a, r) < 1e-5, name


def test_sage_conv_uses_tensor_core_linear_and_matches_torch_layers():
    import torch.nn.functional as F
    g_ = torch.Generator().manual_seed(6)
    n, e, fin, fout = 3000, 20000, 128, 256
    ei = torch.randint(0, n, (2, e), generator=g_).to(DEV)
    x = torch.randn(n, fin, generator=g_).to(DEV)
    torch.manual_seed(0)
    conv = S.MySAGEConv(fin, fout).to(DEV)
    S._lib.TRACE = []
    try:
        out = conv(x, ei)
        names = [r[0] for r in S._lib.TRACE]
    finally:
        S._lib.TRACE = None
    assert "linear" in names
    agg = S.aggregate(x, S.GraphStructure(ei, n), reduce="mean", act="relu")
    ref = F.linear(agg.double(), conv.lin_l.weight.double(), conv.lin_l.bias.double()) + \
        F.linear(x.double(), conv.lin_r.weight.double())
    assert _rel_err(out, ref) < 1e-5


def test_wgrad_large_m_reuses_partial_slots():
    """More work items per CTA than partial-tile slots (the read-modify-write epilogue path):
    620k rows = 1,211+ slabs over 148 CTAs against 6 slots."""
    g_ = torch.Generator().manual_seed(21)
    m, n, k = 620_000, 256, 128
    g = torch.randn(m, n, generator=g_).to(DEV)
    a = torch.randn(m, k, generator=g_).to(DEV)
    (gw,), gb = S.linear_wgrad_raw(g, [a], want_bias=True)
    ref = torch.zeros(n, k, dtype=torch.float64, device=DEV)
    for lo in range(0, m, 100_000):                      # fp64 reference in chunks (memory)
        ref += g[lo:lo + 100_000].double().t() @ a[lo:lo + 100_000].double()
    t32 = g.t() @ a
    err, err32 = _rel_err(gw, ref), _rel_err(t32, ref)
    assert err < 1e-5, (err, err32)
    assert ((gb.double() - g.double().sum(0)).abs().max() / g.double().sum(0).abs().max()).item() < 1e-5
    (gw2,), _ = S.linear_wgrad_raw(g, [a])
    assert torch.equal(gw, gw2)


def test_linear_large_m():
    g_ = torch.Generator().manual_seed(22)
    m, n, k = 1_300_000, 128, 64
    a = torch.randn(m, k, generator=g_).to(DEV)
    w = (torch.randn(n, k, generator=g_) / 8).to(DEV)
    out = S.linear_raw([a], [w])
    idx = torch.randint(0, m, (4096,), generator=g_).to(DEV)
    ref = a[idx].double() @ w.double().t()
    assert _rel_err(out[idx], ref) < 1e-5
    assert _rel_err(out[-300:], a[-300:].double() @ w.double().t()) < 1e-5


@pytest.mark.parametrize("ks", [(4096,), (1024, 1024), (768, 520)])
def test_wide_products_are_slabbed_and_hold_the_bound(ks):
    g = torch.Generator().manual_seed(sum(ks))
    m, n = 3000, 256
    srcs = [torch.randn(m, k, generator=g).to(DEV) for k in ks]
    ws = [(torch.randn(n, k, generator=g) / k ** 0.5).to(DEV) for k in ks]
    bias = torch.randn(n, generator=g).to(DEV)
    out = S.linear_raw(srcs, ws, bias)
    ref = sum(a.double() @ w.double().t() for a, w in zip(srcs, ws)) + bias.double()
    assert _rel_err(out, ref) < 1e-5
    # positive operands: the truncation bias adds up coherently -- the hard case
    srcs = [t.abs() for t in srcs]; ws = [t.abs() for t in ws]
    out = S.linear_raw(srcs, ws, None)
    ref = sum(a.double() @ w.double().t() for a, w in zip(srcs, ws))
    assert ((out.double() - ref) / ref).abs().max().item() < 1e-5
    # transposed weights take the same path
    wt = [w.t().contiguous() for w in ws]
    out_t = S.linear_raw(srcs, wt, None, w_transposed=True)
    assert torch.equal(out_t, out)


def test_row_div_epilogue_equals_a_separate_ieee_division():
    """stemgnn_linear(row_div): out[m, :] = (A @ W) / d[m] in the epilogue (ABI 6) -- the same bits as the product
    followed by stemgnn_row_scale (MeanAggregation's g_out / count of the backward), with and without transposed W."""
    import stemgnn_b200 as S
    from stemgnn_b200 import dense, ops
    g = torch.Generator().manual_seed(4)
    m, n, k = 5003, 256, 128
    a = torch.randn(m, k, generator=g).cuda()
    w = torch.randn(n, k, generator=g).cuda()
    d = torch.randint(1, 40, (m,), generator=g).float().cuda()
    plain = dense.linear_raw([a], [w], None)
    fused = dense.linear_raw([a], [w], None, row_div=d)
    assert torch.equal(fused, ops._row_scale(plain, d))
    gt = torch.randn(m, n, generator=g).cuda()
    assert torch.equal(dense.linear_raw([gt], [w], None, w_transposed=True, row_div=d),
                       ops._row_scale(dense.linear_raw([gt], [w], None, w_transposed=True), d))
    with pytest.raises(S.StemGnnError):
        dense.linear_raw([a], [w], None, out=plain.clone(), accumulate=True, row_div=d)
