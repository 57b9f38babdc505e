"""stemgnn_bn_stats_static -- the entry point that allocates nothing (block partials in a `static __device__`
array inside the library) -- against stemgnn_bn_stats with a caller workspace, through the C-ABI, on two calls
in a row.  Collected last on purpose; passes alone in a fresh process and
after the other tests."""
import ctypes as C

import pytest
import torch

from stemgnn_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _stats_workspace(x):
    n, f = x.shape
    h = _lib.lib()
    nb = C.c_size_t()
    _lib.check(h.stemgnn_bn_workspace(n, f, C.byref(nb)))
    ws = torch.empty(max(nb.value, 1), dtype=torch.uint8, device=x.device)
    mean, var = torch.empty(f, device=x.device), torch.empty(f, device=x.device)
    _lib.check(h.stemgnn_bn_stats(n, f, x.data_ptr(), x.stride(0), mean.data_ptr(), var.data_ptr(), ws.data_ptr(),
                                  nb.value, _lib.stream()))
    return mean, var


def _stats_static(x, mean, var):
    n, f = x.shape
    return _lib.lib().stemgnn_bn_stats_static(n, f, x.data_ptr(), x.stride(0), mean.data_ptr(), var.data_ptr(),
                                              _lib.stream())


def test_static_scratch_statistics_equal_the_workspace_entry_point_twice_in_a_row():
    g = torch.Generator().manual_seed(11)
    # two different shapes back to back: the second call finds the first call's partials in the shared array
    xs = [(torch.randn(169343, 256, generator=g) * 2 + 3).to(DEV), (torch.randn(5000, 128, generator=g) - 7).to(DEV)]
    outs = [(torch.full((x.shape[1],), float("nan"), device=DEV), torch.full((x.shape[1],), float("nan"), device=DEV))
            for x in xs]
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    for x, (mean, var) in zip(xs, outs):                    # the two calls, nothing in between
        _lib.check(_stats_static(x, mean, var))
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before          # no torch allocation on the way
    for x, (mean, var) in zip(xs, outs):
        mean_w, var_w = _stats_workspace(x)
        # same kernels, same block plan, same order of additions: identical bits (tolerance of the K7 tests: 1e-5)
        assert torch.equal(mean, mean_w) and torch.equal(var, var_w)
        ref_mean, ref_var = x.double().mean(0), x.double().var(0, unbiased=False)
        assert ((mean.double() - ref_mean).abs().max() / ref_mean.abs().max()).item() < 1e-5
        assert ((var.double() - ref_var).abs().max() / ref_var.abs().max()).item() < 1e-5


def test_static_scratch_strided_rows_and_repeat():
    g = torch.Generator().manual_seed(12)
    big = torch.randn(3000, 520, generator=g).to(DEV)
    x = big[:, 8:264]                                       # row stride 520, 256 columns
    mean, var = torch.empty(256, device=DEV), torch.empty(256, device=DEV)
    _lib.check(_stats_static(x, mean, var))
    m1, v1 = mean.clone(), var.clone()
    _lib.check(_stats_static(x, mean, var))
    mean_w, var_w = _stats_workspace(x)
    assert torch.equal(mean, m1) and torch.equal(var, v1)
    assert torch.equal(mean, mean_w) and torch.equal(var, var_w)


def test_static_scratch_rejects_sizes_beyond_its_limit():
    f = 4096                                                # 444 row blocks x 2 x 4096 floats = 14.5 MB > 4 MiB
    x = torch.zeros(100000, f, device=DEV)
    mean, var = torch.zeros(f, device=DEV), torch.zeros(f, device=DEV)
    rc = _stats_static(x, mean, var)
    assert rc == -1
    msg = _lib.lib().stemgnn_last_error().decode()
    assert "built-in scratch" in msg and "stemgnn_bn_stats" in msg
    # the same width fits while the row count keeps the block count low (64 rows per block at least)
    xs = torch.randn(1024, f, device=DEV)
    _lib.check(_stats_static(xs, mean, var))
    mean_w, var_w = _stats_workspace(xs)
    assert torch.equal(mean, mean_w) and torch.equal(var, var_w)
