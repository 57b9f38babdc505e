"""CPU: the C-ABI library loads and exports every symbol include/stemgnn_b200.h declares; the
host-side modules keep the reference's state_dict layout.  No compute calls (no GPU here)."""
import os
import re

import pytest
import torch

import stemgnn_b200 as S
from stemgnn_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "stemgnn_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(stemgnn_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    names = _declared_symbols()
    assert len(names) >= 12
    assert set(names) == set(_lib.SIGNATURES), "ctypes table and header disagree"
    h = _lib.lib()
    for n in names:
        assert hasattr(h, n), n
    assert h.stemgnn_abi_version() == _lib.ABI_VERSION


def test_sm_reserve_is_a_host_side_setting_that_returns_the_previous_value():
    h = _lib.lib()
    assert h.stemgnn_set_sm_reserve(32) == 0
    assert h.stemgnn_set_sm_reserve(-5) == 32            # negative values clamp to 0
    assert h.stemgnn_set_sm_reserve(0) == 0


def test_argument_errors_are_reported_without_a_gpu():
    import ctypes as C
    h = _lib.lib()
    need = C.c_size_t(0)
    assert h.stemgnn_csr_build_workspace(-1, 4, C.byref(need)) == -1
    assert b"negative" in h.stemgnn_last_error()
    assert h.stemgnn_csr_build_workspace(1000, 10, C.byref(need)) == 0 and need.value > 0
    # maximum sizes: the structure's indices are int32
    assert h.stemgnn_csr_build_workspace(2 ** 31 - 1, 10, C.byref(need)) == -1
    assert h.stemgnn_csr_build_workspace(10, 2 ** 31, C.byref(need)) == -1
    assert h.stemgnn_csr_build_workspace(2 ** 31 - 2, 2 ** 31 - 2, C.byref(need)) == 0 and need.value > 2 ** 34
    # hub scratch: four counters and one slot per possible row of >= 256 edges
    assert h.stemgnn_aggregate_hub_workspace(0, C.byref(need)) == 0 and need.value == 4 * 5
    assert h.stemgnn_aggregate_hub_workspace(1166243, C.byref(need)) == 0 and need.value == 4 * (4 + 1166243 // 256 + 1)
    assert h.stemgnn_aggregate_hub_workspace(-5, C.byref(need)) == -1
    # a scratch smaller than that is refused before anything is launched
    a = _lib.AggregateArgs(n_rows=10, n_gather_rows=10, num_edges=1000, feat=8, rowptr=16, col=16, gathered=16,
                           out=16, hub_ws=16, hub_ws_bytes=8)
    assert h.stemgnn_aggregate(C.byref(a), None) == -1 and b"hub_ws" in h.stemgnn_last_error()


def test_no_cpu_fallback():
    x = torch.randn(4, 8)
    ei = torch.zeros(2, 3, dtype=torch.int64)
    with pytest.raises(S.StemGnnError):
        S.GraphStructure(ei, 4)
    vq = S.VectorQuantize(dim=8, codebook_size=4)
    with pytest.raises(S.StemGnnError):
        vq(x)


def test_state_dict_layout_matches_reference():
    vq = S.VectorQuantize(dim=768, codebook_size=128, codebook_dim=768, heads=4,
                          separate_codebook_per_head=True, use_cosine_sim=True,
                          orthogonal_reg_weight=1, orthogonal_reg_max_codes=32, ema_update=False,
                          commitment_weight=10, kmeans_init=False)
    sd = {k: tuple(v.shape) for k, v in vq.state_dict().items()}
    assert sd["_codebook.embed"] == (4, 128, 768) and sd["_codebook.embed_avg"] == (4, 128, 768)
    assert sd["_codebook.cluster_size"] == (4, 128) and sd["_codebook.initted"] == (1,)
    assert sd["project_in.weight"] == (3072, 768) and sd["project_out.weight"] == (768, 3072)
    assert isinstance(vq._codebook.embed, torch.nn.Parameter) and vq.codebook.shape == (4, 128, 768)
    vq1 = S.VectorQuantize(dim=16, codebook_size=8)
    assert vq1.codebook.shape == (8, 16) and not isinstance(vq1._codebook.embed, torch.nn.Parameter)
    enc = S.Encoder(20, 24, torch.nn.ReLU, 2, backbone="sage", normalize="batch", moe=True,
                    moe_layers="last")
    keys = set(enc.state_dict())
    assert {"layers.0.lin_l.weight", "layers.0.lin_l.bias", "layers.0.lin_r.weight",
            "layers.1.weights", "env_encoders.0.weight", "norms.1.running_mean"} <= keys
    assert enc.moe_layer_flags == [False, True]
    import copy
    copy.deepcopy(enc)


def test_unsupported_options_raise():
    with pytest.raises(NotImplementedError):
        S.VectorQuantize(dim=8, codebook_size=4, affine_param=True)
    with pytest.raises(NotImplementedError):
        S.VectorQuantize(dim=8, codebook_size=4, stochastic_sample_codes=True)
    with pytest.raises(NotImplementedError):
        S.Encoder(8, 8, torch.nn.ReLU, 1, backbone="gat")


def test_every_entry_point_validates_its_arguments_before_touching_the_device():
    """Error behaviour of the boundary (no GPU needed: every check below precedes the first CUDA call):
    a bad argument returns STEMGNN_EINVAL (-1) and leaves a message for stemgnn_last_error()."""
    import ctypes as C
    h = _lib.lib()
    need = C.c_size_t(0)

    def refused(rc, *words):
        msg = h.stemgnn_last_error().decode()
        assert rc == -1, (rc, msg)
        assert all(w in msg for w in words), msg

    # K3 VQ assignment
    a = _lib.VqAssignArgs(heads=1, embed_heads=1, rows=10, codes=8, dim=16, metric=0, path=0)
    assert h.stemgnn_vq_assign_workspace(C.byref(a), C.byref(need)) == 0
    for field, value, words in [("metric", 2, ("metric",)), ("path", 3, ("path",)), ("embed_heads", 3, ("embed_heads",)),
                                ("codes", 0, ("bad sizes",)), ("codes", 1 << 24, ("too large",)),
                                ("rows", 2 ** 31, ("2^31",))]:
        b = _lib.VqAssignArgs(heads=1, embed_heads=1, rows=10, codes=8, dim=16, metric=0, path=0)
        setattr(b, field, value)
        refused(h.stemgnn_vq_assign_workspace(C.byref(b), C.byref(need)), *words)
    refused(h.stemgnn_vq_assign(C.byref(a), None, 0, None), "NULL")
    # K7 BatchNorm: rows of float4, an empty batch has no statistics, dropout needs its seed
    refused(h.stemgnn_bn_workspace(-1, 8, C.byref(need)), "bad arguments")
    refused(h.stemgnn_bn_stats(16, 6, 16, 6, 16, 16, 16, 1 << 20, None), "multiple of 4")
    refused(h.stemgnn_bn_stats(0, 8, 16, 8, 16, 16, 16, 1 << 20, None), "empty batch")
    refused(h.stemgnn_bn_act_forward(4, 8, 16, 8, 16, 16, 16, 16, 1, 0.5, None, 16, 8, None), "seed")
    refused(h.stemgnn_bn_act_forward(4, 8, 16, 8, 16, 16, 16, 16, 1, 1.0, 16, 16, 8, None), "[0, 1)")
    refused(h.stemgnn_bn_act_backward_apply(8, 8, 16, 8, 16, 8, 16, 16, 16, 16, 0, 0.0, None, 16, 16, 4, 16, 8, None),
            "batch_rows")
    # K9 augmentations
    refused(h.stemgnn_dropout_adj_workspace(2 ** 31, C.byref(need)), "2^31")
    refused(h.stemgnn_dropout_adj_count(16, 10, None, 16, 1.5, 0, 16, 16, 1 << 20, None), "between 0 and 1")
    refused(h.stemgnn_dropout_adj_count(16, 10, None, None, 0.5, 0, 16, 16, 1 << 20, None), "rand or seed")
    refused(h.stemgnn_mask_fill(4, 6, 16, 6, None, None, 0.0, None, 0.0, 16, 6, None, None), "multiples of 4")
    # K10 sampler: fan-outs
    fan = (C.c_int32 * 2)(10, 10)
    nc, ec = C.c_int64(0), C.c_int64(0)
    assert h.stemgnn_neighbor_sample_bounds(1024, 1000, 5000, 2, fan, C.byref(nc), C.byref(ec), C.byref(need)) == 0
    assert nc.value >= 1024 and ec.value >= 1024 * 10 and need.value > 0
    refused(h.stemgnn_neighbor_sample_bounds(1024, 1000, 5000, 17, fan, C.byref(nc), C.byref(ec), C.byref(need)),
            "num_hops")
    refused(h.stemgnn_neighbor_sample_bounds(-1, 1000, 5000, 2, fan, C.byref(nc), C.byref(ec), C.byref(need)),
            "negative")
    # K12 negative sampling
    refused(h.stemgnn_negative_sampling_workspace(-1, 10, C.byref(need)), "negative size")
    refused(h.stemgnn_negative_sampling(16, 10, 1, 10, 10, 16, 16, 16, 16, 1 << 20, None), "two nodes")
    # K1 aggregation
    refused(h.stemgnn_aggregate(None, None), "NULL")
    g = _lib.AggregateArgs(n_rows=4, n_gather_rows=4, num_edges=3, feat=8, rowptr=16, col=16, gathered=16, out=16,
                           edge_feat=16, edge_weight=16, perm=16)
    refused(h.stemgnn_aggregate(C.byref(g), None), "cannot be combined")
    g = _lib.AggregateArgs(n_rows=4, n_gather_rows=4, num_edges=3, feat=8, rowptr=16, col=16, gathered=16, out=16,
                           backward=1, act=1)
    refused(h.stemgnn_aggregate(C.byref(g), None), "row_x")
    g = _lib.AggregateArgs(n_rows=4, n_gather_rows=4, num_edges=3, feat=8, rowptr=16, col=16, gathered=16, out=16,
                           out_ld=4)
    refused(h.stemgnn_aggregate(C.byref(g), None), "out_ld")


def test_model_shim_serves_the_reference_import_names(tmp_path):
    """`shim/` in front of a reference checkout: `model.encoder` / `model.vq` are ours, the rest of the
    reference's `model/` package (pt_model.py, ft_model.py) is still found in the checkout."""
    import subprocess
    import sys
    ref = tmp_path / "checkout" / "model"
    ref.mkdir(parents=True)
    (ref / "__init__.py").write_text("")
    (ref / "encoder.py").write_text("raise ImportError('the reference encoder must be shadowed')\n")
    (ref / "pt_model.py").write_text("from model.encoder import Encoder\nfrom model.vq import VectorQuantize\nMARK = 41\n")
    code = (
        "import sys\n"
        f"sys.path[:0] = [{os.path.join(ROOT, 'shim')!r}, {ROOT!r}, {str(tmp_path / 'checkout')!r}]\n"
        "import stemgnn_b200 as S\n"
        "from model.encoder import Encoder, MySAGEConv, MixtureSageLayer, InnerProductDecoder\n"
        "from model.vq import VectorQuantize, EuclideanCodebook, CosineSimCodebook\n"
        "import model.pt_model as pt\n"
        "assert Encoder is S.Encoder and VectorQuantize is S.VectorQuantize and MySAGEConv is S.MySAGEConv\n"
        "assert pt.MARK == 41 and pt.Encoder is S.Encoder and pt.VectorQuantize is S.VectorQuantize\n"
        "print('shim ok')\n")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "shim ok" in r.stdout, r.stderr[-1500:]


def test_header_is_plain_c_and_links_from_c(tmp_path):
    """include/stemgnn_b200.h compiled as C99 (-pedantic, warnings are errors) by a plain-C consumer that
    links the shared library and makes host-only calls: the boundary has no C++ or torch types in it."""
    import shutil
    import subprocess
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    _lib.lib()                                      # the library exists
    exe = str(tmp_path / "abi_probe")
    libdir = os.path.dirname(_lib.SO_PATH)
    cc = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"),
                         os.path.join(ROOT, "examples", "abi_probe.c"), "-o", exe, "-L", libdir, "-lstemgnn_b200",
                         f"-Wl,-rpath,{libdir}"], capture_output=True, text=True, timeout=120)
    assert cc.returncode == 0, cc.stderr[-2000:]
    run = subprocess.run([exe], capture_output=True, text=True, timeout=60)
    assert run.returncode == 0, run.stdout + run.stderr
    assert f"abi {_lib.ABI_VERSION}" in run.stdout and "launches so far 0" in run.stdout


def test_library_depends_on_the_cuda_runtime_only():
    """The boundary library links libcudart and the C/C++ runtimes -- no libtorch, no libc10, no Python:
    PyTorch sits above the C-ABI (device memory, streams), not inside it."""
    import shutil
    import subprocess
    if shutil.which("ldd") is None:
        pytest.skip("no ldd")
    _lib.lib()
    out = subprocess.run(["ldd", _lib.SO_PATH], capture_output=True, text=True, timeout=60).stdout
    names = [l.split()[0] for l in out.splitlines() if l.strip()]
    assert any(n.startswith("libcudart") for n in names), out
    for bad in ("libtorch", "libc10", "libpython", "libnccl", "libcublas", "libcudnn"):
        assert not any(n.startswith(bad) for n in names), (bad, out)


def test_product_code_never_reaches_the_oracle():
    """oracle/ is test infrastructure: nothing under stem-gnn_b200/ or shim/ mentions it, bench.py imports
    it only inside the CPU-baseline step builder, and __graft_entry__ only inside smoke()."""
    import ast
    import glob
    for f in glob.glob(os.path.join(ROOT, "stem-gnn_b200", "*.py")) + glob.glob(os.path.join(ROOT, "shim", "model", "*.py")) \
            + glob.glob(os.path.join(ROOT, "stemgnn_b200", "*.py")):
        assert "oracle" not in open(f).read(), f

    def oracle_importers(path):
        tree = ast.parse(open(path).read())
        found = set()
        for fn in ast.walk(tree):
            if isinstance(fn, (ast.FunctionDef, ast.Module)):
                for node in (fn.body if isinstance(fn, ast.Module) else ast.walk(fn)):
                    mod = getattr(node, "module", None) if isinstance(node, ast.ImportFrom) else None
                    names = [a.name for a in node.names] if isinstance(node, ast.Import) else []
                    if (mod and mod.split(".")[0] == "oracle") or any(n.split(".")[0] == "oracle" for n in names):
                        found.add(fn.name if isinstance(fn, ast.FunctionDef) else "<module>")
        return found

    assert oracle_importers(os.path.join(ROOT, "bench.py")) == {"cpu_baseline_step_fn"}
    assert oracle_importers(os.path.join(ROOT, "__graft_entry__.py")) == {"smoke"}
