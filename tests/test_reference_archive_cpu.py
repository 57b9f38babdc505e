"""CPU: the archive of the reference's own model/ package (oracle/make_ref.py -> oracle/_ref/, git-ignored, shipped
to the GPU box) holds the UNMODIFIED files, loads under private names without touching the import shim's `model`
package, and is what bench.py's CPU arm runs."""
import hashlib
import os
import subprocess
import sys
import zipfile

import pytest
import torch
import torch.nn as nn

from oracle import make_ref, ref_loader

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HAVE_CHECKOUT = os.path.isdir(make_ref.REF_MODEL)


@pytest.mark.skipif(not HAVE_CHECKOUT, reason="/root/reference is only in the build container")
def test_archive_holds_the_reference_files_byte_for_byte():
    path = make_ref.make()
    assert path and os.path.exists(path)
    with zipfile.ZipFile(path) as z:
        for f in make_ref.FILES:
            want = open(os.path.join(make_ref.REF_MODEL, f), "rb").read()
            assert hashlib.sha256(z.read("model/" + f)).digest() == hashlib.sha256(want).digest(), f
        assert "unmodified" in z.read("MANIFEST.txt").decode()
    ignored = subprocess.run(["git", "check-ignore", "-q", path], cwd=ROOT).returncode == 0
    assert ignored, "the archive must stay out of the repository's history (.gitignore: oracle/_ref/)"


@pytest.mark.skipif(ref_loader.source()[0] is None, reason="neither the checkout nor the archive is present")
def test_loader_serves_the_reference_modules_and_leaves_the_shim_alone(monkeypatch):
    before = {k for k in sys.modules if k == "model" or k.startswith("model.")}
    if os.path.exists(ref_loader.ARCHIVE):
        monkeypatch.setattr(ref_loader, "CHECKOUT", "/nonexistent")      # what the GPU box sees
        ref_loader._cache.clear()
    ns = ref_loader.load()
    assert ns.kind == ("archive" if os.path.exists(ref_loader.ARCHIVE) else "checkout")
    assert {k for k in sys.modules if k == "model" or k.startswith("model.")} == before
    torch.manual_seed(0)
    vq = ns.vq.VectorQuantize(dim=16, codebook_size=8, decay=0.8)
    q, ind, loss, _ = vq(torch.randn(50, 16))
    assert ind.dtype == torch.int64 and tuple(loss.shape) == (1,)
    enc = ns.encoder.Encoder(12, 16, nn.ReLU, 2, backbone="sage", normalize="batch", dropout=0.0)
    z = enc(torch.randn(30, 12), torch.randint(0, 30, (2, 100)))
    assert tuple(z.shape) == (30, 16)
    assert hasattr(ns.pt_model, "PretrainModel") and hasattr(ns.ft_model, "TaskModel")
    ref_loader._cache.clear()


@pytest.mark.skipif(ref_loader.source()[0] is None, reason="neither the checkout nor the archive is present")
def test_bench_cpu_arm_is_the_reference_itself():
    import json
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "cora",
                        "--steps", "1", "--warmup", "0"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-400:]
    d = json.loads(r.stdout.strip().splitlines()[-1])
    assert d["cpu_baseline"]["kind"] == "reference" and "model/vq.py" in d["cpu_baseline"]["sample"]
