"""CPU: bench.py's contract where it can be checked without a GPU -- the reference arm's JSON line, the
loud failure of the product arm without a CUDA device, and the synthetic graph generators."""
import importlib.util
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench_module():
    spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(ROOT, "bench.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _run(*args, timeout=240):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], cwd=ROOT, timeout=timeout,
                          capture_output=True, text=True)


@pytest.mark.timeout(300)
def test_reference_arm_prints_one_contract_line():
    r = _run("--impl", "reference", "--workload", "cora", "--steps", "2", "--warmup", "1")
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
                "scaling", "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e", "impl"):
        assert key in d, key
    assert d["impl"] == "reference" and d["steps"] == 2 and d["warmup"] == 1 and d["value"] > 0
    assert d["unit"] == "edges/s" and d["higher_is_better"] is True and d["vs_baseline"] is None
    assert "workload" in d["config"] and "model" not in d["config"]
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == dict(value=d["value"], unit=d["unit"], h2d_bytes_per_step=0, d2h_bytes_per_step=0)


@pytest.mark.skipif(torch.cuda.is_available(), reason="needs a box without a GPU")
def test_product_arm_refuses_to_run_without_a_gpu():
    r = _run("--steps", "1", "--warmup", "0", "--no-cpu", timeout=120)
    assert r.returncode != 0 and "no CPU path" in (r.stderr + r.stdout)
    assert not any(l.startswith("{") for l in r.stdout.splitlines())


def test_zipf_destinations_are_capped_seeded_and_skewed():
    b = _bench_module()
    w = dict(b.WORKLOADS["arxiv"], degree="zipf")
    x, ei = b.synth_graph(w, 1)
    x2, ei2 = b.synth_graph(w, 1)
    assert torch.equal(ei, ei2) and torch.equal(x, x2)                       # seeded
    assert ei.dtype == torch.int64 and tuple(ei.shape) == (2, w["e"])
    assert int(ei.min()) >= 0 and int(ei.max()) < w["n"]
    deg = torch.bincount(ei[1], minlength=w["n"])
    assert 0.9 * b.ZIPF_CAP < int(deg.max()) < 1.1 * b.ZIPF_CAP              # the cap holds in expectation
    hubs = deg >= 256
    assert int(hubs.sum()) > 100 and int(deg[hubs].sum()) > 0.3 * w["e"]     # a third of the edges sit on hubs
    top = torch.topk(deg, 50).indices
    assert int(top.max()) - int(top.min()) > w["n"] // 2                     # hubs scattered over the node ids
    src_deg = torch.bincount(ei[0], minlength=w["n"])
    assert int(src_deg.max()) < 64                                           # sources stay uniform
    _, eu = b.synth_graph(dict(b.WORKLOADS["arxiv"]), 1)
    assert int(torch.bincount(eu[1], minlength=w["n"]).max()) < 64           # the headline configuration is uniform
