#!/usr/bin/env python
"""bench.py -- the driver-facing benchmark of the STEM-GNN encoder/VQ hot path on B200.

One "step" = one pretrain-style pass of the hot path over one synthetic graph of the
BASELINE.json configs[1] shape (ogbn-arxiv-shaped: 169,343 nodes, 1,166,243 edges, 128 feats;
2-layer SAGE encoder 128->256->256 with BatchNorm + dropout 0.15; VectorQuantize 512 x 256,
Euclidean, EMA decay 0.8): K0 CSR/CSC build of the fresh edge_index, encoder forward,
VQ forward (argmin, gather, commitment loss, EMA update), backward, AdamW step.

    python bench.py --gpus N --steps K --warmup W            # our arm (CUDA kernels, C-ABI)
    python bench.py --impl reference --steps K --warmup W     # the CPU restatement, host cores

* `value`  : edges/s = layers * E * n_gpus / step time, inputs resident in HBM, L2 flushed
             between timed steps, CUDA events, max over ranks.
* `e2e`    : same metric through the public module API with HOST (pinned) inputs: H2D of x and
             edge_index and D2H of the loss inside the timed region.
* `roofline`: the dominant kernel family of this library in the timed region, timed live with
             CUDA events around each C-ABI call (stemgnn_b200._lib.traced).
* N > 1    : one process per GPU (torchrun); every rank runs the step on its own graph (weak
             scaling, the reference's mini-batch replicas); EMA statistics (model/vq.py:494,499)
             and dense gradients are all-reduced over NCCL.
The oracle (oracle/) is executed only for `cpu_baseline` and `--impl reference`.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.nn as nn  # noqa: E402

WORKLOADS = {
    # name: N, E, F_in, hidden, layers, K, D
    "arxiv": dict(n=169343, e=1166243, f=128, hid=256, layers=2, codes=512,
                  desc="ogbn-arxiv-shaped synthetic (169,343 nodes, 1,166,243 edges, 128 feats) "
                       "2-layer SAGE encoder 128-256-256 + VQ 512x256 Euclid/EMA pretrain step"),
    "cora": dict(n=2708, e=10556, f=1433, hid=768, layers=2, codes=128,
                 desc="Cora-shaped synthetic (2,708 nodes, 10,556 edges, 1,433 feats)"),
    "products": dict(n=2449029, e=61859140, f=100, hid=256, layers=3, codes=512,
                     desc="ogbn-products-shaped synthetic (2,449,029 nodes, 61,859,140 edges, 100 feats)"),
}
METRIC = "edges/sec encoder fwd+bwd & VQ nodes/sec"
L2_FLUSH_BYTES = 256 << 20


def bind_to_gpu_numa(local):
    """Pin this rank's threads to the CPUs next to its GPU (NVML's affinity mask) BEFORE the pinned host buffers are
    allocated: first touch puts them on that NUMA node, so the per-step host -> device copies of 8 ranks (105 MB each)
    do not all cross the socket interconnect.  Best effort: any failure leaves the affinity alone."""
    if os.environ.get("STEMGNN_BENCH_NUMA", "1") != "1":
        return None
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:                                                  # noqa: BLE001
        pass
    return None


def init_nccl(dist, dev):
    """One process per GPU over NCCL.  The collectives' stream is created at HIGH priority: in the overlapped halo
    exchange (PartitionedGraph halo_mode='blocks') the send/recv kernels of the next block must get SMs while an
    aggregation launch of thousands of CTAs is running -- measured on 4 B200s (profiles/r02_partition_prio_4gpu.log):
    aggregation pair 5.88 -> 4.72 ms with it, against 5.67 ms for the blocking all-gather."""
    opts = dist.ProcessGroupNCCL.Options(is_high_priority_stream=True)
    dist.init_process_group("nccl", device_id=dev, pg_options=opts)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tc=d["bf16_tflops"], tc_sus=d.get("bf16_tflops_sustained"),
                    src="measured")
    return dict(hbm=6650.0, tc=1590.0, tc_sus=1400.0, src="fallback")


ZIPF_CAP = 17000        # no destination above ~17,000 in-edges (ogbn-products' largest: 17,481)


class ZipfNodes:
    """SURVEY.md 8d's power-law variant: node ids drawn with p(rank k) ~ 1/k (Zipf, alpha = 1), no rank
    above `ZIPF_CAP` edges in expectation, ranks scattered over the ids by a seeded permutation (so the
    hubs are not the first rows a launch starts).  Inverse-CDF sampling on the host, float64."""

    def __init__(self, n, e, g):
        base = 1.0 / torch.arange(1, n + 1, dtype=torch.float64)
        wgt = base
        for _ in range(30):                      # fixed point of "largest probability = ZIPF_CAP / e"
            wgt = torch.minimum(base, wgt.sum() * (ZIPF_CAP / e))
        self.cdf = torch.cumsum(wgt / wgt.sum(), 0)
        self.cdf[-1] = 1.0
        self.ids = torch.randperm(n, generator=g)

    def draw(self, count, g):
        u = torch.rand(count, dtype=torch.float64, generator=g)
        return self.ids[torch.searchsorted(self.cdf, u, right=True).clamp_(max=self.ids.numel() - 1)]


def synth_graph(w, seed, device="cpu"):
    g = torch.Generator(device="cpu").manual_seed(seed)
    ei = torch.randint(0, w["n"], (2, w["e"]), generator=g)
    if w.get("degree") == "zipf":
        ei[1] = ZipfNodes(w["n"], w["e"], g).draw(w["e"], g)
    x = torch.randn(w["n"], w["f"], generator=g)
    return x.to(device), ei.to(device)


# --------------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi polled every 20 ms beside the run; summaries are taken over wall-clock windows (the timed
    region is tens of milliseconds at the driver's --steps 20, so the sampler is started before the
    sustained warm-up that precedes it and the timed window is reported next to the warm-up's)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def summary(self, t0=None, t1=None):
        sm, pw, mx, reasons = [], [], None, set()
        for ts, ln in self.lines:
            if (t0 is not None and ts < t0) or (t1 is not None and ts > t1 + 0.03):
                continue
            c = [t.strip() for t in ln.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx = float(c[2]); pw.append(float(c[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                                "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(pw) if pw else None}

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.06)
        self.proc.terminate()
        out = self.summary(t0, t1)
        if out["samples"] == 0:            # a window shorter than the polling period: nearest samples either side
            out = self.summary(None if t0 is None else t0 - 0.05, None if t1 is None else t1 + 0.05)
        return out


# --------------------------------------------------------------------------------- our arm
def build_models(w, device, S):
    torch.manual_seed(0)
    enc = S.Encoder(w["f"], w["hid"], nn.ReLU, w["layers"], backbone="sage", normalize="batch",
                    dropout=0.15).to(device).train()
    vq = S.VectorQuantize(dim=w["hid"], codebook_size=w["codes"], decay=0.8,
                          commitment_weight=1.0).to(device).train()
    # graph-capturable, one fused multi-tensor kernel per step instead of ~16 foreach launches: same math
    opt = torch.optim.AdamW(enc.parameters(), lr=1e-4, capturable=True, fused=True)
    return enc, vq, opt


class _SumSquares(torch.autograd.Function):
    """sum(q^2) of the stand-in task loss with a ONE-pass backward.  Written as `q.square().mean()` autograd runs
    mean -> expand -> div -> pow backward: five passes over the 173 MB matrix, one of them the strided
    `elementwise_kernel` of the expanded gradient (90 us) -- 0.29 ms of a 3.5 ms step spent on the bench's own
    loss expression (ncu launch list, round 2).  Same value, same gradient 2 * q * g."""

    @staticmethod
    def forward(ctx, q):
        ctx.save_for_backward(q)
        return torch.linalg.vector_norm(q).square()

    @staticmethod
    def backward(ctx, g):
        (q,) = ctx.saved_tensors
        return torch._foreach_mul([q], g * 2.0)[0]       # tensor x 0-dim tensor through the vectorised list kernel


def mean_square(q, denom=None):
    return _SumSquares.apply(q) / (q.numel() if denom is None else denom)


def make_step(enc, vq, opt, S, world):
    import torch.distributed as dist
    params = [p for p in enc.parameters()]

    def step(x, ei):
        S.clear_structure_cache()          # pretraining sees a fresh edge_index every batch
        opt.zero_grad(set_to_none=True)
        z = enc(x, ei)
        q, ind, vq_loss, _ = vq(z)
        loss = mean_square(q) + vq_loss.sum()
        loss.backward()
        if world > 1:
            flat = torch.cat([p.grad.reshape(-1) for p in params])
            dist.all_reduce(flat)
            flat /= world
            o = 0
            for p in params:
                p.grad.copy_(flat[o:o + p.numel()].view_as(p)); o += p.numel()
        opt.step()
        return loss
    return step


def summarize_trace(trace, steps, pk, n_feat_bytes_l2):
    fam = {}
    for name, nbytes, flops, launches, e0, e1 in trace:
        ms = e0.elapsed_time(e1)
        key = "aggregate" if name.startswith("aggregate") else name
        d = fam.setdefault(key, dict(ms=0.0, calls=0, bytes_min=0, bytes_gather=0, bytes_alg=0, flops=0, launches=0))
        d["ms"] += ms; d["calls"] += 1; d["flops"] += flops; d["launches"] += launches
        if isinstance(nbytes, tuple):
            d["bytes_min"] += nbytes[0]; d["bytes_gather"] += nbytes[1]; d["bytes_alg"] += nbytes[2]
        else:
            d["bytes_min"] += nbytes; d["bytes_gather"] += nbytes; d["bytes_alg"] += nbytes
    out = {}
    for k, d in fam.items():
        sec = d["ms"] / 1e3
        o = dict(ms_per_step=d["ms"] / steps, calls_per_step=d["calls"] / steps,
                 launches_per_step=d["launches"] / steps)
        if sec > 0:
            o["gbps_alg"] = d["bytes_alg"] / sec / 1e9
            o["alg_bytes_per_call"] = d["bytes_alg"] / d["calls"]
            o["bytes_min_per_call"] = d["bytes_min"] / d["calls"]
            o["bytes_gather_per_call"] = d["bytes_gather"] / d["calls"]
            o["gbps_alg_min"] = d["bytes_min"] / sec / 1e9
            o["gbps_alg_gather"] = d["bytes_gather"] / sec / 1e9
            if d["flops"]:
                o["tflops"] = d["flops"] / sec / 1e12
        out[k] = o
    return out


def roofline_of(kern, name, pk, x_fits_l2=None, codes=None):
    k = kern[name]
    if name == "vq_assign" and k.get("tflops") is not None and codes is not None and codes <= 512:
        # with fp32 inputs the distance GEMM's intensity is 2K/4 flop per byte, the ridge
        # (1,664 TF / 6.55 TB/s = 254 flop/B) is at K ~ 508: K <= 512 is reported against HBM
        a = k["gbps_alg"]
        return dict(kernel=name, bound="hbm", achieved=a, peak=pk["hbm"], unit="GB/s", frac=a / pk["hbm"],
                    traffic=None, peak_source=pk["src"], alg_bytes_per_launch=k["alg_bytes_per_call"],
                    tflops=k["tflops"],
                    bytes_model="N*D*4 (x) + N*D*4 (quantize) + N*8 (indices) + K*D*4 per call; K <= 512 with "
                                "fp32 inputs sits left of the tensor/HBM ridge (SURVEY.md 8d)")
    if name == "vq_assign" and k.get("tflops") is not None:
        # K >= ~508 with fp32 inputs sits right of the ridge: tensor-bound
        return dict(kernel=name, bound="tensor", achieved=k["tflops"], peak=pk["tc_sus"] or pk["tc"],
                    unit="TFLOP/s", frac=k["tflops"] / (pk["tc_sus"] or pk["tc"]), traffic=None,
                    peak_source=pk["src"] + " (sustained: kernel timed inside a long step)")
    if name in ("linear", "linear_wgrad") and k.get("tflops") is not None:
        # fp32-accurate GEMM as 3xTF32: the tensor pipe executes 3x the algorithmic 2MNK flops at
        # the tf32 rate (half the measured dense bf16 rate), so frac tops out at 1/3
        peak = (pk["tc_sus"] or pk["tc"]) / 2
        return dict(kernel=name, bound="tensor", achieved=k["tflops"], peak=peak, unit="TFLOP/s",
                    frac=k["tflops"] / peak, executed_tflops=3 * k["tflops"], traffic=None,
                    peak_source=pk["src"] + " bf16 dense / 2 = tf32 rate (sustained)",
                    flops_model="algorithmic 2*M*N*K per call; 3 tf32 MMAs per product are executed")
    a = k["gbps_alg"]
    r = dict(kernel=name, bound="hbm", achieved=a, peak=pk["hbm"], unit="GB/s", frac=a / pk["hbm"],
             traffic=None, peak_source=pk["src"], alg_bytes_per_launch=k["alg_bytes_per_call"],
             achieved_basis="algorithmic bytes per call / live CUDA-event time of the call")
    if name == "aggregate":
        # two byte models per call: B_min = every row once (its rule while the gathered matrix fits
        # the 126 MB L2: configs C1, C2), B_gather = one source row per edge.  Neither is what the DRAM moves once
        # part of the gathers hit L2, so when an ncu capture of this configuration exists (`traffic`, attached by
        # the caller) `achieved` / `frac` are restated from the MEASURED DRAM bytes and the models stay beside them.
        r.update(frac_model_min=k["gbps_alg_min"] / pk["hbm"], frac_model_gather=k["gbps_alg_gather"] / pk["hbm"],
                 alg_bytes_per_launch=k["bytes_min_per_call"], alg_bytes_gather_per_launch=k["bytes_gather_per_call"],
                 bytes_model="B_min per call: (gathered rows + produced rows [x2 with the backward relu mask]) * F * 4 "
                             "+ E * 4 + (N + 1) * 4 + N * 4; B_gather: E * F * 4 in place of the gathered rows")
    return r


def attach_traffic(r, k, t, pk):
    """`traffic` = dram__bytes_read.sum + dram__bytes_write.sum per call from the ncu --set full capture of this
    configuration (profiles/traffic.json, written by scripts/ncu_traffic.py from the raw CSV beside it).  For the
    HBM-bound families the line's achieved / frac become measured DRAM bytes over the live CUDA-event time."""
    r["traffic"] = t["bytes_per_launch"]
    r["traffic_source"] = t["source"]
    if r["bound"] == "hbm" and k["ms_per_step"] > 0:
        gbps = t["bytes_per_launch"] * k["calls_per_step"] / (k["ms_per_step"] / 1e3) / 1e9
        r.update(achieved=gbps, frac=gbps / pk["hbm"],
                 achieved_basis="measured DRAM bytes per call (ncu dram__bytes_read + write, `traffic`) / live "
                                "CUDA-event time of the call")
        if r.get("alg_bytes_per_launch"):
            r["traffic_over_alg"] = t["bytes_per_launch"] / r["alg_bytes_per_launch"]
    return r


def cpu_baseline_step_fn(w, threads):
    """The same step on the host cores (fp32, all host threads).  Returns (step, kind, what).

    kind "reference": the reference's OWN modules -- model/vq.py VectorQuantize and model/encoder.py Encoder,
    unmodified, imported from /root/reference or from the archive oracle/make_ref.py packed into the git-ignored
    oracle/_ref/ (the GPU box has no /root/reference).  encoder.py runs on oracle/pyg_shim when torch_geometric is
    not importable (it is not installable here).  kind "port": oracle/'s restatement, when neither is present."""
    torch.set_num_threads(threads)
    torch.manual_seed(0)
    try:
        from oracle import ref_loader
        ref = ref_loader.load(("vq", "encoder"))
    except (FileNotFoundError, ImportError):
        ref = None
    if ref is not None:
        enc = ref.encoder.Encoder(w["f"], w["hid"], nn.ReLU, w["layers"], backbone="sage", normalize="batch",
                                  dropout=0.15).train()
        vq = ref.vq.VectorQuantize(dim=w["hid"], codebook_size=w["codes"], decay=0.8, commitment_weight=1.0).train()
        opt = torch.optim.AdamW(enc.parameters(), lr=1e-4)

        def step(x, ei):
            opt.zero_grad(set_to_none=True)
            z = enc(x, ei)
            q, ind, vq_loss, _ = vq(z)
            loss = mean_square(q) + vq_loss.sum()
            loss.backward()
            opt.step()
            return float(loss.detach())
        return step, "reference", (f"the reference's own model/vq.py + model/encoder.py ({ref.kind}), "
                                   f"message passing underneath from {ref.pyg}")
    from oracle.encoder_ref import EncoderRef
    from oracle.vq_ref import vq_forward_ref
    enc = EncoderRef(w["f"], w["hid"], nn.ReLU, w["layers"], backbone="sage", normalize="batch",
                     dropout=0.15).train()
    opt = torch.optim.AdamW(enc.parameters(), lr=1e-4)
    K, D = w["codes"], w["hid"]
    embed = torch.empty(1, K, D)
    nn.init.kaiming_uniform_(embed)
    state = dict(embed=embed, embed_avg=embed.clone(), cluster_size=torch.zeros(1, K))

    def step(x, ei):
        opt.zero_grad(set_to_none=True)
        z = enc(x, ei)
        q, ind, vq_loss, _ = vq_forward_ref(z, state, training=True, ema_update=True, decay=0.8,
                                            commitment_weight=1.0)
        loss = mean_square(q) + vq_loss.sum()
        loss.backward()
        opt.step()
        return float(loss.detach())
    return step, "port", "oracle/ restatement of the reference step (oracle/_ref absent)"


def cpu_sample(w, frac):
    """Bounded sample of the workload for the CPU arm: the same shape scaled by `frac`."""
    if frac >= 1.0:
        return dict(w), "full workload"
    s = dict(w)
    s["n"], s["e"] = max(int(w["n"] * frac), 1), max(int(w["e"] * frac), 1)
    return s, f"{frac:g} of the nodes and edges (same widths), scaled by edges"


def run_reference(args, w):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    ws, sample = cpu_sample(w, args.cpu_frac)
    step, kind, what = cpu_baseline_step_fn(ws, threads)
    x, ei = synth_graph(ws, 1)
    for _ in range(args.warmup):
        step(x, ei)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step(x, ei)
    dt = time.perf_counter() - t0
    val = ws["layers"] * ws["e"] * args.steps / dt
    nodes = ws["n"] * args.steps / dt
    line = dict(metric=METRIC, value=val, unit="edges/s", impl="reference", n_gpus=args.gpus,
                steps=args.steps, warmup=args.warmup, ms_per_step=dt / args.steps * 1e3,
                higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32",
                data="synthetic", vq_nodes_per_s=nodes,
                config=dict(workload=w["desc"], layers=w["layers"], nodes=w["n"], edges=w["e"], sample=sample),
                cpu_baseline=dict(value=val, unit="edges/s", cores=threads, kind=kind,
                                  sample=sample + "; " + what),
                e2e=dict(value=val, unit="edges/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0),
                gpu_launches=0)
    print(json.dumps(line), flush=True)


def run_ours(args, w):
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; stemgnn_b200 has no CPU path "
                         "(use --impl reference for the CPU restatement)")
    import __graft_entry__ as G
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus and world > 1:
        args.gpus = world
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa_cpus = bind_to_gpu_numa(local) if world > 1 else None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        init_nccl(dist, dev)
    if rank == 0:
        G.build()
    if world > 1:
        dist.barrier()
    import stemgnn_b200 as S
    from stemgnn_b200 import _lib
    _lib.lib()

    enc, vq, opt = build_models(w, dev, S)
    step = make_step(enc, vq, opt, S, world)
    x_h, ei_h = synth_graph(w, 1 + rank)
    x_h, ei_h = x_h.pin_memory(), ei_h.pin_memory()
    x_d, ei_d = x_h.to(dev), ei_h.to(dev)
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=dev)
    pk = peaks()

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---------------- one CUDA graph of the whole step (forward, backward, optimizer)
    # ~170 kernels of 5-200 us each: launched one by one the step is bound by the host's launch
    # rate.  The step reads static input buffers (x_in, ei_in) and writes loss_out; replaying the
    # graph does exactly the work of the eager step (CSR rebuild included).
    x_in, ei_in = x_d.clone(), ei_d.clone()
    for _ in range(max(args.warmup, 3)):
        step(x_in, ei_in)
    sync_all()
    graph, loss_out, graph_err = None, None, None
    if not args.no_graph and (world == 1 or os.environ.get("STEMGNN_BENCH_GRAPH_MULTI", "1") == "1"):
        try:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, capture_error_mode="thread_local"):
                loss_out = step(x_in, ei_in)
            graph = g
        except Exception as e:                                     # fall back to eager launches
            graph_err = f"{type(e).__name__}: {e}".splitlines()[0][:200]
            torch.cuda.synchronize()

    def run_step():
        if graph is not None:
            graph.replay()
            return loss_out
        return step(x_in, ei_in)

    for _ in range(3):
        run_step()
    sync_all()

    # ---------------- resident-input timing (value)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    # sustained warm-up: K timed steps are tens of milliseconds, too short for clocks, power and the 20 ms
    # clock sampler to mean anything on their own, so the same step runs untimed for --warm-seconds first and
    # the timed steps follow without a gap (board at its loaded clocks / power state)
    t_warm0 = time.time()
    warm_steps = 0
    while time.time() - t_warm0 < args.warm_seconds:
        for _ in range(10):
            flush.fill_(1)
            run_step()
        torch.cuda.synchronize()
        warm_steps += 10
    sync_all()
    t_timed0 = time.time()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(args.steps)]
    for a, b in ev:
        flush.fill_(1)                      # L2 flush between timed steps (256 MB > 126 MB L2)
        a.record()
        run_step()
        b.record()
    sync_all()
    t_timed1 = time.time()
    ms = sum(a.elapsed_time(b) for a, b in ev)

    # ---------------- end-to-end timing (host buffers in, loss out)
    # Inputs come from pinned host memory every step.  Double-buffered: the copy of step i+1 runs
    # on a side stream while step i computes; step i waits for its own copy only.
    loss_h = torch.empty(1, dtype=torch.float32).pin_memory()
    copy_stream = torch.cuda.Stream()
    stage = [(torch.empty_like(x_d), torch.empty_like(ei_d)) for _ in range(2)]
    ready = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]

    def issue_copy(i):
        b = i & 1
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[b])
            stage[b][0].copy_(x_h, non_blocking=True)
            stage[b][1].copy_(ei_h, non_blocking=True)
            ready[b].record(copy_stream)

    def e2e_step(i, last):
        b = i & 1
        cur = torch.cuda.current_stream()
        cur.wait_event(ready[b])
        x_in.copy_(stage[b][0], non_blocking=True)        # device-to-device into the graph's inputs
        ei_in.copy_(stage[b][1], non_blocking=True)
        consumed[b].record(cur)
        if not last:
            issue_copy(i + 1)
        loss = run_step()
        loss_h.copy_(loss.detach().reshape(1), non_blocking=True)

    for b in range(2):
        consumed[b].record(torch.cuda.current_stream())
    issue_copy(0)
    e2e_step(0, False); e2e_step(1, True)
    sync_all()
    ev2 = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    for i in range(args.steps):
        flush.fill_(1)
        if i == 0:
            ev2[0].record()
            issue_copy(0)                   # the first step's copy is inside the timed region, exposed
        e2e_step(i, i == args.steps - 1)
        ev2[i + 1].record()
    sync_all()
    # the L2 flushes sit inside this region (they overlap the prefetch copies); per-step times come
    # from consecutive events
    ms_e2e = ev2[0].elapsed_time(ev2[-1])
    if rank == 0:
        clocks = sampler.stop(t_timed0, t_timed1)
        clocks["window"] = "the K timed steps"
        clocks["warmup"] = dict(sampler.summary(t_warm0, t_timed0), steps=warm_steps,
                                seconds=round(t_timed0 - t_warm0, 2))
    else:
        clocks = None

    # ---------------- per-kernel device times (eager, traced): roofline of each kernel family
    _lib.TRACE = []
    l0 = _lib.launch_count()
    for _ in range(args.steps):
        flush.fill_(1)
        step(x_in, ei_in)
    sync_all()
    launches = (_lib.launch_count() - l0) / args.steps
    trace, _lib.TRACE = _lib.TRACE, None
    if args.dump_sequence and rank == 0:
        # the C-ABI calls of ONE step in launch order with the kernels each launched: scripts/ncu_traffic.py
        # aligns an ncu launch list of the same step with it
        per = len(trace) // args.steps
        json.dump([[r[0], r[3], r[4].elapsed_time(r[5])] for r in trace[-per:]], open(args.dump_sequence, "w"))

    t = torch.tensor([ms, ms_e2e], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e = float(t[0]), float(t[1])
    # ---------------- the north_star multi-GPU mode, in the same process: BASELINE configs[3] (products-shaped,
    # 3-layer) on a 1-D destination partition over these `world` ranks, strong scaling.  The top-level line stays
    # the configs[1] step (replicas for N > 1); this record rides along on every line, N = 1 included.
    part = None
    if not args.no_partition:
        del flush
        torch.cuda.empty_cache()
        try:
            part = measure_partition(dict(WORKLOADS["products"]), dev, rank, world, args.partition_steps, 3)
            if part is not None:
                part.pop("_extra", None)
        except Exception as exc:                                      # noqa: BLE001 -- the headline line must still print
            part = dict(error=f"{type(exc).__name__}: {exc}".splitlines()[0][:300])

    def finish():
        if world > 1:
            torch.cuda.synchronize()
            sys.stdout.flush(); sys.stderr.flush()
            # Tearing NCCL down while a captured graph still holds its collectives hung in round 1 (torch 2.11 /
            # NCCL 2.28).  The graph is released first and the teardown runs under a 15 s timer: if it still does not
            # return, the process leaves with the result line already printed instead of hanging the launcher.
            def give_up():
                sys.stderr.write("bench.py: NCCL teardown did not return within 15 s; leaving without it\n")
                sys.stderr.flush()
                os._exit(0)
            timer = threading.Timer(15.0, give_up)
            timer.daemon = True
            timer.start()
            try:
                if graph is not None:
                    graph.reset()
                dist.destroy_process_group()
            finally:
                timer.cancel()

    if rank != 0:
        finish()
        return
    units = w["layers"] * w["e"] * world
    value = units * args.steps / (ms / 1e3)
    e2e_v = units * args.steps / (ms_e2e / 1e3)
    kern = summarize_trace(trace, args.steps, pk, None)
    top = max(kern, key=lambda k: kern[k]["ms_per_step"])
    traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
    traffic = json.load(open(traffic_file)).get(args.workload, {}) if os.path.exists(traffic_file) else {}
    if w.get("degree") == "zipf":
        traffic = {}                      # the ncu captures are of the uniform headline configuration
    rooflines = {}
    for name in kern:
        if "gbps_alg" not in kern[name]:
            continue
        r = roofline_of(kern, name, pk, codes=w["codes"])
        if name in traffic:
            attach_traffic(r, kern[name], traffic[name], pk)
        rooflines[name] = r
    roof = rooflines[top]
    line = dict(metric=METRIC, value=value, unit="edges/s", n_gpus=world, steps=args.steps,
                warmup=max(args.warmup, 3), ms_per_step=ms / args.steps, higher_is_better=True,
                scaling="weak", vs_baseline=None, dtype="f32", data="synthetic",
                vq_nodes_per_s=w["n"] * world * args.steps / (ms / 1e3),
                config=dict(workload=w["desc"], layers=w["layers"], nodes=w["n"], edges=w["e"],
                            l2="flushed between timed steps (256 MB write)",
                            cuda_graph=graph is not None, cuda_graph_error=graph_err,
                            e2e_pipeline="pinned host -> device copy of step i+1 on a side stream during "
                                         "step i (double-buffered); every step's copy and loss read-back "
                                         "are inside the timed region",
                            kernel_times="CUDA events around every C-ABI call in an eager pass of the same "
                                         "step right after the timed region",
                            parallelism=f"replicas x{world}: EMA statistics + dense grads all-reduced"
                            if world > 1 else "single GPU",
                            host_affinity=f"rank bound to the {numa_cpus} CPUs NVML lists next to its GPU"
                            if numa_cpus else "unchanged"),
                e2e=dict(value=e2e_v, unit="edges/s", ms_per_step=ms_e2e / args.steps,
                         h2d_bytes_per_step=x_h.numel() * 4 + ei_h.numel() * 8, d2h_bytes_per_step=4),
                gpu_launches=launches, roofline=roof, rooflines=rooflines, kernels=kern, clocks=clocks,
                partition=part)
    if world == 1 and not args.no_cpu:
        threads = os.cpu_count() or 1
        ws, sample = cpu_sample(w, args.cpu_frac)
        cstep, ckind, cwhat = cpu_baseline_step_fn(ws, threads)
        xc, eic = synth_graph(ws, 1)
        cstep(xc, eic)
        t0, n = time.perf_counter(), 0
        while n < 3 or (time.perf_counter() - t0 < 10 and n < 50):
            cstep(xc, eic); n += 1
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = dict(value=ws["layers"] * ws["e"] * n / dt, unit="edges/s",
                                    cores=threads, kind=ckind,
                                    sample=f"{n} steps of {sample}; {cwhat}; torch CPU fp32")
    print(json.dumps(line), flush=True)
    finish()


PARTITION_N1_FILE = os.path.join(os.environ.get("TMPDIR", "/tmp"), "stemgnn_b200_partition_n1.json")


def partition_modes(world):
    """(label, halo_mode, wire dtype, transport) measured by the partition leg.  One rank has no exchange."""
    if world == 1:
        return [("single GPU (no exchange)", "gather", None, "p2p")]
    modes = [("all-gather, fp32 wire (bit-identical to one GPU)", "gather", None, "p2p"),
             ("all-gather pipelined over 2 row chunks, fp32 wire", "chunks", None, "2"),
             ("all-gather pipelined over 4 row chunks, fp32 wire", "chunks", None, "4"),
             ("all-gather pipelined over 4 row chunks, bf16 wire", "chunks", "bf16", "4")]
    if os.environ.get("STEMGNN_HALO_BLOCKS", "0") == "1":     # owner-block modes: P-fold cut of the row sums (P <= 4)
        modes += [("owner blocks overlapped, pairwise send/recv, fp32 wire", "blocks", None, "p2p"),
                  ("owner blocks overlapped, broadcasts, fp32 wire", "blocks", None, "bcast")]
    return modes


def measure_partition(w, dev, rank, world, steps, warmup, want_e2e=False):
    """Strong scaling of one full-batch step on a 1-D destination-partitioned graph (BASELINE configs[3],
    ogbn-products-shaped): every rank owns N/world nodes, their in-edges (forward) and out-edges (backward); per
    layer one halo exchange of the row blocks forward and one of the g_out blocks backward; BatchNorm statistics,
    EMA code statistics and dense gradients are all-reduced.  The graph is static (full-batch training), so K0
    runs once.  Every mode of `partition_modes` is timed on the same graph and model; returns (on rank 0) the
    record the bench line carries, None elsewhere."""
    import torch.distributed as dist
    import stemgnn_b200 as S
    from stemgnn_b200 import _lib
    g = torch.Generator().manual_seed(4)
    n, e = w["n"], w["e"]
    lo, hi = S.partition_1d(n, world)[rank]
    # same seeded global edge list on every rank, generated in slices to bound host memory
    fwd_parts, bwd_parts, chunk = [], [], 1 << 24
    zipf = ZipfNodes(n, e, g) if w.get("degree") == "zipf" else None
    for c0 in range(0, e, chunk):
        ei = torch.randint(0, n, (2, min(chunk, e - c0)), generator=g)
        if zipf is not None:
            ei[1] = zipf.draw(ei.size(1), g)
        ei = ei.to(dev)
        fwd_parts.append(S.local_edges(ei, lo, hi, 1)[0])
        bwd_parts.append(S.local_edges(ei, lo, hi, 0)[0])
        del ei
    x_all = torch.randn(n, w["f"], generator=g)
    x_h = x_all[lo:hi].contiguous().pin_memory()
    del x_all
    split = (torch.cat(fwd_parts, 1), torch.cat(bwd_parts, 1))
    del fwd_parts, bwd_parts
    # BatchNorm statistics over ALL ranks' rows come from the library (K7 + norm_act.BatchSync: a
    # gather of (mean, var, n) forward, an all-reduce of the two column sums backward), not from
    # nn.SyncBatchNorm -- the fused BatchNorm + ReLU + dropout passes stay in use on every rank
    enc, vq, opt = build_models(w, dev, S)
    params = list(enc.parameters())

    def make_step(pg):
        def step(x):
            opt.zero_grad(set_to_none=True)
            z = enc(x, pg)
            q, ind, vq_loss, _ = vq(z)
            loss = mean_square(q, n * q.size(1)) + vq_loss.sum() * ((hi - lo) / n)
            loss.backward()
            if world > 1:
                flat = torch.cat([p.grad.reshape(-1) for p in params])
                dist.all_reduce(flat)
                o = 0
                for p in params:
                    p.grad.copy_(flat[o:o + p.numel()].view_as(p)); o += p.numel()
            opt.step()
            return loss
        return step

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    x_d = x_h.to(dev)
    pk = peaks()
    out_modes, extra = [], {}
    for label, mode, wire, transport in partition_modes(world):
        pg = S.PartitionedGraph(None, n, rank=rank, world=world, pre_split=split, halo_mode=mode,
                                halo_transport=transport if mode == "blocks" else "p2p",
                                halo_chunks=int(transport) if mode == "chunks" else 4,
                                halo_dtype=torch.bfloat16 if wire == "bf16" else None)
        step = make_step(pg)
        for _ in range(max(warmup, 3)):
            step(x_d)
        sync_all()
        _lib.TRACE = []
        l0 = _lib.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            step(x_d)
        e1.record()
        sync_all()
        launches = (_lib.launch_count() - l0) / steps
        trace, _lib.TRACE = _lib.TRACE, None
        times = [e0.elapsed_time(e1)]
        if want_e2e and not out_modes:
            x_s = torch.empty_like(x_d)
            loss_h = torch.empty(1).pin_memory()
            e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e2.record()
            for _ in range(steps):
                x_s.copy_(x_h, non_blocking=True)
                loss_h.copy_(step(x_s).detach().reshape(1), non_blocking=True)
            e3.record()
            sync_all()
            times.append(e2.elapsed_time(e3))
        t = torch.tensor(times, dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        kern = summarize_trace(trace, steps, pk, False)
        ms_of = lambda *names: sum(kern[k]["ms_per_step"] for k in names if k in kern)
        halo_bytes = sum(r[1] for r in trace if r[0] == "halo_wait") / steps
        rec = dict(mode=label, halo_mode=mode, wire="bf16" if wire == "bf16" else "f32", transport=transport,
                   ms_per_step=float(t[0]) / steps, halo_wait_ms=ms_of("halo_wait"), aggregate_ms=ms_of("aggregate"),
                   dense_ms=ms_of("linear", "linear_wgrad"), batchnorm_ms=ms_of("bn_stats", "bn_act_forward", "bn_act_backward"),
                   vq_ms=ms_of("vq_assign", "vq_code_stats", "vq_ema_apply", "vq_backward"),
                   halo_bytes_received_per_rank_per_step=halo_bytes, gpu_launches=launches,
                   note="per-family times are CUDA events around each C-ABI call on rank 0's compute stream in "
                        "the timed steps themselves; halo_wait is what the compute stream waited for")
        out_modes.append(rec)
        if not extra:
            extra = dict(kern=kern, launches=launches, ms_e2e=float(t[1]) if len(times) > 1 else None,
                         h2d=x_h.numel() * 4 * world)
        del pg, step
        torch.cuda.empty_cache()
    if rank != 0:
        return None
    fp32 = [m for m in out_modes if m["wire"] == "f32"]
    best = min(fp32, key=lambda m: m["ms_per_step"])
    rec = dict(workload=w["desc"], config="BASELINE configs[3]", scaling="strong", n_gpus=world, layers=w["layers"],
               nodes=n, edges=e, steps=steps, warmup=max(warmup, 3), ms_per_step=best["ms_per_step"],
               headline_mode=best["mode"] + " (fastest fp32-wire mode)",
               value=w["layers"] * e / (best["ms_per_step"] / 1e3), unit="edges/s",
               halo_wait_ms=best["halo_wait_ms"], aggregate_ms=best["aggregate_ms"], dense_ms=best["dense_ms"],
               wire=best["wire"], halo_bytes_received_per_rank_per_step=best["halo_bytes_received_per_rank_per_step"],
               modes=out_modes,
               parity="all-gather mode: bit-identical to one GPU; row-chunk / owner-block modes: same terms summed group by group, "
                      "<= 1e-5 of the output scale on the fp32 wire (tests/test_gpu_partition.py, "
                      "scripts/check_partition_multi.py); bf16 wire: 2^-9 relative per gathered feature, opt-in",
               l2="working set (N x F fp32 >= 980 MB) exceeds the 126 MB L2")
    try:
        if world == 1:
            json.dump(dict(ms_per_step=rec["ms_per_step"], nodes=n, edges=e, layers=w["layers"]), open(PARTITION_N1_FILE, "w"))
        elif os.path.exists(PARTITION_N1_FILE):
            n1 = json.load(open(PARTITION_N1_FILE))
            if (n1["nodes"], n1["edges"], n1["layers"]) == (n, e, w["layers"]):
                rec["n1_ms_per_step"] = n1["ms_per_step"]
                rec["efficiency_vs_n1"] = n1["ms_per_step"] / (world * rec["ms_per_step"])
                rec["n1_source"] = "this bench's own 1-GPU partition leg, run earlier on this box (" + PARTITION_N1_FILE + ")"
    except (OSError, ValueError, KeyError):
        pass
    rec["_extra"] = extra
    return rec


def run_partitioned(args, w):
    """`--partition` / `--workload products`: the partition leg as the whole bench line (strong scaling)."""
    import torch.distributed as dist
    import __graft_entry__ as G
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        init_nccl(dist, dev)
    if rank == 0:
        G.build()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    rec = measure_partition(w, dev, rank, world, args.steps, args.warmup, want_e2e=True)
    clocks = sampler.stop() if rank == 0 else None
    if rank == 0:
        pk = peaks()
        extra = rec.pop("_extra")
        kern = extra["kern"]
        roof = roofline_of(kern, "aggregate", pk, False)
        first = rec["modes"][0]
        units = w["layers"] * w["e"]
        line = dict(metric=METRIC, value=rec["value"], unit="edges/s", n_gpus=world,
                    steps=args.steps, warmup=max(args.warmup, 3), ms_per_step=rec["ms_per_step"],
                    higher_is_better=True, scaling="strong", vs_baseline=None, dtype="f32",
                    data="synthetic", vq_nodes_per_s=w["n"] / (rec["ms_per_step"] / 1e3),
                    config=dict(workload=w["desc"], layers=w["layers"], nodes=w["n"], edges=w["e"],
                                l2=rec["l2"], parallelism=f"1-D destination partition x{world}: " + rec["headline_mode"]),
                    e2e=dict(value=units * args.steps / (extra["ms_e2e"] / 1e3), unit="edges/s",
                             ms_per_step=extra["ms_e2e"] / args.steps, mode=first["mode"],
                             h2d_bytes_per_step=extra["h2d"], d2h_bytes_per_step=4 * world),
                    gpu_launches=extra["launches"], roofline=roof, kernels=kern, partition=rec, clocks=clocks)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None,
                    help="timed steps; default 300 (~1 s of timed region at 3.5 ms per step), 20 for --impl reference")
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="arxiv", choices=sorted(WORKLOADS))
    ap.add_argument("--cpu-frac", type=float, default=1.0,
                    help="fraction of the graph the CPU arm runs per step (bounded sample)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--warm-seconds", type=float, default=1.0,
                    help="untimed sustained run of the same step right before the K timed steps")
    ap.add_argument("--no-partition", action="store_true",
                    help="skip the nested `partition` record (configs[3] on a 1-D partition over the same ranks)")
    ap.add_argument("--partition-steps", type=int, default=5, help="timed steps per mode of the partition record")
    ap.add_argument("--dump-sequence", default=None,
                    help="write the C-ABI call sequence of one step (name, kernels launched, ms) to this JSON file")
    ap.add_argument("--no-graph", action="store_true", help="launch the step eagerly instead of replaying a CUDA graph")
    ap.add_argument("--partition", action="store_true",
                    help="1-D destination partition across the ranks (strong scaling, static graph)")
    ap.add_argument("--degree", default="uniform", choices=["uniform", "zipf"],
                    help="in-degree distribution of the synthetic graph: uniform endpoints (the headline "
                         "configuration) or SURVEY.md 8d's power-law variant with hubs of up to ~17,000 in-edges")
    args = ap.parse_args()
    if args.steps is None:
        args.steps = 20 if args.impl == "reference" else (300 if args.workload == "arxiv" and not args.partition else 20)
    w = dict(WORKLOADS[args.workload])
    if args.degree == "zipf":
        w["degree"] = "zipf"
        w["desc"] += f"; destinations Zipf(1) over a random node order, capped at ~{ZIPF_CAP:,} in-edges"
    if args.impl == "reference":
        run_reference(args, w)
    elif args.partition or args.workload == "products":
        run_partitioned(args, w)
    else:
        run_ours(args, w)


if __name__ == "__main__":
    main()
