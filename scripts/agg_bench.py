"""A/B timing of the aggregation kernel (K1/K2) alone on graph shapes of the bench workloads.
CUDA events on the current stream, L2 flushed between launches.  Usage: python scripts/agg_bench.py [tag]"""
import importlib, json, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("stem-gnn_b200")
from importlib import import_module
ops = import_module("stem-gnn_b200.ops"); graph = import_module("stem-gnn_b200.graph")

def skewed_dst(n, e, g, cap=17000):
    """SURVEY.md 8d's power-law variant: destinations drawn with p(rank k) ~ 1/k (Zipf, alpha = 1), no node
    above `cap` edges in expectation (ogbn-products' largest degree is 17,481), ranks scattered over the
    node ids by a random permutation so that hubs are not the first rows the launch starts."""
    w = 1.0 / torch.arange(1, n + 1, dtype=torch.float64)
    w = torch.minimum(w, torch.tensor(cap / e, dtype=torch.float64) * w.sum())
    dst_rank = torch.multinomial(w, e, replacement=True, generator=g)
    return torch.randperm(n, generator=g)[dst_rank]


def run(n, e, f, reps=5, seed=0, skew=False):
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(seed)
    ei = torch.randint(0, n, (2, e), generator=g)
    if skew:
        ei[1] = skewed_dst(n, e, g)
    max_deg = int(torch.bincount(ei[1], minlength=n).max())
    ei = ei.to(dev)
    x = torch.randn(n, f, device=dev, requires_grad=True)
    gs = graph.GraphStructure(ei, n)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    res = {}
    for name in ("fwd", "bwd"):
        ts = []
        for r in range(reps + 2):
            flush.zero_()
            if name == "fwd":
                s, t = torch.cuda.Event(True), torch.cuda.Event(True)
                s.record(); out = ops.aggregate(x, gs, reduce="mean"); t.record()
            else:
                out = ops.aggregate(x, gs, reduce="mean"); go = torch.ones_like(out)
                flush.zero_()
                s, t = torch.cuda.Event(True), torch.cuda.Event(True)
                s.record(); out.backward(go); t.record(); x.grad = None
            torch.cuda.synchronize()
            if r >= 2: ts.append(s.elapsed_time(t) * 1e3)
        res[name] = min(ts)
    gb_gather = (e * f * 4 + n * f * 4 + e * 4 + n * 4) / 1e9
    gb_min = (2 * n * f * 4 + e * 4 + n * 4) / 1e9
    print(json.dumps(dict(n=n, e=e, f=f, skew=skew, max_in_deg=max_deg, fwd_us=round(res["fwd"], 1), bwd_us=round(res["bwd"], 1),
                          fwd_gather_GBps=round(gb_gather / res["fwd"] * 1e6), fwd_min_GBps=round(gb_min / res["fwd"] * 1e6))), flush=True)

if __name__ == "__main__":
    print("tag", sys.argv[1:], "no_hub =", os.environ.get("STEMGNN_AGG_NO_HUB"), "ring =", os.environ.get("STEMGNN_AGG_HUB_RING"))
    if "skew" in sys.argv[1:]:
        for sk in (False, True):
            run(169343, 1166243, 128, skew=sk); run(169343, 1166243, 256, skew=sk)
        if "small" not in sys.argv[1:]:
            run(2449029 // 2, 61859140 // 2, 256, reps=2, skew=True)
        sys.exit(0)
    run(169343, 1166243, 128); run(169343, 1166243, 256)
    run(2449029 // 2, 61859140 // 2, 100); run(2449029 // 2, 61859140 // 2, 256)
    run(2708, 10556, 1432); run(500000, 5000000, 64); run(500000, 2000000, 512)
