"""Timing of K9 (dropout_adj / mask_feature / gather_rows) and K10 (neighbour sampler) alone on the
bench graph shapes, with torch's own ops beside them.  CUDA events, 3 warm-ups, L2 flushed between
repetitions.  Usage: python scripts/aug_bench.py [arxiv|products|mix]"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import stemgnn_b200 as S  # noqa: E402

DEV = "cuda"
SHAPES = {"arxiv": (169343, 1166243, 128), "mix": (195095, 1274551, 768), "products": (2449029, 61859140, 100)}
FLUSH = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(reps):
        FLUSH.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def torch_dropout_adj(ei, p):
    row, col = ei
    mask = torch.rand(row.size(0), device=ei.device) >= p
    mask[row > col] = False
    row, col = row[mask], col[mask]
    return torch.stack([torch.cat([row, col]), torch.cat([col, row])])


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "arxiv"
    n, e, f = SHAPES[name]
    g = torch.Generator().manual_seed(0)
    ei = torch.randint(0, n, (2, e), generator=g).to(DEV)
    p = 0.2
    t = timed(lambda: S.dropout_adj(ei, p=p, force_undirected=True, return_edge_id=True))
    out = S.dropout_adj(ei, p=p, force_undirected=True, return_edge_id=True)[0]
    m = out.size(1)
    nbytes = e * 4 + 2 * e * 20 + m * 24            # rand write + two passes + outputs
    print(f"{name} dropout_adj(undirected, torch.rand) E={e} -> {m}: {t:.3f} ms  {nbytes / t / 1e6:.0f} GB/s (incl. host sync)")
    t2 = timed(lambda: S.dropout_adj(ei, p=p, force_undirected=True, generator='philox', return_edge_id=True))
    print(f"{name} dropout_adj(undirected, philox): {t2:.3f} ms")
    t3 = timed(lambda: torch_dropout_adj(ei, p))
    print(f"{name} torch ops (PyG's sequence): {t3:.3f} ms  -> {t3 / t:.2f}x")

    rows = min(n, 400_000)
    x = torch.randn(rows, f if f % 4 == 0 else f + 4 - f % 4, device=DEV)
    t = timed(lambda: S.mask_feature(x, p=0.2))
    tt = timed(lambda: x.masked_fill(~(torch.rand(x.size(1), device=DEV) >= 0.2).view(1, -1), 0.0))
    print(f"{name} mask_feature col [{rows}x{x.size(1)}]: {t:.3f} ms {2 * x.numel() * 4 / t / 1e6:.0f} GB/s (torch {tt:.3f} ms)")
    ids = torch.randint(0, rows, (rows,), device=DEV)
    t = timed(lambda: S.gather_rows(x, ids))
    tt = timed(lambda: x[ids])
    print(f"{name} gather_rows [{rows}x{x.size(1)}]: {t:.3f} ms {2 * x.numel() * 4 / t / 1e6:.0f} GB/s (torch {tt:.3f} ms)")

    if name != "products":
        smp = S.NeighborSampler(ei, n, [10, 10], seed=1)
        seeds = torch.randperm(n, device=DEV)[:1024]
        t = timed(lambda: smp.sample_padded(seeds), reps=20)
        n_id, sei, e_id = smp.sample(seeds)
        print(f"{name} neighbor_sample 1024 seeds [10,10]: {t:.3f} ms device  ({n_id.numel()} nodes, {e_id.numel()} edges) "
              f"-> {e_id.numel() / t / 1e3:.1f} M sampled edges/s")
        t0 = time.perf_counter()
        for _ in range(50):
            smp.sample(seeds)
        torch.cuda.synchronize()
        print(f"{name} neighbor_sample incl. host (sizes read back): {(time.perf_counter() - t0) / 50 * 1e3:.3f} ms/batch")


if __name__ == "__main__":
    main()
