"""Run-to-run bit stability of the VQ assignment (K3 path 2: vq_tc_kernel -> vq_gather_kernel ->
vq_rescan_kernel) with the workspace pre-filled with 0xFF / 0x00 / random bytes, against the exact scan
(path 1) and the fp64 scores.  Written for the round-1 driver record in which one of two identical
smoke() runs (the one under ncu) reported `vq indices mismatch`.

    python scripts/vq_stability.py [reps]          (B200; prints one line per case, exit 1 on any difference)
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import stemgnn_b200 as S
from stemgnn_b200 import _lib

DEV = "cuda"


def smoke_inputs():
    """the exact VQ inputs of __graft_entry__.smoke()"""
    gen = torch.Generator().manual_seed(0)
    n, e, f = 500, 4000, 64
    torch.randn(n, f, generator=gen)
    torch.randint(0, n, (2, e), generator=gen)
    torch.randn(n, f, generator=gen)
    torch.manual_seed(0)
    vq = S.VectorQuantize(dim=64, codebook_size=128, decay=0.8)
    z = torch.randn(n, 64, generator=gen)
    return z.to(DEV), vq._codebook.embed.detach().clone().to(DEV)


def c2_inputs(rows=169343, K=512, D=256):
    g = torch.Generator(device=DEV).manual_seed(5)
    x = torch.randn(rows, D, generator=g, device=DEV)
    torch.manual_seed(1)
    vq = S.VectorQuantize(dim=D, codebook_size=K, decay=0.8)
    return x, vq._codebook.embed.detach().clone().to(DEV)


def fp64_gap(x, embed, rows, a, b, metric=S.METRIC_EUCLID):
    """fp64 score difference between codes a and b for the given rows"""
    xs, e = x[rows].double(), embed[0].double()
    if metric == S.METRIC_DOT:
        return ((xs * e[a]).sum(-1) - (xs * e[b]).sum(-1)).abs()
    da = ((xs - e[a]) ** 2).sum(-1).sqrt()
    db = ((xs - e[b]) ** 2).sum(-1).sqrt()
    return (da - db).abs()


def run_case(name, x, embed, reps, metric=S.METRIC_EUCLID):
    bad = 0
    _lib.set_workspace_poison("")
    i1, q1, sq1, _ = S.vq_assign_raw(x, embed, 1, metric, S.PATH_EXACT)
    torch.cuda.synchronize()
    first = None
    for mode in ("ff", "00", "rand", ""):
        _lib.set_workspace_poison(mode)
        for rep in range(reps):
            i2, q2, sq2, st = S.vq_assign_raw(x, embed, 1, metric, S.PATH_TENSOR, want_stats=True)
            if first is None:
                first = (i2.clone(), q2.clone(), sq2.clone(), st.clone())
                d = (i1 != i2).view(-1).nonzero().view(-1)
                gaps = fp64_gap(x, embed, d, i1.view(-1)[d], i2.view(-1)[d], metric) if d.numel() else torch.zeros(0)
                print(f"{name}: path2 vs path1: {d.numel()} differing rows of {i1.numel()}, "
                      f"max fp64 gap {gaps.max().item() if d.numel() else 0.0:.3e}, stats {st.tolist()}", flush=True)
                if d.numel() and gaps.max().item() > 1e-6 * 8:
                    bad += 1
                continue
            same = (torch.equal(i2, first[0]) and torch.equal(q2, first[1]) and torch.equal(sq2, first[2])
                    and torch.equal(st, first[3]))
            if not same:
                bad += 1
                d = (i2 != first[0]).view(-1).nonzero().view(-1)
                print(f"{name}: UNSTABLE poison={mode!r} rep={rep}: {d.numel()} rows differ from the first run; "
                      f"rows {d[:8].tolist()} now {i2.view(-1)[d[:8]].tolist()} first {first[0].view(-1)[d[:8]].tolist()} "
                      f"exact {i1.view(-1)[d[:8]].tolist()} stats {st.tolist()} vs {first[3].tolist()} "
                      f"sq {sq2.item()!r} vs {first[2].item()!r}", flush=True)
    _lib.set_workspace_poison("")
    torch.cuda.synchronize()
    print(f"{name}: {4 * reps} runs, {bad} unstable", flush=True)
    return bad


def main():
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 50
    bad = 0
    x, e = smoke_inputs()
    bad += run_case("smoke 500x64 K=128", x, e, reps)
    x, e = c2_inputs()
    bad += run_case("C2 169343x256 K=512", x, e, max(reps // 4, 5))
    x, e = c2_inputs(20000, 4096, 128)
    bad += run_case("20000x128 K=4096", x, e, max(reps // 4, 5))
    g = torch.Generator(device=DEV).manual_seed(9)
    x = torch.nn.functional.normalize(torch.randn(43000, 768, generator=g, device=DEV), dim=-1)
    e = torch.nn.functional.normalize(torch.randn(1, 128, 768, generator=g, device=DEV), dim=-1)
    bad += run_case("cosine 43000x768 K=128", x, e, max(reps // 4, 5), S.METRIC_DOT)
    print("vq_stability:", "FAILED" if bad else "ok")
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
