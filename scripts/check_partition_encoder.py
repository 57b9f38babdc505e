"""torchrun --nproc-per-node P scripts/check_partition_encoder.py : a 2-layer SAGE encoder WITH
BatchNorm on a 1-D destination partition over NCCL (halo all-gathers + the library's full-batch
BatchNorm statistics, norm_act.BatchSync) against the same encoder on the whole graph computed locally:
rows of the output, of the input gradient, and the all-reduced parameter gradients."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn as nn, torch.distributed as dist
import stemgnn_b200 as S


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n, e, f, hid = 50000, 600000, 64, 128
    g = torch.Generator().manual_seed(3)
    ei = torch.randint(0, n, (2, e), generator=g).cuda()
    x = torch.randn(n, f, generator=g).cuda()
    up = torch.randn(n, hid, generator=g).cuda()
    torch.manual_seed(0)
    full = S.Encoder(f, hid, nn.ReLU, 2, backbone="sage", normalize="batch", dropout=0.0).cuda().train()
    part = S.Encoder(f, hid, nn.ReLU, 2, backbone="sage", normalize="batch", dropout=0.0).cuda().train()
    part.load_state_dict(full.state_dict())
    xf = x.clone().requires_grad_(True)
    zf = full(xf, ei)
    (zf * up).sum().backward()
    pg = S.PartitionedGraph(ei, n)
    lo, hi = pg.lo, pg.hi
    xl = x[lo:hi].clone().requires_grad_(True)
    zl = part(xl, pg)
    (zl * up[lo:hi]).sum().backward()
    rel = lambda a, b: ((a - b).abs().max() / (b.abs().max() + 1e-30)).item()
    errs = dict(z=rel(zl.detach(), zf.detach()[lo:hi]), gx=rel(xl.grad, xf.grad[lo:hi]))
    for (k, p), (_, q) in zip(part.named_parameters(), full.named_parameters()):
        gsum = p.grad.clone()
        dist.all_reduce(gsum)                                  # a rank holds its rows' share of every gradient
        errs[k] = rel(gsum, q.grad)
    errs["running_var"] = rel(part.norms[0].running_var, full.norms[0].running_var)
    worst = torch.tensor([max(v for k, v in errs.items() if not k.endswith("lin_l.bias"))], device="cuda")
    dist.all_reduce(worst, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"partitioned encoder with BatchNorm, world={world}: max relative error {worst.item():.2e}  "
              + " ".join(f"{k}={v:.1e}" for k, v in errs.items()), flush=True)
    dist.barrier()
    os._exit(0 if worst.item() < 1e-4 else 1)


if __name__ == "__main__":
    main()
