"""BASELINE configs[2]: one pretrain epoch over the multi-graph mix -- the disjoint
union of Cora / CiteSeer / PubMed / arxiv-sized synthetic graphs (195,095 nodes, 1,274,551 edges,
768-wide text features, 237 edge types), weight-5 seeds = 975,475 seeds -> 953 mini-batches of 1,024
seeds, fan-out [10, 10] (reference STEM-GNN/pretrain.py:28-66,143-154, config/pt_data.yaml).

Every stage of a mini-batch runs on the device through the library: K10 sampler -> K9 row gather of
the text features -> K9 mask_feature + dropout_adj(force_undirected) -> K0 CSR -> 2-layer SAGE
768-768 with the edge-type table (K1/K2, K5/K6, K7) -> shipped VQ 4 x 128 x 768 cosine (K3) ->
backward -> AdamW.  Prints one JSON line (whole-epoch device time, CUDA events).
    python scripts/mix_epoch.py [--batches B] [--cpu-batches C]"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

PARTS = [(2708, 10556), (3327, 9104), (19717, 88648), (169343, 1166243)]
F, T, HID, LAYERS, BATCH, FANOUT, WEIGHT = 768, 237, 768, 2, 1024, [10, 10], 5


def mix_graph(seed):
    g = torch.Generator().manual_seed(seed)
    eis, off = [], 0
    for n, e in PARTS:
        eis.append(torch.randint(0, n, (2, e), generator=g) + off)
        off += n
    ei = torch.cat(eis, 1)
    return off, ei, torch.randn(off, F, generator=g), torch.randn(T, F, generator=g), \
        torch.randint(0, T, (ei.size(1),), generator=g)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batches", type=int, default=0, help="0 = the full epoch (953)")
    ap.add_argument("--cpu-batches", type=int, default=3)
    ap.add_argument("--encoder-vq-only", action="store_true",
                    help="stand-in loss on the encoder + VQ only (the earlier form of this script) instead of the "
                         "reference's five pretrain losses")
    ap.add_argument("--trace", type=int, default=5, help="batches traced call by call after the timed loop")
    args = ap.parse_args()
    import torch.distributed as dist
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:                                                     # replicas over mini-batches (C3)
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    import __graft_entry__ as G
    if rank == 0:
        G.build()
    if world > 1:
        dist.barrier()
    import stemgnn_b200 as S
    from stemgnn_b200 import _lib
    n, ei, text, etable, etype = mix_graph(0)
    data = dict(edge_index=ei.to(dev), x=torch.arange(n, device=dev), xe=etype.to(dev), num_nodes=n)
    text_d, etable_d = text.to(dev), etable.to(dev)
    seeds = torch.arange(n).repeat(WEIGHT)                            # get_train_node_idx, weight 5
    all_batches = -(-seeds.numel() // BATCH)
    if world > 1:                                                     # every rank samples its own share of the seeds
        seeds = seeds[torch.randperm(seeds.numel(), generator=torch.Generator().manual_seed(7))][rank::world]
        seeds = seeds[:(all_batches // world) * BATCH]
    loader = S.NeighborLoader(data, num_neighbors=FANOUT, input_nodes=seeds, batch_size=BATCH, shuffle=True,
                              seed=1 + rank)
    torch.manual_seed(0)
    enc = S.Encoder(F, HID, nn.ReLU, LAYERS, backbone="sage", normalize="batch", dropout=0.15).to(dev).train()
    vq = S.VectorQuantize(dim=HID, codebook_size=128, codebook_dim=HID, heads=4, separate_codebook_per_head=True,
                          decay=0.8, commitment_weight=10, use_cosine_sim=True, orthogonal_reg_weight=1,
                          orthogonal_reg_max_codes=32, orthogonal_reg_active_codes_only=False, kmeans_init=False,
                          ema_update=False).to(dev).train()
    # the rest of PretrainModel (model/pt_model.py:12-23) and its decoders (pretrain.py:121-123)
    import copy
    feat_dec = nn.Linear(HID, F).to(dev)
    topo_dec = S.InnerProductDecoder(hidden_dim=HID, output_dim=HID).to(dev)
    topo_sem_dec = nn.Linear(2 * HID, HID).to(dev)
    sem_enc = copy.deepcopy(enc)
    sem_proj = nn.Linear(HID, HID).to(dev)
    full = not args.encoder_vq_only
    modules = [enc, vq] + ([feat_dec, topo_dec, topo_sem_dec, sem_proj] if full else [])
    params = [p for m in modules for p in m.parameters()]
    opt = torch.optim.AdamW(params, lr=1e-4, capturable=True)
    EPS, RATIO = 1e-15, 0.1                                           # pt_model.py:8, utils/args.py:46
    ar_events = []

    def losses_of(z, quantize, commit, x, ei, et, bs):
        """PretrainModel.forward (pt_model.py:128-157) with the default weights of utils/args.py:44-51."""
        query = quantize
        feat = nn.functional.mse_loss(feat_dec(query[:bs]), x[:bs])                      # :42-43
        e = ei.size(1)
        k = max(int(e * RATIO), 1)
        pos = ei[:, torch.randperm(e, device=dev)[:k]]                                   # :51-57
        neg = S.negative_sampling(pos, z.size(0))                                        # :60
        topo = -torch.log(topo_dec(query, pos, sigmoid=True) + EPS).mean() \
               - torch.log(1 - topo_dec(query, neg, sigmoid=True) + EPS).mean()          # :62-65
        pick = torch.randperm(e, device=dev)[:k]                                         # :72-78
        target = S.gather_rows(etable_d, et[pick])                                       # edge_attr[perm]
        topo_sem = nn.functional.mse_loss(S.edge_concat_linear(query, ei[:, pick], topo_sem_dec), target)   # :80-81
        with torch.no_grad():
            zs = sem_enc(x, ei, (etable_d, et))                                          # :92
        h = sem_proj(quantize)
        zs, h = nn.functional.normalize(zs[:bs], dim=-1), nn.functional.normalize(h[:bs], dim=-1)
        sem = (1 - (zs * h).sum(-1)).mean()                                              # :95-99
        return 100 * feat + 0.01 * topo + 100 * topo_sem + 1 * sem + commit.sum()        # pretrain.py:51-58

    def step(batch):
        x = S.gather_rows(text_d, batch.x)                            # node_text_feat[data.x]
        aug_x, _ = S.mask_feature(x, p=0.2)
        aug_ei, aug_t = S.dropout_adj(batch.edge_index, batch.xe, p=0.2, force_undirected=True,
                                      num_nodes=x.size(0))
        opt.zero_grad(set_to_none=True)
        z = enc(aug_x, aug_ei, (etable_d, aug_t))                     # edge_text_feat[data.xe] as a table
        q, ind, vq_loss, _ = vq(z)
        if full:
            loss = losses_of(z, q, vq_loss, x, batch.edge_index, batch.xe, batch.batch_size)
        else:
            loss = q.square().mean() + vq_loss.sum()
        loss.backward()
        if world > 1:                                                 # one flat all-reduce of the gradients
            ar0 = torch.cuda.Event(enable_timing=True); ar1 = torch.cuda.Event(enable_timing=True)
            ar0.record()
            flat = torch.cat([p.grad.reshape(-1) for p in params])
            dist.all_reduce(flat)
            flat /= world
            o = 0
            for p in params:
                p.grad.copy_(flat[o:o + p.numel()].view_as(p)); o += p.numel()
            ar1.record()
            ar_events.append((ar0, ar1))
        if full:
            nn.utils.clip_grad_norm_(params, 1.0)                     # pretrain.py:62
        opt.step()
        if full:
            with torch.no_grad():                                     # ema_update_sem_encoder, pt_model.py:104-106
                torch._foreach_mul_(list(sem_enc.parameters()), 0.99)
                torch._foreach_add_(list(sem_enc.parameters()), list(enc.parameters()), alpha=0.01)
        return loss, aug_ei.size(1) + (batch.edge_index.size(1) if full else 0), x.size(0)

    total = len(loader) if args.batches <= 0 else min(args.batches, len(loader))
    it = iter(loader)
    for _ in range(3):                                                # warm-up batches
        step(next(it))
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    l0 = _lib.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    edges = nodes = done = 0
    kept = []
    t0 = time.perf_counter()
    e0.record()
    for batch in loader:
        loss, m, nn_ = step(batch)
        edges += m; nodes += nn_; done += 1
        if len(kept) < args.cpu_batches:
            kept.append(batch)
        if done >= total:
            break
    e1.record()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    ms = e0.elapsed_time(e1)
    if world > 1:                                                     # max over ranks; totals over ranks
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)
        tot = torch.tensor([edges, nodes], dtype=torch.float64, device=dev)
        dist.all_reduce(tot)
        edges, nodes = float(tot[0]), float(tot[1])
    launches = _lib.launch_count() - l0
    ar_ms = sum(a.elapsed_time(b) for a, b in ar_events[3:]) / max(len(ar_events) - 3, 1) if ar_events else 0.0
    line = dict(metric="edges/sec encoder fwd+bwd & VQ nodes/sec", value=LAYERS * edges / ms * 1e3, unit="edges/s",
                n_gpus=world, steps=done, ms_per_step=ms / done, epoch_s_extrapolated=ms / done * len(loader) / 1e3,
                scaling="strong (the epoch's mini-batches are dealt over the ranks)",
                wall_s=wall, vq_nodes_per_s=nodes / ms * 1e3, dtype="f32", data="synthetic",
                grad_allreduce_ms_per_step=ar_ms, gpu_launches=launches / done, final_loss=float(loss.detach()),
                config=dict(workload="pretrain mix (195,095 nodes, 1,274,551 edges, 768 feats, 237 edge types), "
                            "mini-batches of 1,024 seeds, fan-out [10,10], 2-layer SAGE 768-768 + shipped VQ 4x128x768" + (", the reference's five pretrain losses (feature / topology / edge-semantics / tree-semantics reconstruction + commitment), clip, AdamW, EMA teacher encoder" if full else ", stand-in loss"),
                            batches_per_epoch_per_rank=len(loader), mean_nodes=nodes / done / world,
                            mean_aug_edges=edges / done / world,
                            parallelism="single GPU" if world == 1 else f"dp{world}: replicas over mini-batches, flat gradient all-reduce (NCCL)"))
    if args.trace > 0 and world == 1:                                                # per-call device times (CUDA events)
        _lib.TRACE = []
        ta, tb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ta.record()
        for _, batch in zip(range(args.trace), loader):
            step(batch)
        tb.record()
        torch.cuda.synchronize()
        fam = {}
        for name, nbytes, flops, launches, a, b in _lib.TRACE:
            d = fam.setdefault(name, [0.0, 0])
            d[0] += a.elapsed_time(b); d[1] += 1
        _lib.TRACE = None
        line["traced_ms_per_step"] = ta.elapsed_time(tb) / args.trace
        line["kernels_ms_per_step"] = {k: round(v[0] / args.trace, 4) for k, v in sorted(fam.items(), key=lambda kv: -kv[1][0])}
    if args.cpu_batches > 0 and rank == 0:
        from oracle.encoder_ref import EncoderRef
        torch.set_num_threads(os.cpu_count() or 1)
        ref = EncoderRef(F, HID, nn.ReLU, LAYERS, backbone="sage", normalize="batch", dropout=0.15).train()
        ropt = torch.optim.AdamW(ref.parameters(), lr=1e-4)
        t0, ce = time.perf_counter(), 0
        for b in kept:                                                # encoder part of the step only
            x = text[b.x.cpu()]
            bei, bt = b.edge_index.cpu(), b.xe.cpu()
            ropt.zero_grad()
            z = ref(x, bei, etable[bt])
            z.square().mean().backward()
            ropt.step()
            ce += bei.size(1)
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = dict(value=LAYERS * ce / dt, unit="edges/s", ms_per_step=dt / len(kept) * 1e3,
                                    cores=os.cpu_count(), kind="port",
                                    sample=f"{len(kept)} sampled mini-batches, encoder fwd+bwd+AdamW only "
                                           "(no sampler, no augmentation, no VQ), oracle/ restatement")
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        os._exit(0)


if __name__ == "__main__":
    main()
