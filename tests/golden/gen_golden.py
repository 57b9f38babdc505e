"""Generates the golden fixtures in this directory by executing the UNMODIFIED reference code
from /root/reference (only available in the build container, never on the GPU box):

* vq_*.npz  -- /root/reference/STEM-GNN/model/vq.py VectorQuantize (imports torch + einops only)
* vq_c2_*.npz, vq_shipped_full.npz -- the same module at BASELINE configs[1]'s codebook shape (512 x 256) and in
               the shipped configuration at its real shape (4 x 128 x 768 + projections); COMPACT fixtures:
               inputs / initial state are regenerated from seeds by tests/golden/cases.py (checked here), outputs
               are stored as indices, loss and row / column sums
* enc_c2.npz -- Encoder at configs[1]'s widths (128 -> 256 -> 256), seeded parameters, compact outputs
* enc_*.npz, dec_*.npz -- /root/reference/STEM-GNN/model/encoder.py Encoder / InnerProductDecoder,
               executed on oracle/pyg_shim
               (torch_geometric / torch_scatter are not installable here; see oracle/__init__.py)

Run:  python tests/golden/gen_golden.py [kmeans | dec | vq_c2]     (CPU, fp32, seeded; a few seconds)
"""
import os
import sys
import warnings

import numpy as np
import torch
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
OUT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle", "pyg_shim"))
sys.path.insert(0, "/root/reference/STEM-GNN")
warnings.filterwarnings("ignore")

from model.vq import VectorQuantize  # noqa: E402  (the reference's own module)
from model.encoder import Encoder, InnerProductDecoder  # noqa: E402  (the reference's own module, on the shim)

from tests.golden.cases import (DEC_CASES, ENC_CASES, KMEANS_CASES, VQ_C2_CASES, VQ_C2_ROWS, VQ_CASES,  # noqa: E402
                                ENC_C2, VQ_SHIPPED_FULL, VQ_SHIPPED_ROWS, c2_inputs, enc_c2_inputs, marginals,
                                seeded_parameters, shipped_inputs, strided_sample_fn)


def gen_vq(name, cfg):
    cfg = dict(cfg)
    is_eval, dup = cfg.pop("_eval", False), cfg.pop("_dup", False)
    torch.manual_seed(0)
    vq = VectorQuantize(**cfg)
    vq.train(not is_eval)
    if dup:
        with torch.no_grad():
            vq._codebook.embed[0, 5] = vq._codebook.embed[0, 2]
            vq._codebook.embed[0, 7] = vq._codebook.embed[0, 2]
    n = 200
    x = torch.randn(n, cfg["dim"])
    if dup:
        x[:40] = vq._codebook.embed[0, 2] + 0.01 * torch.randn(40, cfg["dim"])
    x.requires_grad_(True)
    before = {k: v.detach().clone().numpy() for k, v in vq.state_dict().items()}
    torch.manual_seed(1)  # fixes the randperm inside the orthogonal loss (vq.py:1024)
    q, ind, loss, oq = vq(x)
    out = dict(x=x.detach().numpy(), quantize=q.detach().numpy(), embed_ind=ind.numpy(),
               loss=loss.detach().numpy(), orig_quantize=oq.detach().numpy())
    if not is_eval:
        g = torch.randn(q.shape, generator=torch.Generator().manual_seed(2))
        ((q * g).sum() + loss.sum()).backward()
        out["g_quantize"] = g.numpy()
        out["g_x"] = x.grad.numpy()
        emb = vq._codebook.embed
        if isinstance(emb, nn.Parameter) and emb.grad is not None:
            out["g_embed"] = emb.grad.numpy()
    for k, v in before.items():
        out["before." + k] = v
    for k, v in vq.state_dict().items():
        out["after." + k] = v.detach().numpy()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(name, {k: v.shape for k, v in out.items() if not k.startswith(("before", "after"))})


def gen_vq_c2(name, cfg):
    """The reference module at BASELINE configs[1]'s codebook shape; compact fixture (see cases.py)."""
    torch.manual_seed(0)
    vq = VectorQuantize(**cfg)
    vq.train()
    x = torch.randn(VQ_C2_ROWS, cfg["dim"])
    x2, state = c2_inputs(cfg)
    assert torch.equal(x, x2), "c2_inputs() does not regenerate the reference's input"
    for k in ("embed", "embed_avg", "cluster_size"):
        assert torch.equal(getattr(vq._codebook, k).detach(), state[k]), f"c2_inputs() state mismatch: {k}"
    x.requires_grad_(True)
    q, ind, loss, oq = vq(x)
    g = torch.randn(q.shape, generator=torch.Generator().manual_seed(2))
    ((q * g).sum() + loss.sum()).backward()
    out = dict(embed_ind=ind.numpy(), loss=loss.detach().numpy(),
               x_check=np.array([x.detach().double().sum().item(), x.detach().double().abs().sum().item()]),
               embed_check=np.array([state["embed"].double().sum().item(), state["embed"].double().abs().sum().item()]))
    for key, t in (("quantize", q), ("g_x", x.grad), ("after.embed", vq._codebook.embed),
                   ("after.embed_avg", vq._codebook.embed_avg)):
        out[key + ".rows"], out[key + ".cols"] = marginals(t.detach().numpy())
    out["after.cluster_size"] = vq._codebook.cluster_size.detach().numpy()
    emb = vq._codebook.embed
    if isinstance(emb, nn.Parameter) and emb.grad is not None:
        out["g_embed.rows"], out["g_embed.cols"] = marginals(emb.grad.numpy())
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(name, ind.shape, float(loss), int(np.unique(ind.numpy()).size), "codes used")


def gen_vq_shipped_full():
    """The reference module in the shipped configuration at its real shape; compact fixture."""
    torch.manual_seed(0)
    vq = VectorQuantize(**VQ_SHIPPED_FULL)
    vq.train()
    x = torch.randn(VQ_SHIPPED_ROWS, VQ_SHIPPED_FULL["dim"])
    x2, pin, pout, state = shipped_inputs()
    assert torch.equal(x, x2), "shipped_inputs() does not regenerate the reference's input"
    assert torch.equal(vq.project_in.weight, pin.weight) and torch.equal(vq.project_in.bias, pin.bias)
    assert torch.equal(vq.project_out.weight, pout.weight) and torch.equal(vq.project_out.bias, pout.bias)
    for k in ("embed", "embed_avg", "cluster_size"):
        assert torch.equal(getattr(vq._codebook, k).detach(), state[k]), f"shipped_inputs() state mismatch: {k}"
    x.requires_grad_(True)
    torch.manual_seed(1)                                   # the randperm of the orthogonal loss (vq.py:1024)
    q, ind, loss, oq = vq(x)
    g = torch.randn(q.shape, generator=torch.Generator().manual_seed(2))
    ((q * g).sum() + loss.sum()).backward()
    out = dict(embed_ind=ind.numpy(), loss=loss.detach().numpy(),
               x_check=np.array([x.detach().double().sum().item(), x.detach().double().abs().sum().item()]),
               embed_check=np.array([state["embed"].double().sum().item(), state["embed"].double().abs().sum().item()]),
               pin_check=np.array([pin.weight.double().sum().item(), pout.weight.double().abs().sum().item()]))
    for key, t in (("quantize", q), ("orig_quantize", oq), ("g_x", x.grad), ("g_embed", vq._codebook.embed.grad),
                   ("g_project_in", vq.project_in.weight.grad), ("g_project_out", vq.project_out.weight.grad)):
        out[key + ".rows"], out[key + ".cols"] = marginals(t.detach().numpy())
    np.savez_compressed(os.path.join(OUT, "vq_shipped_full.npz"), **out)
    print("vq_shipped_full", ind.shape, float(loss), [int(np.unique(ind[:, h].numpy()).size) for h in range(4)], "codes used per head")


def gen_enc_c2():
    """The reference Encoder (on oracle/pyg_shim) at configs[1]'s widths; compact fixture."""
    c = ENC_C2
    enc = Encoder(c["in_dim"], c["hid"], nn.ReLU, c["layers"], backbone=c["backbone"], normalize=c["normalize"],
                  dropout=0.0)
    seeded_parameters(enc).train()
    x, ei, gz = enc_c2_inputs()
    x.requires_grad_(True)
    z = enc(x, ei)
    (z * gz).sum().backward()
    out = dict(keys=np.array(sorted(enc.state_dict())))
    out["z.rows"], out["z.cols"] = marginals(z.detach().numpy())
    out["g_x.rows"], out["g_x.cols"] = marginals(x.grad.numpy())
    out["z.sample"] = z.detach().numpy()[[0, 5, 11, 1234, 2999]]
    out["g_x.sample"] = x.grad.numpy()[[0, 5, 11, 1234, 2999]]
    for k, p in enc.named_parameters():
        if p.grad is not None:
            out["grad." + k + ".rows"], out["grad." + k + ".cols"] = marginals(p.grad.numpy().reshape(p.shape[0], -1))
    for k in ("norms.0.running_mean", "norms.0.running_var", "norms.1.running_mean", "norms.1.running_var"):
        out["after." + k] = enc.state_dict()[k].numpy()
    np.savez_compressed(os.path.join(OUT, "enc_c2.npz"), **out)
    print("enc_c2", z.shape, float(z.abs().mean()))


def gen_kmeans(name, cfg):
    """First training forward of a kmeans_init module: init_embed_ runs the reference's own kmeans()."""
    torch.manual_seed(0)
    vq = VectorQuantize(**cfg)
    vq.train()
    vq._codebook.sample_fn = strided_sample_fn
    x = torch.randn(300, cfg["dim"]) + 2.0 * torch.randn(6, cfg["dim"])[torch.randint(0, 6, (300,))]
    before = {k: v.detach().clone().numpy() for k, v in vq.state_dict().items()}
    q, ind, loss, oq = vq(x)
    out = dict(x=x.numpy(), quantize=q.detach().numpy(), embed_ind=ind.numpy(), loss=loss.detach().numpy())
    for k, v in before.items():
        out["before." + k] = v
    for k, v in vq.state_dict().items():
        out["after." + k] = v.detach().numpy()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(name, ind.shape, float(vq._codebook.initted))


def gen_enc(name, c):
    torch.manual_seed(3)
    n, e = 96, 400
    enc = Encoder(c["in_dim"], c["hid"], nn.ReLU, c["layers"], backbone=c["backbone"],
                  normalize=c["normalize"], dropout=0.0, moe=c.get("moe", False),
                  num_experts=3, moe_layers="all")
    enc.train(not c.get("moe", False))  # MoE: eval branch (softmax gate) is deterministic
    x = torch.randn(n, c["in_dim"], requires_grad=True)
    ei = torch.randint(0, n, (2, e))
    ei[:, :8] = ei[:, 8:16]          # duplicate edges
    ei[1, 16:24] = ei[0, 16:24]      # self loops
    ei[1][ei[1] == 5] = 6            # node 5 has no in-edges
    ei[0][ei[0] == 7] = 8            # node 7 has no out-edges
    xe = None
    if c["xe"]:
        table = torch.randn(3, c["in_dim"])
        xe = table[torch.randint(0, 3, (e,))]
    before = {k: v.detach().clone().numpy() for k, v in enc.state_dict().items()}
    z = enc(x, ei, xe)
    g = torch.randn(z.shape, generator=torch.Generator().manual_seed(4))
    (z * g).sum().backward()
    out = dict(x=x.detach().numpy(), edge_index=ei.numpy(), z=z.detach().numpy(), g_z=g.numpy(),
               g_x=x.grad.numpy())
    if xe is not None:
        out["edge_attr"] = xe.numpy()
    for k, v in before.items():
        out["state." + k] = v
    for k, p in enc.named_parameters():
        if p.grad is not None:
            out["grad." + k] = p.grad.numpy()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(name, z.shape, len(before), "state tensors")


def gen_dec(name, c):
    torch.manual_seed(5)
    dec = InnerProductDecoder(c["dim"], c["proj"]) if c["proj"] else InnerProductDecoder()
    z = torch.randn(c["n"], c["dim"], requires_grad=True)
    ei = torch.randint(0, c["n"], (2, c["e"]))
    ei[:, :6] = ei[:, 6:12]           # duplicate edges
    ei[1, 12:18] = ei[0, 12:18]       # self loops
    out = dec(z, ei, sigmoid=c["sigmoid"])
    g = torch.randn(out.shape, generator=torch.Generator().manual_seed(6))
    (out * g).sum().backward()
    res = dict(z=z.detach().numpy(), edge_index=ei.numpy(), out=out.detach().numpy(), g_out=g.numpy(),
               g_z=z.grad.numpy())
    for k, v in dec.state_dict().items():
        res["state." + k] = v.detach().numpy()
    for k, p in dec.named_parameters():
        res["grad." + k] = p.grad.numpy()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **res)
    print(name, out.shape)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "kmeans":   # only the k-means fixtures (added later)
        for k, v in KMEANS_CASES.items():
            gen_kmeans(k, v)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "vq_c2":    # only the configs[1]-shaped VQ fixtures (added later)
        for k, v in VQ_C2_CASES.items():
            gen_vq_c2(k, v)
        gen_vq_shipped_full()
        gen_enc_c2()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "dec":      # only the decoder fixtures (added later)
        for k, v in DEC_CASES.items():
            gen_dec(k, v)
        sys.exit(0)
    for k, v in VQ_CASES.items():
        gen_vq(k, v)
    for k, v in ENC_CASES.items():
        gen_enc(k, v)
    for k, v in DEC_CASES.items():
        gen_dec(k, v)
    for k, v in KMEANS_CASES.items():
        gen_kmeans(k, v)
    for k, v in VQ_C2_CASES.items():
        gen_vq_c2(k, v)
    gen_vq_shipped_full()
    gen_enc_c2()
