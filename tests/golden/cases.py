"""Case tables shared by gen_golden.py (build container) and the tests."""
VQ_CASES = {
    # north_star configuration: Euclidean, 1 head, EMA decay 0.8
    "vq_euclid_ema": dict(dim=32, codebook_size=24, decay=0.8, commitment_weight=1.0),
    "vq_euclid_eval": dict(dim=32, codebook_size=24, _eval=True),
    "vq_cosine_ema": dict(dim=32, codebook_size=24, use_cosine_sim=True, decay=0.9),
    # shipped configuration (pretrain.py:105-119), scaled down
    "vq_shipped": dict(dim=32, codebook_size=16, codebook_dim=32, heads=4,
                       separate_codebook_per_head=True, decay=0.8, commitment_weight=10,
                       use_cosine_sim=True, orthogonal_reg_weight=1, orthogonal_reg_max_codes=8,
                       orthogonal_reg_active_codes_only=False, kmeans_init=False,
                       ema_update=False),
    "vq_shared_heads": dict(dim=32, codebook_size=16, codebook_dim=8, heads=4,
                            separate_codebook_per_head=False),
    "vq_learnable": dict(dim=32, codebook_size=16, ema_update=False, learnable_codebook=True,
                         commitment_weight=0.25),
    # exact duplicate code rows -> ties -> lower index must win (vq.py:79)
    "vq_ties": dict(dim=16, codebook_size=8, _dup=True),
}



ENC_CASES = {
    "enc_sage": dict(backbone="sage", normalize="batch", in_dim=20, hid=24, layers=2, xe=False),
    "enc_sage_xe": dict(backbone="sage", normalize="none", in_dim=24, hid=24, layers=2, xe=True),
    "enc_gcn": dict(backbone="gcn", normalize="none", in_dim=20, hid=24, layers=2, xe=False),
    "enc_gin": dict(backbone="gin", normalize="batch", in_dim=20, hid=24, layers=3, xe=False),
    "enc_moe": dict(backbone="sage", normalize="none", in_dim=24, hid=24, layers=2, xe=False,
                    moe=True),
}


# InnerProductDecoder (model/encoder.py:336-380): the topology-reconstruction decode on edges
DEC_CASES = {
    "dec_inner_sigmoid": dict(n=80, e=300, dim=24, proj=None, sigmoid=True),
    "dec_inner_raw": dict(n=80, e=300, dim=20, proj=None, sigmoid=False),
    "dec_inner_proj": dict(n=64, e=257, dim=24, proj=16, sigmoid=True),
}


# k-means initialisation (model/vq.py:182-222, 320-342) with the initial means fixed by a deterministic
# sample_fn (rows 0, 7, 14, ... of every codebook's samples) on both sides
KMEANS_CASES = {
    "km_euclid": dict(dim=16, codebook_size=12, kmeans_init=True, kmeans_iters=10),
    "km_cosine": dict(dim=16, codebook_size=12, kmeans_init=True, kmeans_iters=6, use_cosine_sim=True),
    "km_heads": dict(dim=16, codebook_size=8, codebook_dim=8, heads=2, separate_codebook_per_head=True,
                     kmeans_init=True, kmeans_iters=5, use_cosine_sim=True),
}


def strided_sample_fn(samples, num):
    """Deterministic stand-in for batched_sample_vectors (model/vq.py:115-116)."""
    import torch
    n = samples.shape[1]
    idx = (torch.arange(num, device=samples.device) * 7) % n
    return samples[:, idx]


# BASELINE configs[1] (ogbn-arxiv-shaped) at its real codebook shape -- K = 512, D = 256, one head, no
# projection -- in both variants SURVEY.md 8 names.  The fixtures are COMPACT (indices, loss, marginal sums
# of every float output): inputs and the initial codebook are regenerated from seeds by c2_inputs(), whose
# result is checked against checksums stored at generation time.
VQ_C2_CASES = {
    "vq_c2_euclid_ema": dict(dim=256, codebook_size=512, decay=0.8, commitment_weight=1.0),
    "vq_c2_cosine_learnable": dict(dim=256, codebook_size=512, use_cosine_sim=True, ema_update=False,
                                   learnable_codebook=True, commitment_weight=0.25),
}
VQ_C2_ROWS = 1536


def c2_inputs(cfg):
    """(x, initial codebook state) exactly as the reference module holds them after
    `torch.manual_seed(0); VectorQuantize(**cfg)` (model/vq.py:54-57 uniform_init = kaiming_uniform_ on an
    empty [1, K, D]; l2-normalised for cosine, vq.py:542; embed_avg = embed.clone(), cluster_size zeros for
    Euclid :291-293 / zeros for cosine :562) followed by `x = torch.randn(rows, dim)`."""
    import torch
    import torch.nn.functional as F
    torch.manual_seed(0)
    t = torch.empty(1, cfg["codebook_size"], cfg["dim"])
    torch.nn.init.kaiming_uniform_(t)
    if cfg.get("use_cosine_sim"):
        t = F.normalize(t, p=2, dim=-1)
    state = dict(embed=t.clone(), embed_avg=t.clone(), cluster_size=torch.zeros(1, cfg["codebook_size"]))
    x = torch.randn(VQ_C2_ROWS, cfg["dim"])
    return x, state


def marginals(a):
    """Row and column sums in float64: a compact stand-in for a [rows, cols] float matrix."""
    import numpy as np
    a = np.asarray(a, dtype=np.float64).reshape(-1, a.shape[-1])
    return a.sum(1), a.sum(0)


# The shipped configuration at its real shape (pretrain.py:105-119 / finetune.py:132-146: cosine, 4 separate
# heads x 128 codes x 768, 768 -> 3072 -> 768 projections, learnable codebook through the orthogonal loss),
# which BASELINE configs[0] and configs[2] use.  Compact fixture like the C2 ones.
VQ_SHIPPED_FULL = dict(dim=768, codebook_size=128, codebook_dim=768, heads=4, separate_codebook_per_head=True,
                       decay=0.8, commitment_weight=10, use_cosine_sim=True, orthogonal_reg_weight=1,
                       orthogonal_reg_max_codes=32, orthogonal_reg_active_codes_only=False, kmeans_init=False,
                       ema_update=False)
VQ_SHIPPED_ROWS = 384


def shipped_inputs():
    """(x, project_in, project_out, codebook state) as `torch.manual_seed(0); VectorQuantize(**VQ_SHIPPED_FULL)`
    leaves them (model/vq.py:738-739 the two nn.Linear, then the codebook's l2norm(uniform_init) :542),
    followed by `x = torch.randn(rows, 768)`."""
    import torch
    import torch.nn as nn
    import torch.nn.functional as F
    c = VQ_SHIPPED_FULL
    torch.manual_seed(0)
    inner = c["codebook_dim"] * c["heads"]
    pin, pout = nn.Linear(c["dim"], inner), nn.Linear(inner, c["dim"])
    t = torch.empty(c["heads"], c["codebook_size"], c["codebook_dim"])
    nn.init.kaiming_uniform_(t)
    t = F.normalize(t, p=2, dim=-1)
    state = dict(embed=t.clone(), embed_avg=t.clone(), cluster_size=torch.zeros(c["heads"], c["codebook_size"]))
    x = torch.randn(VQ_SHIPPED_ROWS, c["dim"])
    return x, pin, pout, state


# The encoder at BASELINE configs[1]'s widths (2-layer SAGE 128 -> 256 -> 256, BatchNorm, dropout off for
# determinism) on a 3,000-node / 21,000-edge graph with a 2,000-edge hub row.  Parameters are OVERWRITTEN with
# seeded values in state_dict order on both sides (seeded_parameters), so the fixture carries outputs only.
ENC_C2 = dict(backbone="sage", normalize="batch", in_dim=128, hid=256, layers=2, n=3000, e=21000)


def enc_c2_inputs():
    import torch
    g = torch.Generator().manual_seed(31)
    c = ENC_C2
    x = torch.randn(c["n"], c["in_dim"], generator=g)
    ei = torch.randint(0, c["n"], (2, c["e"]), generator=g)
    ei[1, :2000] = 11                    # hub destination
    ei[1][ei[1] == 5] = 6                # node 5 has no in-edge
    gz = torch.randn(c["n"], c["hid"], generator=g)
    return x, ei, gz


def seeded_parameters(module, seed=32):
    """Overwrites every floating-point parameter / buffer of `module` in state_dict order with seeded values
    (weights ~ 0.05 N(0,1); running_var kept positive); returns the module."""
    import torch
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for k, v in module.state_dict().items():
            if not v.dtype.is_floating_point:
                continue
            r = torch.randn(v.shape, generator=g) * 0.05
            v.copy_(r.abs() + 0.5 if k.endswith("running_var") or k.endswith("norms.0.weight") or k.endswith("norms.1.weight") else r)
    return module
