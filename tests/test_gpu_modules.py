"""GPU: the drop-in modules against the golden vectors produced by the reference's own code
(tests/golden/gen_golden.py) -- Encoder on every backbone, VectorQuantize in every configuration."""
import os

import numpy as np
import pytest
import torch
import torch.nn as nn

import stemgnn_b200 as S
from tests.golden.cases import ENC_CASES, VQ_CASES

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _load(golden_dir, name):
    return dict(np.load(os.path.join(golden_dir, name + ".npz")))


@pytest.mark.parametrize("name", sorted(ENC_CASES))
def test_encoder_matches_reference_golden(golden_dir, name):
    c, g = ENC_CASES[name], _load(golden_dir, name)
    enc = S.Encoder(c["in_dim"], c["hid"], nn.ReLU, c["layers"], backbone=c["backbone"],
                    normalize=c["normalize"], dropout=0.0, moe=c.get("moe", False), num_experts=3,
                    moe_layers="all")
    sd = {k[len("state."):]: torch.from_numpy(v.copy()) for k, v in g.items() if k.startswith("state.")}
    assert set(sd) == set(enc.state_dict())
    enc.load_state_dict(sd)
    enc.to(DEV).train(not c.get("moe", False))
    x = torch.from_numpy(g["x"].copy()).to(DEV).requires_grad_(True)
    ei = torch.from_numpy(g["edge_index"]).to(DEV)
    xe = torch.from_numpy(g["edge_attr"]).to(DEV) if "edge_attr" in g else None
    z = enc(x, ei, xe)
    (z * torch.from_numpy(g["g_z"]).to(DEV)).sum().backward()
    # dense layers run on cuBLAS here vs CPU GEMM in the golden run: 1e-5 relative (north_star)
    np.testing.assert_allclose(z.detach().cpu().numpy(), g["z"], rtol=1e-5, atol=2e-5)
    np.testing.assert_allclose(x.grad.cpu().numpy(), g["g_x"], rtol=1e-5, atol=2e-5)
    for k, p in enc.named_parameters():
        if "grad." + k in g:
            np.testing.assert_allclose(p.grad.cpu().numpy(), g["grad." + k], rtol=1e-4, atol=5e-5)


@pytest.mark.parametrize("name", sorted(VQ_CASES))
def test_vq_matches_reference_golden(golden_dir, name):
    cfg = dict(VQ_CASES[name])
    g = _load(golden_dir, name)
    training = not cfg.pop("_eval", False)
    cfg.pop("_dup", None)
    vq = S.VectorQuantize(**cfg)
    sd = {k[len("before."):]: torch.from_numpy(v.copy()) for k, v in g.items() if k.startswith("before.")}
    assert set(sd) == set(vq.state_dict())
    vq.load_state_dict(sd)
    vq.to(DEV).train(training)
    x = torch.from_numpy(g["x"].copy()).to(DEV).requires_grad_(training)
    torch.manual_seed(1)
    q, ind, loss, oq = vq(x)
    assert ind.dtype == torch.int64 and tuple(ind.shape) == g["embed_ind"].shape
    assert tuple(loss.shape) == (1,)
    assert np.array_equal(ind.cpu().numpy(), g["embed_ind"])
    np.testing.assert_allclose(q.detach().cpu().numpy(), g["quantize"], rtol=1e-5, atol=2e-6)
    np.testing.assert_allclose(oq.detach().cpu().numpy(), g["orig_quantize"], rtol=1e-5, atol=2e-6)
    if not cfg.get("orthogonal_reg_weight"):
        # (the orthogonal term draws randperm on the device: different ids than the CPU golden run)
        np.testing.assert_allclose(loss.detach().cpu().numpy(), g["loss"], rtol=1e-5, atol=1e-7)
    for k in ("embed", "embed_avg", "cluster_size"):
        np.testing.assert_allclose(getattr(vq._codebook, k).detach().cpu().numpy(),
                                   g["after._codebook." + k], rtol=1e-5, atol=2e-6)
    if training:
        ((q * torch.from_numpy(g["g_quantize"]).to(DEV)).sum() + loss.sum()).backward()
        np.testing.assert_allclose(x.grad.cpu().numpy(), g["g_x"], rtol=1e-5, atol=2e-6)
        if "g_embed" in g and not cfg.get("orthogonal_reg_weight"):
            np.testing.assert_allclose(vq._codebook.embed.grad.cpu().numpy(), g["g_embed"],
                                       rtol=1e-5, atol=2e-6)
    else:
        assert not loss.requires_grad and float(loss) == 0.0


from tests.golden.cases import KMEANS_CASES, strided_sample_fn  # noqa: E402


@pytest.mark.parametrize("name", sorted(KMEANS_CASES))
def test_kmeans_init_matches_reference_golden(golden_dir, name):
    """kmeans_init=True: the first training forward initialises the codebook with k-means run by K3
    (exact scan) and K4 (per-cluster sums); the fixture holds what the reference's own kmeans()
    (vq.py:182-222) produced from the same initial means."""
    cfg = dict(KMEANS_CASES[name])
    g = _load(golden_dir, name)
    vq = S.VectorQuantize(**cfg)
    sd = {k[len("before."):]: torch.from_numpy(v.copy()) for k, v in g.items() if k.startswith("before.")}
    assert set(sd) == set(vq.state_dict())
    vq.load_state_dict(sd)
    vq.to(DEV).train()
    vq._codebook.sample_fn = strided_sample_fn
    x = torch.from_numpy(g["x"].copy()).to(DEV)
    q, ind, loss, _ = vq(x)
    assert float(vq._codebook.initted) == 1.0
    assert np.array_equal(ind.cpu().numpy(), g["embed_ind"])
    np.testing.assert_allclose(q.detach().cpu().numpy(), g["quantize"], rtol=1e-5, atol=2e-6)
    np.testing.assert_allclose(loss.detach().cpu().numpy(), g["loss"], rtol=1e-5, atol=1e-7)
    for k in ("embed", "embed_avg", "cluster_size"):
        np.testing.assert_allclose(getattr(vq._codebook, k).detach().cpu().numpy(),
                                   g["after._codebook." + k], rtol=1e-5, atol=2e-6)
    q2, ind2, _, _ = vq(x)                                   # initialised: the second call does not re-run k-means
    assert float(vq._codebook.initted) == 1.0
