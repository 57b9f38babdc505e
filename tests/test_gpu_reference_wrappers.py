"""GPU: "pretrain.py and finetune.py run unchanged" as a test.  The reference's OWN wrappers --
model/pt_model.py PretrainModel and model/ft_model.py TaskModel, unmodified, loaded from the checkout or from
the archive oracle/make_ref.py packed into oracle/_ref/ -- are built exactly as pretrain.py:91-130 /
finetune.py:117-146,199-203 build them, once over the B200 drop-in modules (CUDA) and once over the
reference's own Encoder / VectorQuantize / InnerProductDecoder (CPU; encoder.py on oracle/pyg_shim), with the
same initial state_dict, and stepped side by side through the loop body of pretrain.py:28-66.

torch_geometric is not installable here, so the one PyG function the wrappers call,
`torch_geometric.utils.negative_sampling` (pt_model.py:6,60), is served by oracle/pyg_shim, whose hook lets
the CUDA arm draw with the library's K12 `negative_sampling` and the CPU arm replay the same pairs."""
import copy

import pytest
import torch
import torch.nn as nn

import stemgnn_b200 as S
from oracle import ref_loader

pytestmark = pytest.mark.gpu
DEV = "cuda"

if ref_loader.source()[0] is None:
    pytest.skip("the reference's model/ package is not on this machine (oracle/_ref archive absent)",
                allow_module_level=True)
REF = ref_loader.load()
import torch_geometric.utils as pyg_utils  # noqa: E402  (oracle/pyg_shim unless the real package is importable)

HID, CODES, HEADS = 64, 16, 4
PARAMS = dict(feat_lambda=100, topo_lambda=0.01, topo_sem_lambda=100, sem_lambda=1, sem_encoder_decay=0.99,
              topo_recon_ratio=0.5)          # config/pretrain.yaml (ratio 0.1 there; 0.5 keeps enough edges here)


def _vq(cls, kmeans_init=False):
    return cls(dim=HID, codebook_size=CODES, codebook_dim=HID, heads=HEADS, separate_codebook_per_head=True,
               decay=0.8, commitment_weight=10, use_cosine_sim=True, orthogonal_reg_weight=1,
               orthogonal_reg_max_codes=32, orthogonal_reg_active_codes_only=False, kmeans_init=kmeans_init,
               ema_update=False)            # pretrain.py:105-119 (max_codes >= codebook_size: no random subset)


def _pretrain_model(enc_cls, vq_cls, dec_cls):
    enc = enc_cls(input_dim=HID, hidden_dim=HID, activation=nn.ReLU, num_layers=2, backbone="sage",
                  normalize="batch", dropout=0.0, moe=False, num_experts=3, tau=1.0, moe_layers="none")
    return REF.pt_model.PretrainModel(
        encoder=enc, vq=_vq(vq_cls), feat_recon_decoder=nn.Linear(HID, HID),
        topo_recon_decoder=dec_cls(hidden_dim=HID, output_dim=HID),
        topo_sem_recon_decoder=nn.Linear(HID * 2, HID))


def _graph(seed):
    g = torch.Generator().manual_seed(seed)
    n, e = 600, 4000
    x = torch.randn(n, HID, generator=g)
    ei = torch.randint(0, n, (2, e), generator=g)
    ei = ei[:, ei[0] != ei[1]]
    xe = torch.randn(ei.size(1), HID, generator=g)
    keep = torch.rand(ei.size(1), generator=g) >= 0.2
    aug_x = x * (torch.rand(1, HID, generator=g) >= 0.2)
    return (x, ei, xe), (aug_x, ei[:, keep].contiguous(), xe[keep].contiguous())


def _loop_body(model, opt, aug_g, g, bs):
    """pretrain.py:47-66"""
    z, quantize, indices, losses = model(aug_g, g, PARAMS["topo_recon_ratio"], bs=bs, no_codebook=False)
    loss = (PARAMS["feat_lambda"] * losses["feat_recon_loss"] + PARAMS["topo_lambda"] * losses["topo_recon_loss"]
            + PARAMS["topo_sem_lambda"] * losses["topo_sem_recon_loss"] + PARAMS["sem_lambda"] * losses["sem_recon_loss"]
            + losses["commit_loss"] + 0.0 * losses["env_reg_loss"])
    opt.zero_grad()
    loss.backward()
    nn.utils.clip_grad_norm_(model.parameters(), 1.0)
    opt.step()
    model.ema_update_sem_encoder(decay=PARAMS["sem_encoder_decay"])
    return {k: float(v.detach().sum()) for k, v in losses.items()}, float(loss.detach().sum()), indices


def test_reference_pretrain_model_runs_on_the_drop_in_modules():
    torch.manual_seed(0)
    ref_model = _pretrain_model(REF.encoder.Encoder, REF.vq.VectorQuantize, REF.encoder.InnerProductDecoder)
    ours = _pretrain_model(S.Encoder, S.VectorQuantize, S.InnerProductDecoder)
    assert set(ours.state_dict()) == set(ref_model.state_dict())          # incl. the deepcopy'd sem_encoder
    ours.load_state_dict(ref_model.state_dict())
    ours.to(DEV).train()
    ref_model.train()
    opt_o = torch.optim.AdamW(ours.parameters(), lr=1e-3, weight_decay=1e-5)
    opt_r = torch.optim.AdamW(ref_model.parameters(), lr=1e-3, weight_decay=1e-5)
    drawn = []
    try:
        for step in range(3):
            g, aug = _graph(10 + step)
            pyg_utils.NEGATIVE_SAMPLING_HOOK = lambda ei, n, k=None: drawn.append(S.negative_sampling(ei, n, k)) or drawn[-1]
            torch.manual_seed(100 + step)                                  # the randperm draws of pt_model.py:53,73
            lo, total_o, ind_o = _loop_body(ours, opt_o, [t.to(DEV) for t in aug], [t.to(DEV) for t in g], 256)
            assert drawn and drawn[-1].is_cuda and drawn[-1].size(0) == 2
            replay = [t.cpu() for t in drawn]
            pyg_utils.NEGATIVE_SAMPLING_HOOK = lambda ei, n, k=None: replay.pop(0)
            torch.manual_seed(100 + step)
            lr_, total_r, ind_r = _loop_body(ref_model, opt_r, list(aug), list(g), 256)
            drawn.clear()
            assert tuple(ind_o.shape) == tuple(ind_r.shape) and ind_o.dtype == ind_r.dtype
            assert float((ind_o.cpu() != ind_r).float().mean()) <= 0.01
            for k in lr_:
                assert abs(lo[k] - lr_[k]) <= 1e-4 * max(1.0, abs(lr_[k])), (step, k, lo[k], lr_[k])
            assert abs(total_o - total_r) <= 1e-4 * max(1.0, abs(total_r)), (step, total_o, total_r)
    finally:
        pyg_utils.NEGATIVE_SAMPLING_HOOK = None
    for (k, p), (_, pr) in zip(ours.named_parameters(), ref_model.named_parameters()):
        if k.endswith("lin_l.bias"):
            continue        # a bias in front of BatchNorm: its true gradient is zero, AdamW normalises rounding noise
        assert float((p.detach().cpu() - pr.detach()).abs().max()) <= 5e-3 * (float(pr.detach().abs().max()) + 1e-3), k


@pytest.mark.parametrize("separate", [True, False])
def test_reference_task_model_runs_on_the_drop_in_modules(separate):
    """ft_model.py TaskModel as finetune.py:199-203 builds it (deepcopy of encoder and VQ, VQ frozen,
    finetune.py:178-180): logits, loss and the decoder / encoder gradients of one finetune step."""
    torch.manual_seed(1)
    enc_r = REF.encoder.Encoder(input_dim=HID, hidden_dim=HID, activation=nn.ReLU, num_layers=2, backbone="sage",
                                normalize="none", dropout=0.0)
    vq_r = _vq(REF.vq.VectorQuantize)
    enc_o = S.Encoder(input_dim=HID, hidden_dim=HID, activation=nn.ReLU, num_layers=2, backbone="sage",
                      normalize="none", dropout=0.0)
    vq_o = _vq(S.VectorQuantize)
    enc_o.load_state_dict(enc_r.state_dict())
    vq_o.load_state_dict(vq_r.state_dict())
    for p in list(vq_r.parameters()) + list(vq_o.parameters()):
        p.requires_grad = False                                            # utils/others.py freeze_params
    params = dict(separate_decoder_for_each_head=separate, decoder_jac_coeff=0.0, use_vq=1)
    tm_r = REF.ft_model.TaskModel(copy.deepcopy(enc_r), copy.deepcopy(vq_r), 7, params)
    tm_o = REF.ft_model.TaskModel(copy.deepcopy(enc_o), copy.deepcopy(vq_o), 7, params)
    tm_o.load_state_dict(tm_r.state_dict())
    tm_o.to(DEV).train()
    tm_r.train()
    (x, ei, xe), _ = _graph(5)
    y = torch.randint(0, 7, (x.size(0),), generator=torch.Generator().manual_seed(6))
    z_r = tm_r.encode(x, ei, xe)
    loss_r = tm_r.compute_activation_loss(z_r, y, task="single")
    loss_r.backward()
    z_o = tm_o.encode(x.to(DEV), ei.to(DEV), xe.to(DEV))
    loss_o = tm_o.compute_activation_loss(z_o, y.to(DEV), task="single")
    loss_o.backward()
    assert float((z_o.detach().cpu() - z_r.detach()).abs().max()) <= 1e-5 * float(z_r.detach().abs().max())
    assert abs(float(loss_o.detach()) - float(loss_r.detach())) <= 1e-5 * max(1.0, abs(float(loss_r.detach())))
    logits_o, logits_r = tm_o(x.to(DEV), ei.to(DEV), xe.to(DEV)), tm_r(x, ei, xe)
    assert tuple(logits_o.shape) == tuple(logits_r.shape)
    assert float((logits_o.detach().cpu() - logits_r.detach()).abs().max()) <= 2e-5 * float(logits_r.detach().abs().max())
    for (k, p), (_, pr) in zip(tm_o.named_parameters(), tm_r.named_parameters()):
        assert (p.grad is None) == (pr.grad is None), k
        if pr.grad is not None:
            assert float((p.grad.cpu() - pr.grad).abs().max()) <= 1e-4 * (float(pr.grad.abs().max()) + 1e-9), k
    pooled_o = tm_o.encode_graph(x.to(DEV), ei.to(DEV), xe.to(DEV),
                                 batch=torch.arange(x.size(0), device=DEV) // 100, pool="mean")
    pooled_r = tm_r.encode_graph(x, ei, xe, batch=torch.arange(x.size(0)) // 100, pool="mean")
    assert float((pooled_o.detach().cpu() - pooled_r.detach()).abs().max()) <= 1e-5 * float(pooled_r.detach().abs().max())
