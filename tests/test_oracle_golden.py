"""CPU: the oracle restatements against the golden vectors produced by the reference's own
code (tests/golden/gen_golden.py), and the torch oracle against the sequential C oracle."""
import glob
import os

import numpy as np
import pytest
import torch
import torch.nn as nn

from oracle import c_agg, graph_ref
from oracle import propagate_ref as P
from oracle.encoder_ref import EncoderRef
from oracle.vq_ref import vq_forward_ref
from oracle.decoder_ref import edge_dot_grad_ref, edge_dot_ref
from tests.golden.cases import (ENC_C2, VQ_C2_CASES, VQ_SHIPPED_FULL, c2_inputs, enc_c2_inputs, marginals,  # noqa: E402
                                seeded_parameters, shipped_inputs)
from tests.golden.cases import DEC_CASES, ENC_CASES, VQ_CASES


def _load(golden_dir, name):
    return dict(np.load(os.path.join(golden_dir, name + ".npz")))


def vq_kwargs(cfg):
    """Reference ctor kwargs (vq.py:693-728) -> oracle.vq_forward_ref kwargs."""
    return dict(heads=cfg.get("heads", 1),
                separate_codebook_per_head=cfg.get("separate_codebook_per_head", False),
                cosine=cfg.get("use_cosine_sim", False), ema_update=cfg.get("ema_update", True),
                decay=cfg.get("decay", 0.8), eps=cfg.get("eps", 1e-5),
                commitment_weight=cfg.get("commitment_weight", 1.0),
                learnable_codebook=cfg.get("learnable_codebook", False),
                orthogonal_reg_weight=cfg.get("orthogonal_reg_weight", 0.0),
                orthogonal_reg_max_codes=cfg.get("orthogonal_reg_max_codes"))


@pytest.mark.parametrize("name", sorted(VQ_CASES))
def test_vq_oracle_matches_reference_golden(golden_dir, name):
    cfg = dict(VQ_CASES[name])
    g = _load(golden_dir, name)
    training = not cfg.pop("_eval", False)
    cfg.pop("_dup", None)
    kw = vq_kwargs(cfg)
    state = {k: torch.from_numpy(g["before._codebook." + k].copy())
             for k in ("embed", "embed_avg", "cluster_size")}
    if "g_embed" in g:
        state["embed"].requires_grad_(True)
    pin = pout = None
    if "before.project_in.weight" in g:
        pin = nn.Linear(*g["before.project_in.weight"].shape[::-1])
        pout = nn.Linear(*g["before.project_out.weight"].shape[::-1])
        with torch.no_grad():
            pin.weight.copy_(torch.from_numpy(g["before.project_in.weight"]))
            pin.bias.copy_(torch.from_numpy(g["before.project_in.bias"]))
            pout.weight.copy_(torch.from_numpy(g["before.project_out.weight"]))
            pout.bias.copy_(torch.from_numpy(g["before.project_out.bias"]))
    x = torch.from_numpy(g["x"].copy()).requires_grad_(training)
    torch.manual_seed(1)
    q, ind, loss, oq = vq_forward_ref(x, state, training=training, project_in=pin,
                                      project_out=pout, **kw)
    assert ind.dtype == torch.int64 and tuple(ind.shape) == g["embed_ind"].shape
    assert np.array_equal(ind.numpy(), g["embed_ind"])
    np.testing.assert_allclose(q.detach().numpy(), g["quantize"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(oq.detach().numpy(), g["orig_quantize"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(loss.detach().numpy(), g["loss"], rtol=1e-5, atol=1e-7)
    for k in ("embed", "embed_avg", "cluster_size"):
        np.testing.assert_allclose(state[k].detach().numpy(), g["after._codebook." + k],
                                   rtol=1e-5, atol=1e-6)
    if training:
        ((q * torch.from_numpy(g["g_quantize"])).sum() + loss.sum()).backward()
        np.testing.assert_allclose(x.grad.numpy(), g["g_x"], rtol=1e-5, atol=1e-6)
        if "g_embed" in g:
            np.testing.assert_allclose(state["embed"].grad.numpy(), g["g_embed"],
                                       rtol=1e-5, atol=1e-6)


def test_vq_ties_lower_index_wins(golden_dir):
    g = _load(golden_dir, "vq_ties")
    ind = g["embed_ind"]
    assert (ind[:40] == 2).all() and not np.isin(ind, [5, 7]).any()


@pytest.mark.parametrize("name", sorted(ENC_CASES))
def test_encoder_oracle_matches_reference_golden(golden_dir, name):
    c = ENC_CASES[name]
    g = _load(golden_dir, name)
    enc = EncoderRef(c["in_dim"], c["hid"], nn.ReLU, c["layers"], backbone=c["backbone"],
                     normalize=c["normalize"], dropout=0.0, moe=c.get("moe", False),
                     num_experts=3, moe_layers="all")
    sd = {k[len("state."):]: torch.from_numpy(v.copy()) for k, v in g.items()
          if k.startswith("state.")}
    assert set(sd) == set(enc.state_dict()), "state_dict keys must match the reference"
    enc.load_state_dict(sd)
    enc.train(not c.get("moe", False))
    x = torch.from_numpy(g["x"].copy()).requires_grad_(True)
    ei = torch.from_numpy(g["edge_index"])
    xe = torch.from_numpy(g["edge_attr"]) if "edge_attr" in g else None
    z = enc(x, ei, xe)
    np.testing.assert_array_equal(z.detach().numpy(), g["z"])  # same ops, same order: bit-exact
    (z * torch.from_numpy(g["g_z"])).sum().backward()
    np.testing.assert_array_equal(x.grad.numpy(), g["g_x"])
    for k, p in enc.named_parameters():
        if "grad." + k in g:
            np.testing.assert_allclose(p.grad.numpy(), g["grad." + k], rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("with_xe", [False, True])
@pytest.mark.parametrize("reduce", ["mean", "sum"])
def test_torch_oracle_is_sequential_in_edge_order(with_xe, reduce):
    """scatter_add_/index_add_ on CPU == add the messages of a row one by one in original edge
    order (the C oracle): bit-for-bit, forward and backward."""
    gen = torch.Generator().manual_seed(11)
    n, e, f = 300, 4000, 37
    x = torch.randn(n, f, generator=gen, requires_grad=True)
    ei = torch.randint(0, n, (2, e), generator=gen)
    ei[1, :600] = 3                                            # a hub row
    xe = torch.randn(e, f, generator=gen) if with_xe else None
    out = P.sage_propagate_ref(x, ei, xe, n, reduce)
    ref = c_agg.aggregate_fwd(ei.numpy(), x.detach().numpy(), n, None if xe is None else xe.numpy(),
                              act=1, reduce=int(reduce == "mean"))
    np.testing.assert_array_equal(out.detach().numpy(), ref)
    go = torch.randn(n, f, generator=gen)
    out.backward(go)
    gref = c_agg.aggregate_bwd(ei.numpy(), x.detach().numpy(), go.numpy(),
                               None if xe is None else xe.numpy(), act=1,
                               reduce=int(reduce == "mean"))
    np.testing.assert_array_equal(x.grad.numpy(), gref)


def test_gcn_oracle_vs_c():
    gen = torch.Generator().manual_seed(12)
    n, e, f = 200, 1500, 19
    xw = torch.randn(n, f, generator=gen, requires_grad=True)
    ei = torch.randint(0, n, (2, e), generator=gen)
    out = P.gcn_propagate_ref(xw, ei, n)
    ei2, w = P.gcn_norm_ref(ei, n)
    ref = c_agg.aggregate_fwd(ei2.numpy(), xw.detach().numpy(), n, None, w.numpy(), act=0, reduce=0)
    np.testing.assert_array_equal(out.detach().numpy(), ref)
    assert np.array_equal(ei2.numpy(), graph_ref.add_remaining_self_loops_ref(ei.numpy(), n))


def test_csr_ref_properties():
    rng = np.random.default_rng(0)
    n, e = 50, 400
    ei = rng.integers(0, n, (2, e))
    for by in (0, 1):
        rowptr, col, perm, deg = graph_ref.csr_build_ref(ei, n, by)
        assert rowptr[0] == 0 and rowptr[-1] == e and (np.diff(rowptr) == deg).all()
        assert sorted(perm.tolist()) == list(range(e))
        for i in range(n):
            seg = perm[rowptr[i]:rowptr[i + 1]]
            assert (ei[by][seg] == i).all() and (np.diff(seg) > 0).all()  # stable
            assert (col[rowptr[i]:rowptr[i + 1]] == ei[1 - by][seg]).all()


@pytest.mark.parametrize("name", sorted(DEC_CASES))
def test_decoder_oracle_matches_reference_golden(golden_dir, name):
    """oracle/decoder_ref.py against the reference's own InnerProductDecoder (encoder.py:336-380)."""
    c = DEC_CASES[name]
    g = _load(golden_dir, name)
    z = torch.from_numpy(g["z"]).requires_grad_(True)
    ei = torch.from_numpy(g["edge_index"])
    h = z
    if c["proj"]:
        lin = nn.Linear(c["dim"], c["proj"])
        with torch.no_grad():
            lin.weight.copy_(torch.from_numpy(g["state.lin.weight"]))
            lin.bias.copy_(torch.from_numpy(g["state.lin.bias"]))
        h = lin(z)
    out = edge_dot_ref(h, ei, c["sigmoid"])
    np.testing.assert_array_equal(out.detach().numpy(), g["out"])
    (out * torch.from_numpy(g["g_out"])).sum().backward()
    np.testing.assert_array_equal(z.grad.numpy(), g["g_z"])
    if not c["proj"]:     # the closed form used for the large GPU cases
        g64 = edge_dot_grad_ref(g["z"], g["edge_index"], g["g_out"], c["sigmoid"])
        np.testing.assert_allclose(g["g_z"], g64, rtol=1e-5, atol=1e-6)


from tests.golden.cases import KMEANS_CASES, strided_sample_fn  # noqa: E402


@pytest.mark.parametrize("name", sorted(KMEANS_CASES))
def test_kmeans_init_oracle_matches_reference_golden(golden_dir, name):
    """The reference's own kmeans() (vq.py:182-222) ran inside init_embed_ when the fixture was made;
    the restatement must reproduce the initialised codebook and the forward that follows it."""
    from oracle.vq_ref import kmeans_init_ref, l2norm
    cfg = dict(KMEANS_CASES[name])
    g = _load(golden_dir, name)
    x = torch.from_numpy(g["x"].copy())
    heads, cosine = cfg.get("heads", 1), cfg.get("use_cosine_sim", False)
    pin = pout = None
    h = x
    if "before.project_in.weight" in g:
        D = cfg["codebook_dim"] * heads
        pin, pout = nn.Linear(cfg["dim"], D), nn.Linear(D, cfg["dim"])
        pin.load_state_dict({k: torch.from_numpy(g["before.project_in." + k].copy()) for k in ("weight", "bias")})
        pout.load_state_dict({k: torch.from_numpy(g["before.project_out." + k].copy()) for k in ("weight", "bias")})
        h = pin(x).detach()
    flatten = h.view(x.size(0), heads, -1).permute(1, 0, 2).contiguous()
    if cosine:
        flatten = l2norm(flatten)
    state = {k: torch.from_numpy(g["before._codebook." + k].copy()) for k in ("embed", "embed_avg", "cluster_size")}
    assert float(g["before._codebook.initted"][0]) == 0.0 and float(g["after._codebook.initted"][0]) == 1.0
    kmeans_init_ref(flatten, state, cfg["kmeans_iters"], cosine,
                    strided_sample_fn(flatten, cfg["codebook_size"]))
    q, ind, loss, _ = vq_forward_ref(x, state, project_in=pin, project_out=pout, **vq_kwargs(cfg))
    assert np.array_equal(ind.numpy(), g["embed_ind"])
    np.testing.assert_allclose(q.detach().numpy(), g["quantize"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(loss.detach().numpy(), g["loss"], rtol=1e-5, atol=1e-7)
    for k in ("embed", "embed_avg", "cluster_size"):
        np.testing.assert_allclose(state[k].numpy(), g["after._codebook." + k], rtol=1e-5, atol=1e-6)


def test_oracle_against_an_independent_sparse_formulation():
    """The restated PyG arithmetic against scipy.sparse in float64 (a formulation that shares no code with
    the oracle): mean aggregation = D^-1 A relu(x), its adjoint for the gradient, GCN = D^-1/2 (A + I') D^-1/2.
    The graph has a 6,000-edge hub row and isolated nodes; fp32 sequential sums against float64: 1e-5 of the
    output scale (the north-star bound), which also bounds what the order of additions can change."""
    import scipy.sparse as sp
    gen = torch.Generator().manual_seed(21)
    n, e, f = 900, 20000, 24
    ei = torch.randint(0, n, (2, e), generator=gen)
    ei[1, :6000] = 17                                    # hub destination
    ei[1][ei[1] == 5] = 6                                # node 5 isolated as a destination
    x = torch.randn(n, f, generator=gen)
    go = torch.randn(n, f, generator=gen)
    src, dst = ei[0].numpy(), ei[1].numpy()
    A = sp.coo_matrix((np.ones(e), (dst, src)), shape=(n, n)).tocsr()      # duplicates are summed: multi-edges count
    deg = np.maximum(np.asarray(A.sum(1)).ravel(), 1.0)
    x64 = x.double().numpy()
    want = (A @ np.maximum(x64, 0.0)) / deg[:, None]
    xc = x.clone().requires_grad_(True)
    out = P.sage_propagate_ref(xc, ei, None, n, "mean")
    out.backward(go)
    scale = np.abs(want).max()
    assert np.abs(out.detach().double().numpy() - want).max() <= 1e-5 * scale
    assert np.all(out.detach().numpy()[5] == 0.0)
    gwant = (A.T @ (go.double().numpy() / deg[:, None])) * (x64 > 0)
    assert np.abs(xc.grad.double().numpy() - gwant).max() <= 1e-5 * np.abs(gwant).max()
    # GCN: remaining self loops (existing loops keep their weight, the others get 1), symmetric normalisation
    ei_sl, w = P.gcn_norm_ref(ei, n)
    loops = src == dst
    A2 = sp.coo_matrix((np.ones(int((~loops).sum())), (dst[~loops], src[~loops])), shape=(n, n)).tocsr()
    A2 = A2 + sp.identity(n, format="csr")               # every node ends up with exactly one loop of weight 1
    d2 = np.asarray(A2.sum(1)).ravel()
    dis = 1.0 / np.sqrt(d2)
    want_gcn = (sp.diags(dis) @ A2 @ sp.diags(dis)) @ x64
    got = P.gcn_propagate_ref(x, ei, n)
    assert np.abs(got.double().numpy() - want_gcn).max() <= 1e-5 * np.abs(want_gcn).max()
    assert ei_sl.size(1) == int((~loops).sum()) + n and abs(float(w.double().sum()) - (sp.diags(dis) @ A2 @ sp.diags(dis)).sum()) < 1e-3


@pytest.mark.parametrize("name", sorted(VQ_C2_CASES))
def test_vq_oracle_at_the_baseline_codebook_shape(golden_dir, name):
    """BASELINE configs[1]'s real codebook shape (K = 512, D = 256; Euclid + EMA 0.8 and cosine +
    learnable): the oracle against the reference's own VectorQuantize.  Compact fixture: code indices and
    loss in full, row / column sums of every float matrix; inputs regenerated from seeds and checked."""
    cfg = dict(VQ_C2_CASES[name])
    g = _load(golden_dir, name)
    x, state = c2_inputs(cfg)
    np.testing.assert_allclose([x.double().sum().item(), x.double().abs().sum().item()], g["x_check"], rtol=0, atol=0)
    np.testing.assert_allclose([state["embed"].double().sum().item(), state["embed"].double().abs().sum().item()],
                               g["embed_check"], rtol=0, atol=0)
    kw = vq_kwargs(cfg)
    if cfg.get("learnable_codebook"):
        state["embed"].requires_grad_(True)
    x.requires_grad_(True)
    q, ind, loss, _ = vq_forward_ref(x, state, training=True, **kw)
    assert np.array_equal(ind.numpy(), g["embed_ind"])
    np.testing.assert_allclose(loss.detach().numpy(), g["loss"], rtol=1e-5, atol=1e-8)
    gq = torch.randn(q.shape, generator=torch.Generator().manual_seed(2))
    ((q * gq).sum() + loss.sum()).backward()

    def close(t, key):
        rows, cols = marginals(t.detach().numpy())
        scale = max(np.abs(g[key + ".rows"]).max(), 1e-12)
        np.testing.assert_allclose(rows, g[key + ".rows"], rtol=0, atol=2e-5 * scale)
        np.testing.assert_allclose(cols, g[key + ".cols"], rtol=0, atol=2e-5 * max(np.abs(g[key + ".cols"]).max(), 1e-12))

    close(q, "quantize")
    close(x.grad, "g_x")
    close(state["embed"], "after.embed")
    close(state["embed_avg"], "after.embed_avg")
    np.testing.assert_allclose(state["cluster_size"].detach().numpy(), g["after.cluster_size"], rtol=1e-6, atol=1e-7)
    if "g_embed.rows" in g:
        close(state["embed"].grad, "g_embed")
    assert np.unique(g["embed_ind"]).size > 400            # the case exercises most of the codebook


def test_vq_oracle_in_the_shipped_configuration_at_full_shape(golden_dir):
    """The configuration the reference ships (pretrain.py:105-119): cosine, 4 separate heads x 128 codes x 768,
    768 -> 3072 -> 768 projections, orthogonal loss on 32 random codes, commitment weight 10 -- at its real
    shape, the oracle against the reference's own module (compact golden `vq_shipped_full.npz`)."""
    cfg = dict(VQ_SHIPPED_FULL)
    g = _load(golden_dir, "vq_shipped_full")
    x, pin, pout, state = shipped_inputs()
    np.testing.assert_allclose([x.double().sum().item(), x.double().abs().sum().item()], g["x_check"], rtol=0, atol=0)
    np.testing.assert_allclose([state["embed"].double().sum().item(), state["embed"].double().abs().sum().item()],
                               g["embed_check"], rtol=0, atol=0)
    np.testing.assert_allclose([pin.weight.double().sum().item(), pout.weight.double().abs().sum().item()],
                               g["pin_check"], rtol=0, atol=0)
    state["embed"].requires_grad_(True)
    x.requires_grad_(True)
    torch.manual_seed(1)
    q, ind, loss, oq = vq_forward_ref(x, state, training=True, project_in=pin, project_out=pout, **vq_kwargs(cfg))
    assert tuple(ind.shape) == (x.size(0), 4) and np.array_equal(ind.numpy(), g["embed_ind"])
    np.testing.assert_allclose(loss.detach().numpy(), g["loss"], rtol=1e-5, atol=1e-8)
    gq = torch.randn(q.shape, generator=torch.Generator().manual_seed(2))
    ((q * gq).sum() + loss.sum()).backward()
    for key, t in (("quantize", q), ("orig_quantize", oq), ("g_x", x.grad), ("g_embed", state["embed"].grad),
                   ("g_project_in", pin.weight.grad), ("g_project_out", pout.weight.grad)):
        rows, cols = marginals(t.detach().numpy())
        np.testing.assert_allclose(rows, g[key + ".rows"], rtol=0, atol=2e-5 * max(np.abs(g[key + ".rows"]).max(), 1e-12))
        np.testing.assert_allclose(cols, g[key + ".cols"], rtol=0, atol=2e-5 * max(np.abs(g[key + ".cols"]).max(), 1e-12))


def test_encoder_oracle_at_the_baseline_widths(golden_dir):
    """The reference's Encoder (model/encoder.py on oracle/pyg_shim) at BASELINE configs[1]'s widths -- 2-layer
    SAGE 128 -> 256 -> 256 with BatchNorm -- on 3,000 nodes / 21,000 edges with a 2,000-edge hub row: the oracle
    gives the same bits (sampled rows) and the same marginal sums; parameters seeded identically on both sides."""
    c = ENC_C2
    g = _load(golden_dir, "enc_c2")
    enc = EncoderRef(c["in_dim"], c["hid"], nn.ReLU, c["layers"], backbone=c["backbone"], normalize=c["normalize"],
                     dropout=0.0)
    assert sorted(enc.state_dict()) == g["keys"].tolist(), "state_dict keys (and so the seeded values) must match"
    seeded_parameters(enc).train()
    x, ei, gz = enc_c2_inputs()
    x.requires_grad_(True)
    z = enc(x, ei)
    (z * gz).sum().backward()
    rows = [0, 5, 11, 1234, 2999]
    np.testing.assert_array_equal(z.detach().numpy()[rows], g["z.sample"])          # same ops, same order
    np.testing.assert_array_equal(x.grad.numpy()[rows], g["g_x.sample"])
    for key, t in (("z", z.detach().numpy()), ("g_x", x.grad.numpy())):
        r, cl = marginals(t)
        np.testing.assert_allclose(r, g[key + ".rows"], rtol=0, atol=1e-9 * max(np.abs(g[key + ".rows"]).max(), 1))
        np.testing.assert_allclose(cl, g[key + ".cols"], rtol=0, atol=1e-9 * max(np.abs(g[key + ".cols"]).max(), 1))
    for k, p in enc.named_parameters():
        if "grad." + k + ".rows" in g:
            r, cl = marginals(p.grad.numpy().reshape(p.shape[0], -1))
            np.testing.assert_allclose(r, g["grad." + k + ".rows"], rtol=0, atol=1e-5 * max(np.abs(g["grad." + k + ".rows"]).max(), 1e-9))
            np.testing.assert_allclose(cl, g["grad." + k + ".cols"], rtol=0, atol=1e-5 * max(np.abs(g["grad." + k + ".cols"]).max(), 1e-9))
    for k in ("norms.0.running_mean", "norms.0.running_var", "norms.1.running_mean", "norms.1.running_var"):
        np.testing.assert_allclose(enc.state_dict()[k].numpy(), g["after." + k], rtol=1e-6, atol=1e-7)
