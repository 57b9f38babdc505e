"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference VectorQuantize forward
(/root/reference/STEM-GNN/model/vq.py:849-1064) for 2-D inputs (N, dim), with the Euclidean
(vq.py:438-514) and cosine (vq.py:623-688) codebooks, deterministic argmax sampling
(vq.py:79-85), straight-through (vq.py:937), MSE commitment loss (vq.py:1007-1009), optional
orthogonal regularisation (vq.py:232-237,1011-1028) and the EMA update (vq.py:492-506 /
665-681).  Pinned against the reference's own vq.py through tests/golden/vq_*.npz.

State is passed explicitly as a dict of tensors with the reference's buffer names
(`embed`, `embed_avg`, `cluster_size`; shapes (Hc,K,D),(Hc,K,D),(Hc,K) with Hc = heads if
separate_codebook_per_head else 1).
"""
import torch
import torch.nn.functional as F


def l2norm(t):
    return F.normalize(t, p=2, dim=-1)                                   # vq.py:28-29


def cdist_ref(x, y):
    """vq.py:31-35: sqrt(clamp(x2 + y2 - 2xy, 0)) for (h,n,d),(h,c,d)."""
    x2 = (x ** 2).sum(-1)
    y2 = (y ** 2).sum(-1)
    xy = torch.einsum("bid,bjd->bij", x, y) * -2
    return (x2.unsqueeze(-1) + y2.unsqueeze(1) + xy).clamp(min=0).sqrt()


def codebook_forward_ref(flatten, state, *, cosine, training, ema_update, decay, eps,
                         freeze_codebook=False, learnable=False):
    """{Euclidean,CosineSim}Codebook.forward on flatten (Hc, M, D).  Returns (quantize, ind,
    dist) and updates `state` in place when the EMA branch is active."""
    embed = state["embed"] if learnable else state["embed"].detach()
    K = embed.shape[1]
    if cosine:
        dist = torch.einsum("hnd,hcd->hnc", flatten, embed)               # vq.py:650
    else:
        dist = -cdist_ref(flatten, embed)                                 # vq.py:472
    ind = dist.argmax(dim=-1)                                             # vq.py:79 (first max)
    onehot = F.one_hot(ind, K).type(flatten.dtype)                        # vq.py:80
    if training:
        quantize = torch.einsum("hnc,hcd->hnd", onehot, embed)            # vq.py:480/657
    else:
        quantize = torch.gather(embed, 1, ind.unsqueeze(-1).expand(-1, -1, embed.shape[-1]))
    if training and ema_update and not freeze_codebook:
        with torch.no_grad():
            cs = onehot.sum(dim=1)                                        # vq.py:492/665
            state["cluster_size"].lerp_(cs, 1 - decay)                    # vq.py:495 -> 40-46
            es = torch.einsum("hnd,hnc->hcd", flatten.detach(), onehot)   # vq.py:497/670
            state["embed_avg"].lerp_(es.contiguous(), 1 - decay)          # vq.py:501/674
            c = state["cluster_size"]
            sm = (c + eps) / (c.sum(-1, keepdim=True) + K * eps) * c.sum(-1, keepdim=True)
            new = state["embed_avg"] / sm.unsqueeze(-1)                   # vq.py:503-505
            if cosine:
                new = l2norm(l2norm(new))                                 # vq.py:679-681
            state["embed"].data.copy_(new)                                # vq.py:506/681
    return quantize, ind, dist


def orthogonal_loss_ref(t):
    h, n = t.shape[:2]                                                    # vq.py:232-237
    nc = l2norm(t)
    cs = torch.einsum("hid,hjd->hij", nc, nc)
    return (cs ** 2).sum() / (h * n ** 2) - (1 / n)


def vq_forward_ref(x, state, *, heads=1, separate_codebook_per_head=False, cosine=False,
                   training=True, ema_update=True, decay=0.8, eps=1e-5,
                   commitment_weight=1.0, freeze_codebook=False, learnable_codebook=False,
                   orthogonal_reg_weight=0.0, orthogonal_reg_max_codes=None,
                   project_in=None, project_out=None):
    """VectorQuantize.forward for x of shape (N, dim).  Returns (quantize, embed_ind, loss,
    orig_quantize) exactly as vq.py:1064.  `learnable_codebook` is the VectorQuantize-level
    flag (vq.py:747); the codebook's embed is a Parameter also when the orthogonal loss is on
    (vq.py:785)."""
    N = x.shape[0]
    xi = x.unsqueeze(1)                                                   # 'b d -> b 1 d' :863
    if project_in is not None:
        xi = project_in(xi)                                               # :881
    if heads > 1:
        D = xi.shape[-1] // heads
        if separate_codebook_per_head:
            xi = xi.view(N, 1, heads, D).permute(2, 0, 1, 3)              # 'b n (h d) -> h b n d'
        else:
            xi = xi.view(N, 1, heads, D).permute(0, 2, 1, 3).reshape(1, N * heads, 1, D)
    if cosine:
        xi = l2norm(xi)                                                   # :891
    xin = xi.float()
    h4 = xin if xin.dim() == 4 else xin.unsqueeze(0)                      # vq.py:451-452
    Hc, D = h4.shape[0], h4.shape[-1]
    flatten = h4.reshape(Hc, -1, D)
    cb_learn = learnable_codebook or orthogonal_reg_weight > 0
    q, ind, _ = codebook_forward_ref(flatten, state, cosine=cosine, training=training,
                                     ema_update=ema_update, decay=decay, eps=eps,
                                     freeze_codebook=freeze_codebook, learnable=cb_learn)
    q = q.reshape(h4.shape)
    ind = ind.reshape(h4.shape[:-1])
    if xin.dim() != 4:
        q, ind = q[0], ind[0]
    xq = xin
    loss = torch.zeros(1, requires_grad=training)                        # :983
    if training:
        commit_q = q if (learnable_codebook and not freeze_codebook) else q.detach()  # :931-933
        q = xq + (q - xq).detach()                                        # :937
    if heads > 1:
        if separate_codebook_per_head:
            ind = ind.permute(1, 2, 0)                                    # 'h b n -> b n h'
        else:
            ind = ind.reshape(N, heads, 1).permute(0, 2, 1)
    ind = ind.squeeze(1)                                                  # 'b 1 ... -> b ...'
    if training:
        if commitment_weight > 0:
            loss = loss + F.mse_loss(commit_q, xq) * commitment_weight    # :1007-1009
        if orthogonal_reg_weight > 0:
            cb = state["embed"]
            n_codes = cb.shape[-2]
            if orthogonal_reg_max_codes is not None and n_codes > orthogonal_reg_max_codes:
                rand_ids = torch.randperm(n_codes)[:orthogonal_reg_max_codes]  # :1024
                cb = cb[:, rand_ids]
            loss = loss + orthogonal_loss_ref(cb) * orthogonal_reg_weight  # :1027-1028
    if heads > 1:
        if separate_codebook_per_head:
            q = q.permute(1, 2, 0, 3).reshape(N, 1, -1)                   # 'h b n d -> b n (h d)'
        else:
            q = q.reshape(N, heads, 1, D).permute(0, 2, 1, 3).reshape(N, 1, -1)
    orig_q = q
    if project_out is not None:
        q = project_out(q)                                                # :1041
    return q.squeeze(1), ind, loss, orig_q.squeeze(1)


def kmeans_ref(samples, num_clusters, num_iters, cosine, means):
    """model/vq.py:182-222 with the initial means passed in (the reference draws them with
    `sample_fn`, vq.py:192).  samples (Hc, n, D).  Returns (means, bins)."""
    Hc, n, D = samples.shape
    bins = None
    for _ in range(num_iters):
        dists = samples @ means.transpose(1, 2) if cosine else -cdist_ref(samples, means)   # :195-198
        buckets = dists.argmax(dim=-1)                                                     # :200
        bins = torch.zeros(Hc, num_clusters, dtype=buckets.dtype)
        bins.scatter_add_(-1, buckets, torch.ones_like(buckets))                            # :201
        zero = bins == 0
        new = torch.zeros(Hc, num_clusters, D, dtype=samples.dtype)
        new.scatter_add_(1, buckets.unsqueeze(-1).expand(-1, -1, D), samples)               # :209
        new = new / bins.masked_fill(zero, 1).unsqueeze(-1)                                 # :210
        if cosine:
            new = l2norm(new)                                                               # :214
        means = torch.where(zero.unsqueeze(-1), means, new)                                 # :216-220
    return means, bins


def kmeans_init_ref(flatten, state, num_iters, cosine, means):
    """{Euclidean,CosineSim}Codebook.init_embed_ (vq.py:320-342 / 573-595): k-means on the first batch,
    embed = means, embed_avg = means * cluster_size, cluster_size = bins."""
    K = state["embed"].shape[1]
    embed, bins = kmeans_ref(flatten, K, num_iters, cosine, means)
    bins = bins.to(embed.dtype)
    state["embed"].data.copy_(embed)
    state["embed_avg"].data.copy_(embed * bins.unsqueeze(-1))
    state["cluster_size"].data.copy_(bins)
