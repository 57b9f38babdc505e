"""Drop-in `VectorQuantize` (reference model/vq.py:692-1064) with the codebook search, gather,
commitment partials, per-code statistics and EMA update running in the sm_100a kernels (K3, K4).
Same constructor arguments, forward signature, return tuple and state_dict keys
(`project_in.*`, `project_out.*`, `_codebook.{initted,cluster_size,embed_avg,embed}`).
Option combinations outside the hot path raise NotImplementedError -- there is no CPU or eager
fallback."""
from typing import Callable, Optional

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F

from . import dense, ops


def _l2norm(t):
    return F.normalize(t, p=2, dim=-1)


def _kaiming_uniform(*shape):
    t = torch.empty(shape)
    nn.init.kaiming_uniform_(t)          # model/vq.py:54-57
    return t


def _batched_sample_vectors(samples, num):
    """model/vq.py:106-116: `num` rows of every codebook's samples [Hc, n, D], without replacement
    when there are enough of them."""
    out = []
    for rows in samples.unbind(dim=0):
        n = rows.size(0)
        pick = torch.randperm(n, device=rows.device)[:num] if n >= num \
            else torch.randint(0, n, (num,), device=rows.device)
        out.append(rows[pick])
    return torch.stack(out, dim=0)


def _sample_multinomial(total_count, probs):
    """model/vq.py:120-134: split `total_count` draws over the ranks in proportion to `probs` (sequential
    binomials on the CPU, like the reference)."""
    device = probs.device
    probs = probs.cpu()
    total = probs.new_full((), total_count)
    remainder = probs.new_ones(())
    sample = torch.empty_like(probs, dtype=torch.long)
    for i, p in enumerate(probs):
        s = torch.binomial(total, (p / remainder).clamp(0, 1))
        sample[i] = s
        total = total - s
        remainder = remainder - p
    return sample.to(device)


def _sample_vectors_distributed(samples, num, group=None):
    """model/vq.py:158-180 (sample_vectors_distributed): the `num` initial means are drawn from ALL ranks' rows --
    rank 0 splits `num` multinomially by the ranks' row counts, every rank draws its share locally, the shares
    are exchanged (one broadcast per rank, sizes known from the split) and concatenated in rank order, so every
    rank ends with the same [1, num, D] tensor.  One codebook only (the reference asserts the same, vq.py:285)."""
    if samples.size(0) != 1:
        raise NotImplementedError("distributed kmeans init with several codebooks (model/vq.py:285)")
    local = samples[0]
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    ranks = dist.get_process_group_ranks(group) if group is not None else list(range(world))
    size = torch.tensor([local.size(0)], dtype=torch.long, device=local.device)
    sizes = [torch.empty_like(size) for _ in range(world)]
    dist.all_gather(sizes, size, group=group)
    sizes = torch.cat(sizes)
    if rank == 0:
        per_rank = _sample_multinomial(num, sizes / sizes.sum())
    else:
        per_rank = torch.empty_like(sizes)
    dist.broadcast(per_rank, src=ranks[0], group=group)
    per_rank = per_rank.tolist()
    mine = _batched_sample_vectors(local.unsqueeze(0), per_rank[rank])[0] if per_rank[rank] > 0 \
        else local.new_empty((0, local.size(1)))
    parts = []
    for r, cnt in enumerate(per_rank):
        t = mine.contiguous() if r == rank else local.new_empty((cnt, local.size(1)))
        if cnt > 0:
            dist.broadcast(t, src=ranks[r], group=group)
        parts.append(t)
    return torch.cat(parts, dim=0).unsqueeze(0)


class _Codebook(nn.Module):
    """State holder shared by the Euclidean (model/vq.py:241-514) and cosine (model/vq.py:516-688)
    codebooks: buffers initted (1,), cluster_size (Hc,K), embed_avg (Hc,K,D) and embed (Hc,K,D;
    a Parameter iff learnable)."""
    cosine = False

    def __init__(self, dim, codebook_size, num_codebooks=1, kmeans_init=False, kmeans_iters=10,
                 decay=0.8, eps=1e-5, use_ddp=False, learnable_codebook=False, ema_update=True,
                 path=ops.PATH_AUTO, sync_kmeans=True):
        super().__init__()
        self.dim, self.codebook_size, self.num_codebooks = dim, codebook_size, num_codebooks
        self.kmeans_iters, self.decay, self.eps = kmeans_iters, decay, eps
        self.use_ddp, self.learnable_codebook, self.ema_update = use_ddp, learnable_codebook, ema_update
        self.path = path
        # model/vq.py:287-288: with use_ddp and sync_kmeans the initial means are drawn from every rank's rows and
        # the k-means statistics are all-reduced; with sync_kmeans off the init is local (no collective at all)
        self.sync_kmeans = bool(use_ddp and sync_kmeans)
        self.sample_fn = _sample_vectors_distributed if self.sync_kmeans else _batched_sample_vectors
        if kmeans_init:
            embed = torch.zeros(num_codebooks, codebook_size, dim)
        else:
            embed = _kaiming_uniform(num_codebooks, codebook_size, dim)
            if self.cosine:
                embed = _l2norm(embed)
        self.register_buffer("initted", torch.Tensor([not kmeans_init]))
        self._initted_host = None
        self.register_buffer("cluster_size", torch.zeros(num_codebooks, codebook_size))
        self.register_buffer("embed_avg", embed.clone())
        if learnable_codebook:
            self.embed = nn.Parameter(embed)
        else:
            self.register_buffer("embed", embed)

    @property
    def metric(self):
        return ops.METRIC_DOT if self.cosine else ops.METRIC_EUCLID

    def transform_input(self, x):
        return _l2norm(x) if self.cosine else x

    def _all_reduce(self, t):
        if self.use_ddp:
            dist.all_reduce(t)

    @torch.no_grad()
    def init_embed_(self, x, heads):
        """k-means initialisation (model/vq.py:182-222,320-342) on the first batch, with the
        assignment and the per-cluster sums done by K3 / K4."""
        # `initted` is a device buffer (it travels in the state_dict like the reference's); reading it
        # every forward would stall the launch queue, so its value is cached on the host and re-read
        # only after a state_dict load
        if self._initted_host is None:
            self._initted_host = bool(self.initted.item())
        if self._initted_host:
            return
        Hc, K, D = self.embed.shape
        rows = x.view(x.size(0), heads, D).permute(1, 0, 2) if Hc == heads else x.reshape(1, -1, D)
        means = self.sample_fn(rows, K).contiguous()              # vq.py:192
        counts = None
        for _ in range(self.kmeans_iters):
            ind, _, _, _ = ops.vq_assign_raw(x, means, heads, self.metric, ops.PATH_EXACT,
                                             want_sq_err=False)
            counts, sums, flat = ops.vq_code_stats(x, ind, (Hc, K, D), heads)
            if self.sync_kmeans:                                   # kmeans_all_reduce_fn (vq.py:288)
                dist.all_reduce(flat)
            empty = counts == 0
            new = sums / counts.masked_fill(empty, 1).unsqueeze(-1)
            if self.cosine:
                new = _l2norm(new)
            means = torch.where(empty.unsqueeze(-1), means, new).contiguous()
        self.embed.data.copy_(means)
        self.embed_avg.data.copy_(means * counts.unsqueeze(-1))
        self.cluster_size.data.copy_(counts)
        self.initted.data.fill_(1.0)
        self._initted_host = True

    def _load_from_state_dict(self, *args, **kwargs):
        self._initted_host = None
        return super()._load_from_state_dict(*args, **kwargs)

    def forward(self, x, heads, training, freeze_codebook, commit_to_embed):
        """x [M, heads*D] fp32 (already l2-normalised for cosine).  Returns (quantize with the
        straight-through gradient when training, ind [M, heads] int64, mse scalar)."""
        x = x.float()
        self.init_embed_(x, heads)
        embed = self.embed if self.learnable_codebook else self.embed.detach()
        q, ind, mse = ops.vq_quantize(x, embed, heads, self.metric, self.path,
                                      straight_through=training,
                                      embed_grad=commit_to_embed and self.learnable_codebook)
        if training and self.ema_update and not freeze_codebook:
            with torch.no_grad():
                counts, sums, flat = ops.vq_code_stats(x.detach(), ind, tuple(self.embed.shape), heads)
                self._all_reduce(flat)                            # model/vq.py:494,499 / 666,672
                ops.vq_ema_apply(counts, sums, self.cluster_size, self.embed_avg, self.embed.data,
                                 self.decay, self.eps, self.cosine)
        return q, ind, mse


class EuclideanCodebook(_Codebook):
    cosine = False


class CosineSimCodebook(_Codebook):
    cosine = True


def _orthogonal_loss(t):
    """model/vq.py:232-237 -- tiny (<= max_codes x D) and stays in torch."""
    h, n = t.shape[:2]
    nc = _l2norm(t)
    cs = torch.einsum("hid,hjd->hij", nc, nc)
    return (cs ** 2).sum() / (h * n ** 2) - (1 / n)


class VectorQuantize(nn.Module):
    def __init__(self, dim, codebook_size, codebook_dim=None, heads=1,
                 separate_codebook_per_head=False, decay=0.8, eps=1e-5, freeze_codebook=False,
                 kmeans_init=False, kmeans_iters=10, sync_kmeans=True, use_cosine_sim=False,
                 threshold_ema_dead_code=0, channel_last=True, accept_image_fmap=False,
                 commitment_weight=1., commitment_use_cross_entropy_loss=False,
                 orthogonal_reg_weight=0., orthogonal_reg_active_codes_only=False,
                 orthogonal_reg_max_codes=None, stochastic_sample_codes=False,
                 sample_codebook_temp=1., straight_through=False, reinmax=False,
                 sync_codebook=None, sync_affine_param=False, ema_update=True,
                 learnable_codebook=False,
                 in_place_codebook_optimizer: Optional[Callable] = None, affine_param=False,
                 affine_param_batch_decay=0.99, affine_param_codebook_decay=0.9,
                 sync_update_v=0., assign_path=ops.PATH_AUTO):
        super().__init__()
        unsupported = dict(accept_image_fmap=accept_image_fmap,
                           commitment_use_cross_entropy_loss=commitment_use_cross_entropy_loss,
                           stochastic_sample_codes=stochastic_sample_codes,
                           straight_through=straight_through, reinmax=reinmax,
                           affine_param=affine_param, sync_affine_param=sync_affine_param,
                           in_place_codebook_optimizer=in_place_codebook_optimizer is not None,
                           threshold_ema_dead_code=threshold_ema_dead_code > 0,
                           sync_update_v=sync_update_v > 0, channel_first=not channel_last)
        bad = [k for k, v in unsupported.items() if v]
        if bad:
            raise NotImplementedError(f"VectorQuantize options outside the B200 hot path: {bad}")
        assert not (ema_update and learnable_codebook), "learnable codebook not compatible with EMA update"
        self.dim, self.heads = dim, heads
        self.separate_codebook_per_head = separate_codebook_per_head
        codebook_dim = dim if codebook_dim is None else codebook_dim
        inner = codebook_dim * heads
        self.has_projections = inner != dim
        self.project_in = nn.Linear(dim, inner) if self.has_projections else nn.Identity()
        self.project_out = nn.Linear(inner, dim) if self.has_projections else nn.Identity()
        self.eps, self.commitment_weight = eps, commitment_weight
        self.commitment_use_cross_entropy_loss = False
        self.learnable_codebook = learnable_codebook
        self.has_codebook_orthogonal_loss = orthogonal_reg_weight > 0
        self.orthogonal_reg_weight = orthogonal_reg_weight
        self.orthogonal_reg_active_codes_only = orthogonal_reg_active_codes_only
        self.orthogonal_reg_max_codes = orthogonal_reg_max_codes
        self.sync_update_v = sync_update_v
        if sync_codebook is None:
            sync_codebook = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
        if kmeans_init and sync_codebook and heads > 1 and separate_codebook_per_head:
            raise NotImplementedError("distributed kmeans init with several codebooks (model/vq.py:281)")
        cls = CosineSimCodebook if use_cosine_sim else EuclideanCodebook
        self._codebook = cls(dim=codebook_dim, codebook_size=codebook_size,
                             num_codebooks=heads if separate_codebook_per_head else 1,
                             kmeans_init=kmeans_init, kmeans_iters=kmeans_iters, decay=decay, eps=eps,
                             use_ddp=sync_codebook, sync_kmeans=sync_kmeans,
                             learnable_codebook=self.has_codebook_orthogonal_loss or learnable_codebook,
                             ema_update=ema_update, path=assign_path)
        self.in_place_codebook_optimizer = None
        self.codebook_size = codebook_size
        self.accept_image_fmap, self.channel_last = False, True

    @property
    def codebook(self):
        cb = self._codebook.embed
        return cb if self.separate_codebook_per_head else cb[0]

    @codebook.setter
    def codebook(self, codes):
        if not self.separate_codebook_per_head:
            codes = codes.unsqueeze(0)
        self._codebook.embed.copy_(codes)

    def get_codes_from_indices(self, indices):
        cb = self._codebook.embed                                  # (Hc, K, D)
        if not self.separate_codebook_per_head:
            return cb[0][indices].flatten(-2) if self.heads > 1 else cb[0][indices]
        lead = indices.shape[:-1]
        flat = indices.reshape(-1, self.heads)
        codes = torch.stack([cb[h][flat[:, h]] for h in range(self.heads)], dim=1)
        return codes.reshape(*lead, -1)

    def get_output_from_indices(self, indices):
        return self.project_out(self.get_codes_from_indices(indices))

    def forward(self, x, indices=None, mask=None, sample_codebook_temp=None, freeze_codebook=False):
        if indices is not None or mask is not None:
            raise NotImplementedError("`indices` (CE loss) and `mask` are outside the B200 hot path")
        if x.dim() not in (2, 3):
            raise ValueError("x must be (N, dim) or (B, n, dim)")
        lead = x.shape[:-1]
        heads, cb = self.heads, self._codebook
        h = self._project(self.project_in, x).reshape(-1, heads * cb.dim)   # 'b n (h d)' rows, heads in columns
        if cb.cosine:                                              # l2norm per head (vq.py:891)
            hv = ops.l2norm_rows(dense.pad4(h.float().contiguous().view(-1, cb.dim)))             # K11 (zeros keep the norm)
            h = (hv[:, :cb.dim] if hv.size(1) != cb.dim else hv).reshape(-1, heads * cb.dim)
        commit_to_embed = self.learnable_codebook and not freeze_codebook
        q, ind, mse = cb(h, heads, self.training, freeze_codebook, commit_to_embed)
        loss = torch.zeros(1, device=x.device, requires_grad=self.training)    # vq.py:983
        if self.training:
            if self.commitment_weight > 0:
                loss = loss + mse * self.commitment_weight         # vq.py:1007-1009
            if self.has_codebook_orthogonal_loss:
                codes = cb.embed
                if self.orthogonal_reg_active_codes_only:
                    assert not (heads > 1 and self.separate_codebook_per_head)
                    codes = codes[:, torch.unique(ind)]
                n_codes = codes.shape[-2]
                if self.orthogonal_reg_max_codes is not None and n_codes > self.orthogonal_reg_max_codes:
                    pick = torch.randperm(n_codes, device=x.device)[:self.orthogonal_reg_max_codes]
                    codes = codes[:, pick]
                loss = loss + _orthogonal_loss(codes) * self.orthogonal_reg_weight
        embed_ind = ind.view(*lead, heads) if heads > 1 else ind.view(*lead)
        orig_quantize = q.view(*lead, heads * cb.dim)
        quantize = self._project(self.project_out, orig_quantize)
        return quantize, embed_ind, loss, orig_quantize

    @staticmethod
    def _project(lin, t):
        """project_in / project_out (model/vq.py:738-739,881,1041) on the tensor cores at fp32 accuracy
        (K5/K6): with the shipped 768 <-> 4 x 768 widths the two fp32 SGEMMs and their
        gradients were 2/3 of a pretrain mini-batch.  Odd widths are zero-padded, weight matrices of more than 64
        tiles of 256 x 256 are cut into output blocks for the gradient product (dense.py)."""
        if not isinstance(lin, nn.Linear):
            return lin(t)
        n, k = lin.weight.shape
        out = dense.linear_sum([t.float().reshape(-1, k)], [lin.weight], lin.bias)
        return out.view(*t.shape[:-1], n)
