// k4_vq_ema.cu -- K4: per-code statistics and the EMA codebook update.
//
// Replaces onehot.sum(1) and einsum('h n d, h n c -> h c d') (reference STEM-GNN/model/vq.py:
// 492,497 Euclidean / 665,670 cosine) -- an N x K one-hot and an N x D x K GEMM in the
// reference -- by a segmented reduction: items (head,row) are STABLE-sorted by code (the K0
// radix sort), then every fixed chunk of the sorted list is walked by one thread group that
// adds the rows one by one; a code that spans chunks gets its later pieces added in chunk order
// by a second kernel.  Deterministic, no atomics.  HBM traffic: N*D*4 B (x, read once) +
// N*(8+4*4) B index work + 2*(K*D+K)*4 B.
// stemgnn_vq_ema_apply is lerp_ + laplace smoothing + divide (+ double l2norm), vq.py:495-506 /
// 668-681, 40-46, 102-104.
#include "common.cuh"

namespace stemgnn {

struct StatsView {
    const float* x; int64_t x_row_stride, x_head_stride;
    const int64_t* ind; int64_t ind_row_stride, ind_head_stride;
    int64_t heads, embed_heads, rows;
    int K, D;
};

// item t -> (h, m): separate codebooks: t = h*rows + m; shared codebook: t = m*heads + h
__device__ __forceinline__ void item_hm(const StatsView& v, int64_t t, int64_t& h, int64_t& m) {
    if (v.embed_heads == v.heads) { h = t / v.rows; m = t - h * v.rows; }
    else { m = t / v.heads; h = t - m * v.heads; }
}

__global__ void stats_keys_kernel(StatsView v, int64_t T, int32_t* __restrict__ keys) {
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T) return;
    int64_t h, m;
    item_hm(v, t, h, m);
    int64_t code = v.ind[m * v.ind_row_stride + h * v.ind_head_stride];
    keys[t] = (int32_t)((v.embed_heads == v.heads ? h * v.K : 0) + code);
}

__global__ void stats_counts_kernel(const int32_t* __restrict__ off, int64_t KK,
                                    float* __restrict__ counts) {
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k < KK) counts[k] = (float)(off[k + 1] - off[k]);
}

// One group of G threads walks CHUNK consecutive sorted items; thread j owns VEC-wide feature
// columns j, j+G, ...  (NCOL of them).  A piece that continues a code begun in an earlier chunk
// goes to head_partial[chunk]; pieces of codes that begin in this chunk go straight to sums.
template <int VEC, int NCOL>
__global__ void __launch_bounds__(256)
stats_chunk_kernel(StatsView v, int64_t T, int chunk, int G, const int32_t* __restrict__ skeys,
                   const int32_t* __restrict__ sitems, const int32_t* __restrict__ off,
                   float* __restrict__ sums, float* __restrict__ head_partial) {
    const int groups = blockDim.x / G;
    const int g = threadIdx.x / G, j = threadIdx.x - g * G;
    const int64_t c = (int64_t)blockIdx.x * groups + g;
    const int64_t p0 = c * chunk;
    if (g >= groups || p0 >= T) return;
    const int64_t p1 = p0 + chunk < T ? p0 + chunk : T;
    const int D = v.D;
    int colo[NCOL];
    bool ok[NCOL];
#pragma unroll
    for (int q = 0; q < NCOL; ++q) { colo[q] = (q * G + j) * VEC; ok[q] = colo[q] < D; }
    float acc[NCOL][VEC];
#pragma unroll
    for (int q = 0; q < NCOL; ++q)
#pragma unroll
        for (int e = 0; e < VEC; ++e) acc[q][e] = 0.f;

    int cur = skeys[p0];
    bool is_head = off[cur] < p0;
    constexpr int U = 4;
    for (int64_t p = p0; p < p1; p += U) {
        int key[U];
        const float* rp[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (p + u < p1) {
                key[u] = __ldg(skeys + p + u);
                int64_t h, m;
                item_hm(v, (int64_t)__ldg(sitems + p + u), h, m);
                rp[u] = v.x + m * v.x_row_stride + h * v.x_head_stride;
            } else { key[u] = -1; rp[u] = nullptr; }
        }
        float val[U][NCOL][VEC];
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int q = 0; q < NCOL; ++q) {
                if (key[u] >= 0 && ok[q]) {
                    if (VEC == 4) {
                        float4 t4 = __ldg(reinterpret_cast<const float4*>(rp[u] + colo[q]));
                        val[u][q][0] = t4.x; val[u][q][1 % VEC] = t4.y;
                        val[u][q][2 % VEC] = t4.z; val[u][q][3 % VEC] = t4.w;
                    } else val[u][q][0] = __ldg(rp[u] + colo[q]);
                }
            }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (key[u] < 0) continue;
            if (key[u] != cur) {
                float* dst = is_head ? head_partial + c * D : sums + (int64_t)cur * D;
#pragma unroll
                for (int q = 0; q < NCOL; ++q)
                    if (ok[q])
#pragma unroll
                        for (int e = 0; e < VEC; ++e) { dst[colo[q] + e] = acc[q][e]; acc[q][e] = 0.f; }
                cur = key[u];
                is_head = false;
            }
#pragma unroll
            for (int q = 0; q < NCOL; ++q)
                if (ok[q])
#pragma unroll
                    for (int e = 0; e < VEC; ++e) acc[q][e] = __fadd_rn(acc[q][e], val[u][q][e]);
        }
    }
    float* dst = is_head ? head_partial + c * D : sums + (int64_t)cur * D;
#pragma unroll
    for (int q = 0; q < NCOL; ++q)
        if (ok[q])
#pragma unroll
            for (int e = 0; e < VEC; ++e) dst[colo[q] + e] = acc[q][e];
}

// sums[k] += head_partial[c] for the chunks c after the one in which code k starts, in order.
__global__ void stats_combine_kernel(const int32_t* __restrict__ off, int64_t KK, int D, int chunk,
                                     const float* __restrict__ head_partial, float* __restrict__ sums) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= KK * D) return;
    int64_t k = i / D;
    int d = (int)(i - k * D);
    int b = off[k], e = off[k + 1];
    if (e <= b) return;
    int64_t c0 = b / chunk, c1 = (e - 1) / chunk;
    if (c1 == c0) return;
    float s = sums[i];
    for (int64_t c = c0 + 1; c <= c1; ++c) s = __fadd_rn(s, head_partial[c * D + d]);
    sums[i] = s;
}

// torch lerp_(end, w): w < 0.5 ? start + w*(end-start) : end - (end-start)*(1-w)
__device__ __forceinline__ float lerp_ref(float a, float b, float w) {
    float d = __fsub_rn(b, a);
    return w < 0.5f ? __fadd_rn(a, __fmul_rn(w, d)) : __fsub_rn(b, __fmul_rn(d, __fsub_rn(1.f, w)));
}

__device__ __forceinline__ float block_sum_256(float v, float* sh) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += sh[i];
    return t;
}

// pass A, one CTA per codebook head: cluster_size.lerp_(counts), total = sum(cluster_size)
__global__ void __launch_bounds__(256)
ema_cluster_kernel(const float* __restrict__ counts, int K, float w, float* __restrict__ cluster_size,
                   float* __restrict__ totals) {
    __shared__ float sh[8];
    float* cs = cluster_size + (int64_t)blockIdx.x * K;
    const float* cn = counts + (int64_t)blockIdx.x * K;
    float part = 0.f;
    for (int k = threadIdx.x; k < K; k += 256) {
        float v = lerp_ref(cs[k], cn[k], w);
        cs[k] = v;
        part += v;
    }
    float tot = block_sum_256(part, sh);
    if (threadIdx.x == 0) totals[blockIdx.x] = tot;
}

// pass B, one warp per code row: embed_avg.lerp_(sums); embed = embed_avg / smoothed (+2x l2norm)
__global__ void __launch_bounds__(256)
ema_embed_kernel(const float* __restrict__ sums, const float* __restrict__ cluster_size,
                 const float* __restrict__ totals, int64_t KK, int K, int D, float w, float eps,
                 float keps, int cosine, float* __restrict__ embed_avg, float* __restrict__ embed) {
    int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (row >= KK) return;
    const int lane = lane_id();
    const float tot = totals[row / K];
    // laplace_smoothing(cs, K, eps) * cs.sum(): ((cs + eps) / (tot + K*eps)) * tot
    const float sm = __fmul_rn(__fdiv_rn(__fadd_rn(cluster_size[row], eps),
                                         __fadd_rn(tot, keps)), tot);
    float* ea = embed_avg + row * D;
    float* em = embed + row * D;
    const float* su = sums + row * D;
    float ss = 0.f;
    for (int d = lane; d < D; d += 32) {
        float a = lerp_ref(ea[d], su[d], w);
        ea[d] = a;
        float e = __fdiv_rn(a, sm);
        em[d] = e;
        ss = fmaf(e, e, ss);
    }
    if (!cosine) return;
    for (int rep = 0; rep < 2; ++rep) {   // F.normalize twice (vq.py:679-681): v / max(||v||, 1e-12)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
        float nrm = fmaxf(__fsqrt_rn(ss), 1e-12f);
        ss = 0.f;
        __syncwarp();
        for (int d = lane; d < D; d += 32) {
            float e = __fdiv_rn(em[d], nrm);
            em[d] = e;
            ss = fmaf(e, e, ss);
        }
    }
}

}  // namespace stemgnn

using namespace stemgnn;

namespace {
struct StatsPlan {
    int64_t T, KK, nchunks;
    int chunk, G, vec, ncol;
};

bool make_plan(const stemgnn_vq_code_stats_args* a, StatsPlan& p) {
    p.T = a->heads * a->rows;
    p.KK = a->embed_heads * a->codes;
    int64_t c = p.T / ((int64_t)148 * 64);
    p.chunk = (int)(c < 32 ? 32 : (c > 256 ? 256 : c));
    p.nchunks = (p.T + p.chunk - 1) / p.chunk;
    const bool v4 = a->dim % 4 == 0 && a->x_row_stride % 4 == 0 && a->x_head_stride % 4 == 0 &&
                    (reinterpret_cast<uintptr_t>(a->x) & 15) == 0;
    p.vec = v4 ? 4 : 1;
    int64_t cols = (a->dim + p.vec - 1) / p.vec;
    p.G = 1;
    while (p.G < 32 && p.G < cols) p.G <<= 1;
    int64_t nc = (cols + p.G - 1) / p.G;
    p.ncol = 1;
    while (p.ncol < nc) p.ncol <<= 1;
    return p.ncol <= (v4 ? 8 : 32);
}

template <int VEC, int NCOL>
void launch_chunks(const StatsView& v, const StatsPlan& p, const int32_t* skeys, const int32_t* sitems,
                   const int32_t* off, float* sums, float* hp, cudaStream_t s) {
    const int groups = 256 / p.G;
    unsigned grid = (unsigned)((p.nchunks + groups - 1) / groups);
    SG_LAUNCH(stats_chunk_kernel<VEC, NCOL><<<grid, 256, 0, s>>>(v, p.T, p.chunk, p.G, skeys, sitems, off, sums, hp));
}
}  // namespace

extern "C" int stemgnn_vq_code_stats_workspace(const stemgnn_vq_code_stats_args* a, size_t* bytes) {
    SG_CHECK_ARG(a && bytes, "NULL pointer");
    SG_CHECK_ARG(a->heads >= 1 && a->rows >= 0 && a->codes >= 1 && a->dim >= 1, "bad sizes");
    SG_CHECK_ARG(a->embed_heads == 1 || a->embed_heads == a->heads, "embed_heads must be 1 or heads");
    StatsPlan p;
    SG_CHECK_ARG(make_plan(a, p), "dim too large for the statistics kernel");
    SG_CHECK_ARG(p.T < 2147483647 && p.KK < 2147483647, "int32 index range");
    size_t b = 0;
    b += align_up((size_t)p.T * 4, 256);                    // keys
    b += align_up((size_t)p.T * 4, 256);                    // sorted items
    b += align_up((size_t)(p.KK + 1) * 4, 256);             // offsets
    b += align_up((size_t)p.nchunks * a->dim * 4, 256);     // head partials
    b += radix_sort_workspace(p.T);
    *bytes = b;
    return STEMGNN_OK;
}

extern "C" int stemgnn_vq_code_stats(const stemgnn_vq_code_stats_args* a, void* workspace,
                                     size_t workspace_bytes, stemgnn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    size_t need = 0;
    int rc = stemgnn_vq_code_stats_workspace(a, &need);
    if (rc) return rc;
    SG_CHECK_ARG(a->counts && a->sums, "counts/sums is NULL");
    StatsPlan p;
    make_plan(a, p);
    cudaMemsetAsync(a->sums, 0, (size_t)p.KK * a->dim * 4, stream);
    if (p.T == 0) {
        cudaMemsetAsync(a->counts, 0, (size_t)p.KK * 4, stream);
        SG_CHECK_LAUNCH();
        return STEMGNN_OK;
    }
    SG_CHECK_ARG(a->x && a->ind, "x/ind is NULL");
    if (workspace == nullptr || workspace_bytes < need) {
        set_error("stemgnn_vq_code_stats: workspace too small (%zu < %zu)", workspace_bytes, need);
        return STEMGNN_ENOMEM_WS;
    }
    Workspace ws(workspace, workspace_bytes);
    int32_t* keys = ws.take<int32_t>(p.T);
    int32_t* sitems = ws.take<int32_t>(p.T);
    int32_t* off = ws.take<int32_t>(p.KK + 1);
    float* hp = ws.take<float>((size_t)p.nchunks * a->dim);
    StatsView v;
    v.x = a->x; v.x_row_stride = a->x_row_stride; v.x_head_stride = a->x_head_stride;
    v.ind = a->ind; v.ind_row_stride = a->ind_row_stride; v.ind_head_stride = a->ind_head_stride;
    v.heads = a->heads; v.embed_heads = a->embed_heads; v.rows = a->rows;
    v.K = (int)a->codes; v.D = (int)a->dim;
    SG_LAUNCH(stats_keys_kernel<<<(unsigned)((p.T + 255) / 256), 256, 0, stream>>>(v, p.T, keys));
    const int32_t* skeys = radix_sort_pairs(nullptr, keys, p.T, p.KK, sitems, ws, stream);
    launch_lower_bound(skeys, p.T, p.KK, off, stream);
    SG_LAUNCH(stats_counts_kernel<<<(unsigned)((p.KK + 255) / 256), 256, 0, stream>>>(off, p.KK, a->counts));
#define SG_STATS_CASE(V, N) \
    if (p.vec == V && p.ncol == N) launch_chunks<V, N>(v, p, skeys, sitems, off, a->sums, hp, stream)
    SG_STATS_CASE(4, 1); SG_STATS_CASE(4, 2); SG_STATS_CASE(4, 4); SG_STATS_CASE(4, 8);
    SG_STATS_CASE(1, 1); SG_STATS_CASE(1, 2); SG_STATS_CASE(1, 4); SG_STATS_CASE(1, 8);
    SG_STATS_CASE(1, 16); SG_STATS_CASE(1, 32);
#undef SG_STATS_CASE
    int64_t tot = p.KK * a->dim;
    SG_LAUNCH(stats_combine_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, stream>>>(off, p.KK, (int)a->dim,
                                                                            p.chunk, hp, a->sums));
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}

extern "C" int stemgnn_vq_ema_apply(const float* counts, const float* sums, int64_t embed_heads,
                                    int64_t codes, int64_t dim, double decay, double eps, int cosine,
                                    float* cluster_size, float* embed_avg, float* embed,
                                    float* totals_scratch, stemgnn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SG_CHECK_ARG(embed_heads >= 1 && codes >= 1 && dim >= 1, "bad sizes");
    SG_CHECK_ARG(counts && sums && cluster_size && embed_avg && embed && totals_scratch, "NULL pointer");
    const float w = (float)(1.0 - decay);   // python double 1-decay, rounded to fp32 by lerp_
    SG_LAUNCH(ema_cluster_kernel<<<(unsigned)embed_heads, 256, 0, stream>>>(counts, (int)codes, w, cluster_size,
                                                                  totals_scratch));
    const int64_t KK = embed_heads * codes;
    SG_LAUNCH(ema_embed_kernel<<<(unsigned)((KK * 32 + 255) / 256), 256, 0, stream>>>(
        sums, cluster_size, totals_scratch, KK, (int)codes, (int)dim, w, (float)eps,
        (float)((double)codes * eps), cosine, embed_avg, embed));
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}
