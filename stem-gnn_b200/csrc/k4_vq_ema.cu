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
