// k3_vq_simt.cu -- exact fp32 CUDA-core pieces of the VQ assignment (K3 path 1) and the
// gather / commitment-partial / backward kernels shared with the tcgen05 path (k3_vq_tc.cu).
//
// Replaces model/vq.py:472-482 (Euclidean) / 650-659 (cosine): dist, argmax (first index on
// ties, vq.py:79), quantize = embed[ind]; vq.py:1007 mse partial; vq.py:937 straight-through.
// The N x K distance matrix and the one-hot are never materialised.
#include "k3_vq_common.cuh"

namespace stemgnn {

// ---- squared row norms: out[h*rows + m] = sum_d v(h,m,d)^2, one warp per row ------------------
__global__ void vq_sqnorm_kernel(const float* __restrict__ v, int64_t rows, int64_t heads, int D,
                                 int64_t row_stride, int64_t head_stride, float* __restrict__ out) {
    int64_t gw = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (gw >= rows * heads) return;
    int64_t h = gw / rows, m = gw - h * rows;
    const float* p = v + m * row_stride + h * head_stride;
    float s = 0.f;
    for (int d = lane_id(); d < D; d += 32) { float t = __ldg(p + d); s = fmaf(t, t, s); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane_id() == 0) out[gw] = s;
}

// ---- exact scan: 64 rows x all codes per CTA, 4x4 register tile per thread --------------------
constexpr int SC_TM = 64, SC_TN = 64, SC_DK = 32;

__global__ void __launch_bounds__(256)
vq_scan_kernel(VqView q, const float* __restrict__ x2, const float* __restrict__ e2,
               const int32_t* __restrict__ row_list, const int32_t* __restrict__ row_count,
               int32_t* __restrict__ best_out) {
    // row_list == NULL: CTA b handles rows [64b, 64b+64) of head blockIdx.y.
    // row_list != NULL: entries are flat ids h*rows+m; *row_count of them are valid.
    __shared__ float Xs[SC_DK][SC_TM + 1];
    __shared__ float Es[SC_DK][SC_TN + 1];
    __shared__ float red_s[SC_TM][17];
    __shared__ int red_i[SC_TM][17];
    __shared__ int64_t row_base[SC_TM];   // element offset of the row in x, or -1
    __shared__ int row_flat[SC_TM];

    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    int64_t total = row_list ? (int64_t)*row_count : q.rows;
    int64_t r0 = (int64_t)blockIdx.x * SC_TM;
    if (r0 >= total) return;
    if (tid < SC_TM) {
        int64_t r = r0 + tid;
        int64_t flat = -1;
        if (r < total) flat = row_list ? (int64_t)row_list[r] : (int64_t)blockIdx.y * q.rows + r;
        row_flat[tid] = (int)flat;
        if (flat >= 0) {
            int64_t h = flat / q.rows, m = flat - h * q.rows;
            row_base[tid] = m * q.x_row_stride + h * q.x_head_stride;
        } else row_base[tid] = -1;
    }
    __syncthreads();
    // all rows of one CTA share a head only when row_list == NULL; with a row list the head
    // can differ per row, so the codebook head is resolved per row when embed_heads > 1.
    const int D = q.D, K = q.K;
    float best_s[4];
    int best_i[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) { best_s[i] = -INFINITY; best_i[i] = 0; }

    // With a row list and per-head codebooks, process one head at a time (rare fallback path).
    const int n_head_pass = (row_list && q.embed_heads > 1) ? (int)q.heads : 1;
    for (int hp = 0; hp < n_head_pass; ++hp) {
        int64_t eh;
        if (q.embed_heads == 1) eh = 0;
        else eh = row_list ? hp : blockIdx.y;
        const float* E = q.embed + eh * (int64_t)K * D;
        const float* e2h = e2 + eh * (int64_t)K;
        for (int c0 = 0; c0 < K; c0 += SC_TN) {
            float acc[4][4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
            for (int d0 = 0; d0 < D; d0 += SC_DK) {
                const int k = tid & 31, rr = tid >> 5;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    int r = rr + 8 * i;
                    int64_t rb = row_base[r];
                    Xs[k][r] = (rb >= 0 && d0 + k < D) ? __ldg(q.x + rb + d0 + k) : 0.f;
                    int c = c0 + r;
                    Es[k][r] = (c < K && d0 + k < D) ? __ldg(E + (int64_t)c * D + d0 + k) : 0.f;
                }
                __syncthreads();
#pragma unroll 8
                for (int kk = 0; kk < SC_DK; ++kk) {
                    float a[4], b[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) { a[i] = Xs[kk][ty * 4 + i]; b[i] = Es[kk][tx * 4 + i]; }
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
                }
                __syncthreads();
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                int flat = row_flat[ty * 4 + i];
                if (flat < 0) continue;
                if (n_head_pass > 1 && flat / q.rows != hp) continue;
                float xx = q.metric == 0 ? x2[flat] : 0.f;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    int c = c0 + tx * 4 + j;
                    if (c >= K) continue;
                    float s = vq_score(q.metric, xx, e2h[c], acc[i][j]);
                    if (s > best_s[i]) { best_s[i] = s; best_i[i] = c; }
                }
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) { red_s[ty * 4 + i][tx] = best_s[i]; red_i[ty * 4 + i][tx] = best_i[i]; }
    __syncthreads();
    if (tid < SC_TM && row_flat[tid] >= 0) {
        float bs = red_s[tid][0];
        int bi = red_i[tid][0];
#pragma unroll
        for (int t = 1; t < 16; ++t) {
            float s = red_s[tid][t];
            int i2 = red_i[tid][t];
            if (s > bs || (s == bs && i2 < bi)) { bs = s; bi = i2; }
        }
        best_out[row_flat[tid]] = bi;
    }
}

// ---- gather + commitment partial: one warp per (head,row) --------------------------------------
// quantize(h,m,:) = embed[eh, best, :]; ind(h,m) = best; partials[h*nblk + blockIdx.x] = sum (q-x)^2
// over this CTA's rows, reduced in a fixed order (shuffle tree, then warps in index order).
// Rows the tcgen05 path left open are settled first, while the warp holds the row anyway:
// best == -1: re-score the row's candidate list (cand[flat][0..CAP), -1 padded) with the
// reference's fp32 formula, first index winning ties.  Rows that need the exact scan over all K
// codes (best == -2) are NOT scanned here -- one such row would hold its CTA for K dot products
// (430 us at K=16384, D=256: the kernel's tail) -- the CTA appends itself to open_list and
// vq_rescan_kernel spreads those rows over the whole GPU.
__device__ __forceinline__ float warp_dot(const float* __restrict__ a, const float* __restrict__ b,
                                          int D, bool vec4, int lane) {
    float s = 0.f;
    if (vec4) {
        for (int d = lane * 4; d < D; d += 128) {
            const float4 u = __ldg(reinterpret_cast<const float4*>(a + d));
            const float4 w = __ldg(reinterpret_cast<const float4*>(b + d));
            s = fmaf(u.x, w.x, s); s = fmaf(u.y, w.y, s); s = fmaf(u.z, w.z, s); s = fmaf(u.w, w.w, s);
        }
    } else {
        for (int d = lane; d < D; d += 32) s = fmaf(__ldg(a + d), __ldg(b + d), s);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    return s;
}

// G dot products of one row against G consecutive codes, each summed in warp_dot's order (so a
// (row, code) pair scores the same wherever it is evaluated); x is loaded once, the G code rows
// are G independent loads in flight.
template <int G>
__device__ __forceinline__ void warp_dot_group(const float* __restrict__ a, const float* __restrict__ b,
                                               int D, int ncodes, bool vec4, int lane, float (&out)[G]) {
    float s[G];
#pragma unroll
    for (int i = 0; i < G; ++i) s[i] = 0.f;
    if (vec4) {
        for (int d = lane * 4; d < D; d += 128) {
            const float4 u = __ldg(reinterpret_cast<const float4*>(a + d));
            float4 w[G];
#pragma unroll
            for (int i = 0; i < G; ++i)
                w[i] = i < ncodes ? __ldg(reinterpret_cast<const float4*>(b + (int64_t)i * D + d))
                                  : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int i = 0; i < G; ++i) {
                s[i] = fmaf(u.x, w[i].x, s[i]); s[i] = fmaf(u.y, w[i].y, s[i]);
                s[i] = fmaf(u.z, w[i].z, s[i]); s[i] = fmaf(u.w, w[i].w, s[i]);
            }
        }
    } else {
        for (int d = lane; d < D; d += 32) {
            const float u = __ldg(a + d);
#pragma unroll
            for (int i = 0; i < G; ++i)
                if (i < ncodes) s[i] = fmaf(u, __ldg(b + (int64_t)i * D + d), s[i]);
        }
    }
#pragma unroll
    for (int i = 0; i < G; ++i) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s[i] += __shfl_xor_sync(0xffffffffu, s[i], o);
        out[i] = s[i];
    }
}

// quantize row <- code row, returns this lane's share of |code - x|^2 (the warp sums it)
__device__ __forceinline__ float gather_row(const float* __restrict__ xp, const float* __restrict__ e,
                                            float* __restrict__ qp, int D, int lane) {
    float s = 0.f;
    const bool vec4 = (D & 3) == 0 &&
                      ((reinterpret_cast<uintptr_t>(xp) | reinterpret_cast<uintptr_t>(e) |
                        reinterpret_cast<uintptr_t>(qp)) & 15) == 0;
    if (vec4) {
        // two 16-byte loads of each operand in flight per lane (D = 256: the whole row)
        for (int d = lane * 4; d < D; d += 256) {
            const bool two = d + 128 < D;
            const float4 e0 = __ldg(reinterpret_cast<const float4*>(e + d));
            const float4 x0 = __ldg(reinterpret_cast<const float4*>(xp + d));
            float4 e1 = make_float4(0.f, 0.f, 0.f, 0.f), x1 = e1;
            if (two) {
                e1 = __ldg(reinterpret_cast<const float4*>(e + d + 128));
                x1 = __ldg(reinterpret_cast<const float4*>(xp + d + 128));
            }
            float df;
            df = e0.x - x0.x; s = fmaf(df, df, s); df = e0.y - x0.y; s = fmaf(df, df, s);
            df = e0.z - x0.z; s = fmaf(df, df, s); df = e0.w - x0.w; s = fmaf(df, df, s);
            df = e1.x - x1.x; s = fmaf(df, df, s); df = e1.y - x1.y; s = fmaf(df, df, s);
            df = e1.z - x1.z; s = fmaf(df, df, s); df = e1.w - x1.w; s = fmaf(df, df, s);
            if (qp) {
                *reinterpret_cast<float4*>(qp + d) = e0;
                if (two) *reinterpret_cast<float4*>(qp + d + 128) = e1;
            }
        }
    } else {
        for (int d = lane; d < D; d += 32) {
            const float ev = __ldg(e + d);
            const float df = ev - __ldg(xp + d);
            s = fmaf(df, df, s);
            if (qp) qp[d] = ev;
        }
    }
    return s;
}

// best == -1: settle the row from its candidate list (about 1 row in 20) with the reference's fp32
// formula, first index winning ties.  Returns the code, or -2 when only the exact scan can decide.
// The candidate ids arrive in one 32-byte load and all candidate code rows are in flight together
// (walked one by one -- id, then row, then the next id -- a settled row took three times as long as a
// plain one and held its warp's other rows back).  Every (row, code) dot product is summed in
// warp_dot's order, so a pair scores the same wherever it is evaluated.  Out of line: inlined four
// times it pushed the gather kernel into spills.
__device__ __noinline__ int settle_row(const VqView& q, const float* __restrict__ xp, const float* __restrict__ E,
                                       const float* __restrict__ e2h, const int32_t* __restrict__ cand_row, int lane) {
    const int mine = lane < VQ_CAND_CAP ? cand_row[lane] : -1;
    const unsigned invalid = __ballot_sync(0xffffffffu, mine < 0);
    const int n = __ffs(invalid) - 1;                      // ids are -1 padded behind the last one; lanes >= CAP are -1
    int c[VQ_CAND_CAP];
#pragma unroll
    for (int j = 0; j < VQ_CAND_CAP; ++j) c[j] = __shfl_sync(0xffffffffu, mine, j);
    float dot[VQ_CAND_CAP], x2 = 0.f;
#pragma unroll
    for (int j = 0; j < VQ_CAND_CAP; ++j) dot[j] = 0.f;
    const bool v4 = (q.D & 3) == 0 && ((reinterpret_cast<uintptr_t>(xp) | reinterpret_cast<uintptr_t>(E)) & 15) == 0;
    if (v4) {
        for (int d = lane * 4; d < q.D; d += 128) {
            const float4 u = __ldg(reinterpret_cast<const float4*>(xp + d));
            float4 w[VQ_CAND_CAP];
#pragma unroll
            for (int j = 0; j < VQ_CAND_CAP; ++j)
                w[j] = j < n ? __ldg(reinterpret_cast<const float4*>(E + (int64_t)c[j] * q.D + d))
                             : make_float4(0.f, 0.f, 0.f, 0.f);
            x2 = fmaf(u.x, u.x, x2); x2 = fmaf(u.y, u.y, x2); x2 = fmaf(u.z, u.z, x2); x2 = fmaf(u.w, u.w, x2);
#pragma unroll
            for (int j = 0; j < VQ_CAND_CAP; ++j) {
                dot[j] = fmaf(u.x, w[j].x, dot[j]); dot[j] = fmaf(u.y, w[j].y, dot[j]);
                dot[j] = fmaf(u.z, w[j].z, dot[j]); dot[j] = fmaf(u.w, w[j].w, dot[j]);
            }
        }
    } else {
        for (int d = lane; d < q.D; d += 32) {
            const float u = __ldg(xp + d);
            x2 = fmaf(u, u, x2);
#pragma unroll
            for (int j = 0; j < VQ_CAND_CAP; ++j)
                if (j < n) dot[j] = fmaf(u, __ldg(E + (int64_t)c[j] * q.D + d), dot[j]);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x2 += __shfl_xor_sync(0xffffffffu, x2, o);
    if (q.metric != 0) x2 = 0.f;
    float bs = -INFINITY;
    int bi = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < VQ_CAND_CAP; ++j) {
        float t = dot[j];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if (j < n) {
            const float sc = vq_score(q.metric, x2, e2h[c[j]], t);
            if (sc > bs || (sc == bs && c[j] < bi)) { bs = sc; bi = c[j]; }
        }
    }
    return bi < q.K ? bi : -2;
}

// GR rows per warp: the chain best[row] -> code row address -> loads -> stores is two dependent memory
// latencies long; all GR index loads go first, then the x and code segments of every row are in
// flight together.  A warp is its own unit of bookkeeping (one commitment partial, one open-list
// entry per warp) and the kernel has NO CTA barrier: with per-CTA partials the top stall was
// `barrier` (7.6 warps per issue, ncu) -- every warp waited for the slowest of its CTA, the one that
// had a row to settle, and the idle warps held the SM's occupancy.
__global__ void __launch_bounds__(256, 3)
vq_gather_kernel(VqView q, int32_t* __restrict__ best, const int32_t* __restrict__ cand,
                 const float* __restrict__ e2, double* __restrict__ partials,
                 int32_t* __restrict__ open_list, int32_t* __restrict__ open_count,
                 unsigned long long* __restrict__ stats, int rev) {
    constexpr int GR = VQ_GATHER_ROWS;
    const int w = threadIdx.x >> 5, lane = lane_id();
    const int64_t h = blockIdx.y;
    const int64_t nent = (q.rows + GR - 1) / GR;
    // rev: the first CTAs take the last rows -- the search kernel before this one swept x front to back, so the tail
    // of x is what the L2 still holds (l2_pingpong); entries keep their index, the result does not depend on it
    const int64_t ent = (int64_t)(rev ? gridDim.x - 1 - blockIdx.x : blockIdx.x) * 8 + w;
    if (ent >= nent) return;
    const int64_t m0 = ent * GR;
    const int64_t eh = q.embed_heads == 1 ? 0 : h;
    const float* E = q.embed + eh * (int64_t)q.K * q.D;
    const float* e2h = e2 + eh * (int64_t)q.K;
    const float* x0 = q.x + m0 * q.x_row_stride + h * q.x_head_stride;
    float* q0 = q.quantize ? q.quantize + m0 * q.q_row_stride + h * q.q_head_stride : nullptr;
    int b[GR];
#pragma unroll
    for (int r = 0; r < GR; ++r) b[r] = m0 + r < q.rows ? best[h * q.rows + m0 + r] : -3;
    bool open = false;
#pragma unroll
    for (int r = 0; r < GR; ++r) {
        if (b[r] == -1) {
            const int64_t flat = h * q.rows + m0 + r;
            b[r] = settle_row(q, x0 + r * q.x_row_stride, E, e2h, cand + flat * VQ_CAND_CAP, lane);
            if (lane == 0 && stats) atomicAdd(stats + 0, 1ull);
            if (b[r] == -2 && lane == 0) best[flat] = -2;      // NaN scores: the exact scan decides
        }
        open = open || b[r] == -2;
    }
    float s[GR];
#pragma unroll
    for (int r = 0; r < GR; ++r) s[r] = 0.f;
    const bool vec4 = (q.D & 3) == 0 && ((q.x_row_stride | q.q_row_stride) & 3) == 0 &&
                      ((reinterpret_cast<uintptr_t>(E) | reinterpret_cast<uintptr_t>(x0) | reinterpret_cast<uintptr_t>(q0)) & 15) == 0;
    if (vec4) {
        for (int d = lane * 4; d < q.D; d += 128) {
            float4 ev[GR], xv[GR];
#pragma unroll
            for (int r = 0; r < GR; ++r)
                if (b[r] >= 0) {
                    ev[r] = __ldg(reinterpret_cast<const float4*>(E + (int64_t)b[r] * q.D + d));
                    xv[r] = __ldg(reinterpret_cast<const float4*>(x0 + r * q.x_row_stride + d));
                }
#pragma unroll
            for (int r = 0; r < GR; ++r)
                if (b[r] >= 0) {
                    float df;
                    df = ev[r].x - xv[r].x; s[r] = fmaf(df, df, s[r]); df = ev[r].y - xv[r].y; s[r] = fmaf(df, df, s[r]);
                    df = ev[r].z - xv[r].z; s[r] = fmaf(df, df, s[r]); df = ev[r].w - xv[r].w; s[r] = fmaf(df, df, s[r]);
                    if (q0) *reinterpret_cast<float4*>(q0 + r * q.q_row_stride + d) = ev[r];
                }
        }
    } else {
#pragma unroll
        for (int r = 0; r < GR; ++r)
            if (b[r] >= 0)
                s[r] = gather_row(x0 + r * q.x_row_stride, E + (int64_t)b[r] * q.D,
                                  q0 ? q0 + r * q.q_row_stride : nullptr, q.D, lane);
    }
    double part = 0.0;
#pragma unroll
    for (int r = 0; r < GR; ++r) {          // rows in index order: the commitment sum stays deterministic
        if (b[r] < 0) continue;
        double pr = (double)s[r];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) pr += __shfl_xor_sync(0xffffffffu, pr, o);
        part += pr;
        if (lane == 0 && q.ind) q.ind[(m0 + r) * q.ind_row_stride + h * q.ind_head_stride] = (int64_t)b[r];
    }
    if (lane == 0) {
        if (open) open_list[atomicAdd(open_count, 1)] = (int32_t)(h * nent + ent);
        if (partials) partials[h * nent + ent] = part;
    }
}

// ---- exact scan of the rows left open: one CTA per row at a time, over the whole GPU ----------
// An entry of open_list is a gather warp (VQ_GATHER_ROWS consecutive rows of one head) with at least one row at
// best == -2.  A CTA of 32 warps scans all K codes for each such row (warp w takes the groups of
// eight codes w, w+32, ...: the scan is bound by L2 latency, so 256 code rows are in flight per
// SM; first index wins ties; an all-NaN row takes index 0 like argmax does), writes the row's
// index / quantize row and adds its squared error to the gather CTA's partial -- rows in index
// order, no atomics, so the commitment sum stays deterministic.
constexpr int RS_WARPS = 32, RS_GROUP = 8;

__global__ void __launch_bounds__(RS_WARPS * 32)
vq_rescan_kernel(VqView q, int32_t* __restrict__ best, const float* __restrict__ e2,
                 double* __restrict__ partials, int nblk, const int32_t* __restrict__ open_list,
                 const int32_t* __restrict__ open_count, unsigned long long* __restrict__ stats) {
    __shared__ float s_bs[RS_WARPS];
    __shared__ int s_bi[RS_WARPS];
    __shared__ int s_pick;
    const int w = threadIdx.x >> 5, lane = lane_id();
    const int n_open = *open_count;
    for (int entry = blockIdx.x; entry < n_open; entry += gridDim.x) {
        const int blk = open_list[entry];
        const int64_t h = blk / nblk, bx = blk - h * nblk;
        const int64_t eh = q.embed_heads == 1 ? 0 : h;
        const float* E = q.embed + eh * (int64_t)q.K * q.D;
        const float* e2h = e2 + eh * (int64_t)q.K;
        double extra = 0.0;
        for (int r = 0; r < VQ_GATHER_ROWS; ++r) {
            const int64_t m = bx * VQ_GATHER_ROWS + r;
            if (m >= q.rows) break;
            const int64_t flat = h * q.rows + m;
            if (best[flat] != -2) continue;                   // CTA-uniform
            const float* xp = q.x + m * q.x_row_stride + h * q.x_head_stride;
            const bool vec4 = (q.D & 3) == 0 && ((reinterpret_cast<uintptr_t>(xp) | reinterpret_cast<uintptr_t>(E)) & 15) == 0;
            const float x2 = q.metric == 0 ? warp_dot(xp, xp, q.D, vec4, lane) : 0.f;
            float bs = -INFINITY;
            int bi = 0x7fffffff;
            for (int c0 = w * RS_GROUP; c0 < q.K; c0 += RS_WARPS * RS_GROUP) {
                const int nc = q.K - c0 < RS_GROUP ? q.K - c0 : RS_GROUP;
                float dot[RS_GROUP];
                warp_dot_group<RS_GROUP>(xp, E + (int64_t)c0 * q.D, q.D, nc, vec4, lane, dot);
#pragma unroll
                for (int i = 0; i < RS_GROUP; ++i) {
                    if (i >= nc) break;
                    const float s = vq_score(q.metric, x2, e2h[c0 + i], dot[i]);
                    if (s > bs || (s == bs && c0 + i < bi)) { bs = s; bi = c0 + i; }
                }
            }
            if (lane == 0) { s_bs[w] = bs; s_bi[w] = bi; }
            __syncthreads();
            if (threadIdx.x == 0) {
                float fs = s_bs[0];
                int fi = s_bi[0];
                for (int k = 1; k < RS_WARPS; ++k)
                    if (s_bs[k] > fs || (s_bs[k] == fs && s_bi[k] < fi)) { fs = s_bs[k]; fi = s_bi[k]; }
                s_pick = fi < q.K ? fi : 0;
                best[flat] = s_pick;
                if (stats) atomicAdd(stats + 1, 1ull);
            }
            __syncthreads();
            if (w == 0) {
                const int pick = s_pick;
                float* qp = q.quantize ? q.quantize + m * q.q_row_stride + h * q.q_head_stride : nullptr;
                double part = (double)gather_row(xp, E + (int64_t)pick * q.D, qp, q.D, lane);
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
                extra += part;
                if (lane == 0 && q.ind) q.ind[m * q.ind_row_stride + h * q.ind_head_stride] = (int64_t)pick;
            }
            __syncthreads();
        }
        if (threadIdx.x == 0 && partials) partials[blk] += extra;
    }
}

// fixed-order reduction of the per-warp partials of one head, in two stages: VQ_SQ_CHUNKS CTAs sum
// contiguous chunks (thread-strided, then a tree), one CTA sums the chunk sums.  (One CTA walking all
// partials of a 2.5M-row call took 116 us.)

__device__ __forceinline__ double block_sum_256(double t) {
    __shared__ double sm[256];
    sm[threadIdx.x] = t;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) sm[threadIdx.x] += sm[threadIdx.x + s];
        __syncthreads();
    }
    return sm[0];
}

__global__ void __launch_bounds__(256)
vq_sqerr_chunk_kernel(const double* __restrict__ partials, int nblk, double* __restrict__ chunk_sums) {
    const double* p = partials + (int64_t)blockIdx.y * nblk;
    const int per = (nblk + VQ_SQ_CHUNKS - 1) / VQ_SQ_CHUNKS;
    const int lo = blockIdx.x * per, hi = lo + per < nblk ? lo + per : nblk;
    double t = 0.0;
    for (int i = lo + threadIdx.x; i < hi; i += 256) t += p[i];
    t = block_sum_256(t);
    if (threadIdx.x == 0) chunk_sums[blockIdx.y * VQ_SQ_CHUNKS + blockIdx.x] = t;
}

__global__ void __launch_bounds__(256)
vq_sqerr_reduce_kernel(const double* __restrict__ chunk_sums, int nchunks, double* __restrict__ sq_err) {
    const double* p = chunk_sums + (int64_t)blockIdx.x * nchunks;
    double t = 0.0;
    for (int i = threadIdx.x; i < nchunks; i += 256) t += p[i];
    t = block_sum_256(t);
    if (threadIdx.x == 0) sq_err[blockIdx.x] = t;
}

// ---- backward: g_x = g_q + g_loss*coef*(x - q) ---------------------------------------------------
__global__ void vq_backward_kernel(const float* __restrict__ g_q, const float* __restrict__ x,
                                   const float* __restrict__ qv, const float* __restrict__ g_loss,
                                   float coef, int64_t n, float* __restrict__ g_x) {
    const float gl = g_loss ? __ldg(g_loss) * coef : 0.f;
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        float g = g_q ? g_q[i] : 0.f;
        g_x[i] = fmaf(gl, x[i] - qv[i], g);
    }
}

}  // namespace stemgnn

using namespace stemgnn;

extern "C" int stemgnn_vq_backward(const float* g_q, const float* x, const float* q,
                                   const float* g_loss, float coef, int64_t numel, float* g_x,
                                   stemgnn_stream_t stream) {
    SG_CHECK_ARG(numel >= 0, "negative numel");
    if (numel == 0) return STEMGNN_OK;
    SG_CHECK_ARG(x && q && g_x, "NULL pointer");
    int64_t blocks = (numel + 255) / 256;
    int64_t cap = (int64_t)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    SG_LAUNCH(vq_backward_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(g_q, x, q, g_loss, coef,
                                                                           numel, g_x));
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}
