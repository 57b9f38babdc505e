// k7_bn_act.cu -- K7: BatchNorm1d + ReLU + dropout of the encoder layers as three passes over the
// activations instead of torch's seven and more.
//
// Replaces `z = self.norms[i](z); z = self.dropout(self.activation(z))` (reference
// STEM-GNN/model/encoder.py:313-317) and its autograd:
//
//   forward : column statistics (one read)  ->  y = drop(relu((x - mean) * invstd * gamma + beta))
//             (one read, one write; the dropout mask is a counter-based Philox4x32-7 stream keyed
//              by a device-resident seed and the element index, so it is never stored)
//   backward: g = g_y * mask/(1-p) * [pre > 0];  sums of g and g * xhat per column (reads g_y, x)
//             ->  g_x = gamma * invstd * (g - sum_g/N - xhat * sum_gxhat/N)  (reads g_y, x, writes g_x)
//
// Statistics are shifted sums (shift = row 0 of the column) accumulated in a fixed order:
// deterministic, no atomics.  HBM-bound: N*F*4 B per pass.
#include "common.cuh"
#include "philox.cuh"

namespace stemgnn {
namespace bn {

struct Cols {                 // per-column constants, all [F]
    const float* mean; const float* invstd; const float* gamma; const float* beta;
};

// keep-mask scale factors of the 4 elements of float4 number `idx4`: 1/(1-p) or 0
__device__ __forceinline__ float4 drop_scale4(uint64_t seed, int64_t idx4, float p, float inv_keep) {
    const uint4 r = philox4x32_7(make_uint4((uint32_t)idx4, (uint32_t)((uint64_t)idx4 >> 32), 0u, 0u),
                                  make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
    const float s = 1.f / 16777216.f;
    float4 m;
    m.x = (float)(r.x >> 8) * s >= p ? inv_keep : 0.f;
    m.y = (float)(r.y >> 8) * s >= p ? inv_keep : 0.f;
    m.z = (float)(r.z >> 8) * s >= p ? inv_keep : 0.f;
    m.w = (float)(r.w >> 8) * s >= p ? inv_keep : 0.f;
    return m;
}

__device__ __forceinline__ float4 ld4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

// ---- column reductions: block = 64 column units (256 columns) x 4 row lanes -----------------------
// MODE 0: sums of (x - k), (x - k)^2 with k = x[0, c]        (forward statistics)
// MODE 1: sums of g, g * xhat with g = g_y * mask * relu'     (backward)
template <int MODE>
__global__ void __launch_bounds__(256)
col_reduce_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ gy, int64_t ldg,
                  int64_t rows, int feat, int64_t rows_per_block, Cols cc, int relu, float p, float inv_keep,
                  const uint64_t* __restrict__ seed_p, float* __restrict__ part, int rev) {
    __shared__ float4 sh[2][4][64];
    const int cu = threadIdx.x & 63, rl = threadIdx.x >> 6;
    // rev: the CTAs that start first take the LAST row ranges -- the tail of the matrix the previous kernel wrote
    // front to back is still in L2 (l2_pingpong); the partial of a range keeps its place, so the sums do not change
    const int bid = rev ? (int)gridDim.x - 1 - (int)blockIdx.x : (int)blockIdx.x;
    const int64_t r0 = (int64_t)bid * rows_per_block;
    const int64_t r1 = r0 + rows_per_block < rows ? r0 + rows_per_block : rows;
    const uint64_t seed = (MODE == 1 && p > 0.f) ? *seed_p : 0;
    const int f4 = feat >> 2;
    for (int c0 = 0; c0 < feat; c0 += 256) {
        const int c = c0 + 4 * cu;
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
        if (c < feat) {
            float4 k = make_float4(0.f, 0.f, 0.f, 0.f), mu = k, is = k, ga = k, be = k;
            if (MODE == 0) k = ld4(x + c);
            else { mu = ld4(cc.mean + c); is = ld4(cc.invstd + c); ga = ld4(cc.gamma + c); be = ld4(cc.beta + c); }
            // U rows per trip: all loads of a trip are issued before the first add (the kernel is bound by
            // memory latency x loads in flight: 2 in flight gave 2.8 TB/s)
            constexpr int U = MODE == 0 ? 8 : 4;
            auto consume = [&](const float4 v, float4 g, const int64_t r) {
                if (MODE == 0) {
                    const float dx = v.x - k.x, dy = v.y - k.y, dz = v.z - k.z, dw = v.w - k.w;
                    a.x += dx; a.y += dy; a.z += dz; a.w += dw;
                    b.x = fmaf(dx, dx, b.x); b.y = fmaf(dy, dy, b.y); b.z = fmaf(dz, dz, b.z); b.w = fmaf(dw, dw, b.w);
                } else {
                    const float hx = (v.x - mu.x) * is.x, hy = (v.y - mu.y) * is.y, hz = (v.z - mu.z) * is.z, hw = (v.w - mu.w) * is.w;
                    if (p > 0.f) {
                        const float4 m = drop_scale4(seed, r * f4 + (c >> 2), p, inv_keep);
                        g.x *= m.x; g.y *= m.y; g.z *= m.z; g.w *= m.w;
                    }
                    if (relu) {
                        g.x = fmaf(hx, ga.x, be.x) > 0.f ? g.x : 0.f; g.y = fmaf(hy, ga.y, be.y) > 0.f ? g.y : 0.f;
                        g.z = fmaf(hz, ga.z, be.z) > 0.f ? g.z : 0.f; g.w = fmaf(hw, ga.w, be.w) > 0.f ? g.w : 0.f;
                    }
                    a.x += g.x; a.y += g.y; a.z += g.z; a.w += g.w;
                    b.x = fmaf(g.x, hx, b.x); b.y = fmaf(g.y, hy, b.y); b.z = fmaf(g.z, hz, b.z); b.w = fmaf(g.w, hw, b.w);
                }
            };
            int64_t r = r0 + rl;
            for (; r + 4 * (U - 1) < r1; r += 4 * U) {
                float4 v[U], g[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    v[u] = ld4(x + (r + 4 * u) * ldx + c);
                    g[u] = MODE == 1 ? ld4(gy + (r + 4 * u) * ldg + c) : v[u];
                }
#pragma unroll
                for (int u = 0; u < U; ++u) consume(v[u], g[u], r + 4 * u);
            }
            for (; r < r1; r += 4) {
                const float4 v = ld4(x + r * ldx + c);
                consume(v, MODE == 1 ? ld4(gy + r * ldg + c) : v, r);
            }
        }
        sh[0][rl][cu] = a; sh[1][rl][cu] = b;
        __syncthreads();
        if (rl < 2 && c < feat) {       // row lane 0 finishes the first sum, lane 1 the second
            const float4 s0 = sh[rl][0][cu], s1 = sh[rl][1][cu], s2 = sh[rl][2][cu], s3 = sh[rl][3][cu];
            float* o = part + ((int64_t)bid * 2 + rl) * feat + c;
            o[0] = (s0.x + s1.x) + (s2.x + s3.x); o[1] = (s0.y + s1.y) + (s2.y + s3.y);
            o[2] = (s0.z + s1.z) + (s2.z + s3.z); o[3] = (s0.w + s1.w) + (s2.w + s3.w);
        }
        __syncthreads();
    }
}

// sums the block partials per column in block order.  MODE 0 writes mean and biased variance,
// MODE 1 writes the two sums.
// 32 warps per 32 columns: with ~444 block partials a warp sums 14 of them, all 28 loads in flight at once (8
// warps walked 55 partials each in dependent batches: 14 us per launch, six launches per bench step).
constexpr int CF_WARPS = 32;
// What nn.BatchNorm1d does with the batch statistics in training, as the tail of the statistics kernel instead of ten
// [F]-sized torch kernels per layer: invstd = rsqrt(var + eps); running_mean = running_mean * (1 - m) + mean * m;
// running_var = running_var * (1 - m) + var * n / (n - 1) * m -- every product and sum rounded separately like the
// tensor ops they replace; m = momentum, or 1 / num_batches_tracked (cumulative average, momentum None) read from
// the module's device counter (the caller increments it first).
struct TrainTail {
    float* invstd; float eps;
    float* running_mean; float* running_var;   // NULL: no running statistics
    float momentum, keep;                      // m and 1 - m (rounded from the double difference); m < 0: cumulative average
    const int64_t* num_batches_tracked;        // device counter (already incremented by the caller) or NULL
    float unbias;                              // n / max(n - 1, 1)
};
template <int MODE>
__global__ void __launch_bounds__(CF_WARPS * 32)
col_final_kernel(const float* __restrict__ part, int blocks, int feat, int64_t rows, const float* __restrict__ x,
                 float* __restrict__ out0, float* __restrict__ out1, const TrainTail tt) {
    __shared__ float sh[2][CF_WARPS][32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + lane;
    float a = 0.f, b = 0.f;
    if (c < feat) {
#pragma unroll 16
        for (int i = w; i < blocks; i += CF_WARPS) { a += part[((int64_t)i * 2) * feat + c]; b += part[((int64_t)i * 2 + 1) * feat + c]; }
    }
    sh[0][w][lane] = a; sh[1][w][lane] = b;
    __syncthreads();
    if (w == 0 && c < feat) {
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int i = 0; i < CF_WARPS; ++i) { s1 += sh[0][i][lane]; s2 += sh[1][i][lane]; }
        if (MODE == 0) {
            const float n = (float)rows, m1 = s1 / n;
            const float mean = __ldg(x + c) + m1, var = fmaxf(s2 / n - m1 * m1, 0.f);
            out0[c] = mean;
            out1[c] = var;
            if (tt.invstd) {
                tt.invstd[c] = rsqrtf(__fadd_rn(var, tt.eps));
                if (tt.running_mean) {
                    const bool cma = tt.momentum < 0.f;
                    const float m = cma ? __fdiv_rn(1.0f, (float)*tt.num_batches_tracked) : tt.momentum;
                    const float keep = cma ? __fsub_rn(1.0f, m) : tt.keep;
                    tt.running_mean[c] = __fadd_rn(__fmul_rn(tt.running_mean[c], keep), __fmul_rn(mean, m));
                    tt.running_var[c] = __fadd_rn(__fmul_rn(tt.running_var[c], keep),
                                                  __fmul_rn(__fmul_rn(var, tt.unbias), m));
                }
            }
        } else {
            out0[c] = s1; out1[c] = s2;
        }
    }
}

// ---- elementwise passes --------------------------------------------------------------------------
// BWD = false: out = drop(relu(xhat * gamma + beta))
// BWD = true : out = gamma * invstd * (g - sg/N - xhat * sgx/N)   (batch statistics)
//              out = gamma * invstd * g                            (running statistics: sg = sgx = NULL)
template <bool BWD>
__global__ void __launch_bounds__(256)
apply_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ gy, int64_t ldg, int64_t rows, int feat,
             Cols cc, const float* __restrict__ sg, const float* __restrict__ sgx, int relu, float p, float inv_keep,
             const uint64_t* __restrict__ seed_p, float* __restrict__ out, int64_t ldo, int64_t batch_rows) {
    const int f4 = feat >> 2;
    const int64_t total = rows * f4;
    const uint64_t seed = p > 0.f ? *seed_p : 0;
    // batch_rows: rows the statistics (and sg / sgx) were taken over -- all ranks' rows when the batch
    // is split over several processes, else this call's rows
    const float inv_n = 1.f / (float)(batch_rows > 0 ? batch_rows : rows);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / f4;
        const int c = (int)(i - r * f4) * 4;
        const float4 v = ld4(x + r * ldx + c);
        const float4 mu = ld4(cc.mean + c), is = ld4(cc.invstd + c), ga = ld4(cc.gamma + c), be = ld4(cc.beta + c);
        const float hx = (v.x - mu.x) * is.x, hy = (v.y - mu.y) * is.y, hz = (v.z - mu.z) * is.z, hw = (v.w - mu.w) * is.w;
        float4 m = make_float4(1.f, 1.f, 1.f, 1.f);
        if (p > 0.f) m = drop_scale4(seed, i, p, inv_keep);
        float4 o;
        if (!BWD) {
            o.x = fmaf(hx, ga.x, be.x); o.y = fmaf(hy, ga.y, be.y); o.z = fmaf(hz, ga.z, be.z); o.w = fmaf(hw, ga.w, be.w);
            if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
            o.x *= m.x; o.y *= m.y; o.z *= m.z; o.w *= m.w;
        } else {
            float4 g = ld4(gy + r * ldg + c);
            g.x *= m.x; g.y *= m.y; g.z *= m.z; g.w *= m.w;
            if (relu) {
                g.x = fmaf(hx, ga.x, be.x) > 0.f ? g.x : 0.f; g.y = fmaf(hy, ga.y, be.y) > 0.f ? g.y : 0.f;
                g.z = fmaf(hz, ga.z, be.z) > 0.f ? g.z : 0.f; g.w = fmaf(hw, ga.w, be.w) > 0.f ? g.w : 0.f;
            }
            if (sg) {
                const float4 a = ld4(sg + c), b = ld4(sgx + c);
                g.x -= fmaf(hx, b.x * inv_n, a.x * inv_n); g.y -= fmaf(hy, b.y * inv_n, a.y * inv_n);
                g.z -= fmaf(hz, b.z * inv_n, a.z * inv_n); g.w -= fmaf(hw, b.w * inv_n, a.w * inv_n);
            }
            o.x = ga.x * is.x * g.x; o.y = ga.y * is.y * g.y; o.z = ga.z * is.z * g.z; o.w = ga.w * is.w * g.w;
        }
        *reinterpret_cast<float4*>(out + r * ldo + c) = o;
    }
}

}  // namespace bn
}  // namespace stemgnn

using namespace stemgnn;

namespace {
struct BnPlan { int blocks; int64_t rows_per_block; size_t bytes; };
BnPlan bn_plan(int64_t rows, int64_t feat) {
    BnPlan pl;
    const int64_t want = (int64_t)sm_count() * 3;     // 8 x 16 B loads in flight per thread: few fat blocks, few partials
    int64_t rpb = (rows + want - 1) / want;
    rpb = rpb < 64 ? 64 : (rpb + 3) / 4 * 4;
    pl.rows_per_block = rpb;
    pl.blocks = (int)((rows + rpb - 1) / rpb);
    pl.bytes = (size_t)pl.blocks * 2 * feat * 4 + 256;
    return pl;
}
int bn_check(int64_t rows, int64_t feat, const void* x, int64_t ldx) {
    SG_CHECK_ARG(rows >= 0 && feat >= 0, "negative size");
    SG_CHECK_ARG(feat % 4 == 0 && feat < (1 << 24), "feat must be a multiple of 4");
    SG_CHECK_ARG(rows == 0 || feat == 0 || (x && ldx >= feat && ldx % 4 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0),
                 "x is NULL, misaligned, or its row stride is not a multiple of 4 floats");
    return STEMGNN_OK;
}
unsigned ew_grid(int64_t units) {
    const int64_t want = (units + 255) / 256, cap = (int64_t)sm_count() * 16;
    return (unsigned)(want < cap ? (want > 0 ? want : 1) : cap);
}
}  // namespace

extern "C" int stemgnn_bn_workspace(int64_t rows, int64_t feat, size_t* bytes) {
    SG_CHECK_ARG(bytes != nullptr && rows >= 0 && feat >= 0, "bad arguments");
    *bytes = bn_plan(rows, feat).bytes;
    return STEMGNN_OK;
}

extern "C" int stemgnn_bn_stats(int64_t rows, int64_t feat, const float* x, int64_t ldx, float* mean, float* var,
                                void* workspace, size_t workspace_bytes, stemgnn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (int rc = bn_check(rows, feat, x, ldx)) return rc;
    SG_CHECK_ARG(rows > 0, "batch statistics of an empty batch");
    if (feat == 0) return STEMGNN_OK;
    const BnPlan pl = bn_plan(rows, feat);
    SG_CHECK_ARG(mean && var && workspace && workspace_bytes >= pl.bytes, "mean/var NULL or workspace too small");
    float* part = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~(uintptr_t)255);
    bn::Cols none{nullptr, nullptr, nullptr, nullptr};
    SG_LAUNCH(bn::col_reduce_kernel<0><<<pl.blocks, 256, 0, stream>>>(x, ldx, nullptr, 0, rows, (int)feat, pl.rows_per_block,
                                                                      none, 0, 0.f, 1.f, nullptr, part, l2_pingpong() ? 1 : 0));
    SG_LAUNCH(bn::col_final_kernel<0><<<(unsigned)((feat + 31) / 32), bn::CF_WARPS * 32, 0, stream>>>(part, pl.blocks, (int)feat, rows, x, mean, var, bn::TrainTail{}));
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}

extern "C" int stemgnn_bn_train_stats(int64_t rows, int64_t feat, const float* x, int64_t ldx, float eps,
                                      double momentum, const int64_t* num_batches_tracked, float* running_mean,
                                      float* running_var, float* mean, float* var, float* invstd, void* workspace,
                                      size_t workspace_bytes, stemgnn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (int rc = bn_check(rows, feat, x, ldx)) return rc;
    SG_CHECK_ARG(rows > 1, "Expected more than 1 value per channel when training");
    if (feat == 0) return STEMGNN_OK;
    const BnPlan pl = bn_plan(rows, feat);
    SG_CHECK_ARG(mean && var && invstd && workspace && workspace_bytes >= pl.bytes, "mean/var/invstd NULL or workspace too small");
    SG_CHECK_ARG((running_mean == nullptr) == (running_var == nullptr), "running_mean and running_var go together");
    SG_CHECK_ARG(!(running_mean && momentum < 0.0 && !num_batches_tracked), "cumulative average needs num_batches_tracked");
    float* part = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~(uintptr_t)255);
    bn::Cols none{nullptr, nullptr, nullptr, nullptr};
    bn::TrainTail tt{invstd, eps, running_mean, running_var, (float)momentum, (float)(1.0 - momentum), num_batches_tracked,
                     (float)((double)rows / (double)(rows - 1))};
    SG_LAUNCH(bn::col_reduce_kernel<0><<<pl.blocks, 256, 0, stream>>>(x, ldx, nullptr, 0, rows, (int)feat, pl.rows_per_block,
                                                                      none, 0, 0.f, 1.f, nullptr, part, l2_pingpong() ? 1 : 0));
    SG_LAUNCH(bn::col_final_kernel<0><<<(unsigned)((feat + 31) / 32), bn::CF_WARPS * 32, 0, stream>>>(part, pl.blocks, (int)feat, rows, x, mean, var, tt));
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}

extern "C" int stemgnn_bn_act_forward(int64_t rows, int64_t feat, const float* x, int64_t ldx, const float* mean,
                                      const float* invstd, const float* gamma, const float* beta, int relu, float p,
                                      const uint64_t* seed, float* out, int64_t ldo, stemgnn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (int rc = bn_check(rows, feat, x, ldx)) return rc;
    if (rows == 0 || feat == 0) return STEMGNN_OK;
    SG_CHECK_ARG(mean && invstd && gamma && beta && out && ldo >= feat && ldo % 4 == 0, "NULL column vector / out");
    SG_CHECK_ARG(p >= 0.f && p < 1.f && (p == 0.f || seed), "dropout p must be in [0, 1) and needs a seed");
    bn::Cols cc{mean, invstd, gamma, beta};
    SG_LAUNCH(bn::apply_kernel<false><<<ew_grid(rows * (feat / 4)), 256, 0, stream>>>(
        x, ldx, nullptr, 0, rows, (int)feat, cc, nullptr, nullptr, relu, p, 1.f / (1.f - p), seed, out, ldo, 0));
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}

extern "C" int stemgnn_bn_act_backward(int64_t rows, int64_t feat, const float* x, int64_t ldx, const float* gy, int64_t ldg,
                                       const float* mean, const float* invstd, const float* gamma, const float* beta,
                                       int relu, float p, const uint64_t* seed, int batch_stats, float* sum_g,
                                       float* sum_gx, float* gx, int64_t ldgx, void* workspace, size_t workspace_bytes,
                                       stemgnn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (int rc = bn_check(rows, feat, x, ldx)) return rc;
    if (int rc = bn_check(rows, feat, gy, ldg)) return rc;
    if (feat == 0) return STEMGNN_OK;
    SG_CHECK_ARG(mean && invstd && gamma && beta && sum_g && sum_gx, "NULL column vector");
    SG_CHECK_ARG(p >= 0.f && p < 1.f && (p == 0.f || seed), "dropout p must be in [0, 1) and needs a seed");
    if (rows == 0) {
        SG_CHECK_CUDA(cudaMemsetAsync(sum_g, 0, feat * 4, stream));
        SG_CHECK_CUDA(cudaMemsetAsync(sum_gx, 0, feat * 4, stream));
        return STEMGNN_OK;
    }
    const BnPlan pl = bn_plan(rows, feat);
    SG_CHECK_ARG(workspace && workspace_bytes >= pl.bytes, "workspace too small (stemgnn_bn_workspace)");
    float* part = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~(uintptr_t)255);
    bn::Cols cc{mean, invstd, gamma, beta};
    const float inv_keep = 1.f / (1.f - p);
    SG_LAUNCH(bn::col_reduce_kernel<1><<<pl.blocks, 256, 0, stream>>>(x, ldx, gy, ldg, rows, (int)feat, pl.rows_per_block, cc,
                                                                      relu, p, inv_keep, seed, part, l2_pingpong() ? 1 : 0));
    SG_LAUNCH(bn::col_final_kernel<1><<<(unsigned)((feat + 31) / 32), bn::CF_WARPS * 32, 0, stream>>>(part, pl.blocks, (int)feat, rows, x, sum_g, sum_gx, bn::TrainTail{}));
    if (gx) {
        SG_CHECK_ARG(ldgx >= feat && ldgx % 4 == 0, "ldgx");
        SG_LAUNCH(bn::apply_kernel<true><<<ew_grid(rows * (feat / 4)), 256, 0, stream>>>(
            x, ldx, gy, ldg, rows, (int)feat, cc, batch_stats ? sum_g : nullptr, batch_stats ? sum_gx : nullptr, relu, p,
            inv_keep, seed, gx, ldgx, 0));
    }
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}

extern "C" int stemgnn_bn_act_backward_apply(int64_t rows, int64_t feat, const float* x, int64_t ldx, const float* gy,
                                             int64_t ldg, const float* mean, const float* invstd, const float* gamma,
                                             const float* beta, int relu, float p, const uint64_t* seed,
                                             const float* sum_g, const float* sum_gx, int64_t batch_rows, float* gx,
                                             int64_t ldgx, stemgnn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (int rc = bn_check(rows, feat, x, ldx)) return rc;
    if (int rc = bn_check(rows, feat, gy, ldg)) return rc;
    if (feat == 0 || rows == 0) return STEMGNN_OK;
    SG_CHECK_ARG(mean && invstd && gamma && beta && sum_g && sum_gx && gx, "NULL pointer");
    SG_CHECK_ARG(ldgx >= feat && ldgx % 4 == 0, "ldgx");
    SG_CHECK_ARG(batch_rows >= rows, "batch_rows must cover this call's rows");
    SG_CHECK_ARG(p >= 0.f && p < 1.f && (p == 0.f || seed), "dropout p must be in [0, 1) and needs a seed");
    bn::Cols cc{mean, invstd, gamma, beta};
    SG_LAUNCH(bn::apply_kernel<true><<<ew_grid(rows * (feat / 4)), 256, 0, stream>>>(
        x, ldx, gy, ldg, rows, (int)feat, cc, sum_g, sum_gx, relu, p, 1.f / (1.f - p), seed, gx, ldgx, batch_rows));
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}

// ---- stemgnn_bn_stats_static: the same statistics with the block partials in library-owned memory ----
// For callers that cannot allocate (CUDA graph capture, serving loops, harnesses without cudaMalloc).
// The scratch is a file-scope __device__ array: every GPU that loads the library holds its 4 MiB until
// the process exits.  In host code `bn_static_scratch` is only a host-side stand-in for the symbol, not
// a device address, so the address each device gave it is looked up once per device (with that device
// current) through cudaGetSymbolAddress and that pointer is what the kernels get.
#define STEMGNN_BN_STATIC_SCRATCH_BYTES (4u << 20)
static __device__ __align__(256) float bn_static_scratch[STEMGNN_BN_STATIC_SCRATCH_BYTES / sizeof(float)];

namespace {
constexpr int kMaxDevices = 64;
float* bn_static_ptr[kMaxDevices];   // device address of bn_static_scratch, per device; written once
int bn_static_scratch_address(float** out) {
    int dev = -1;
    SG_CHECK_CUDA(cudaGetDevice(&dev));
    SG_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "device ordinal beyond 63");
    float* p = __atomic_load_n(&bn_static_ptr[dev], __ATOMIC_ACQUIRE);
    if (p == nullptr) {
        // set-up, once per device, the device being current; two threads racing here get the same address
        SG_CHECK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), bn_static_scratch));
        SG_CHECK_ARG(p != nullptr, "cudaGetSymbolAddress returned NULL");
        __atomic_store_n(&bn_static_ptr[dev], p, __ATOMIC_RELEASE);
    }
    *out = p;
    return STEMGNN_OK;
}
}  // namespace

extern "C" int stemgnn_bn_stats_static(int64_t rows, int64_t feat, const float* x, int64_t ldx, float* mean,
                                       float* var, stemgnn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (int rc = bn_check(rows, feat, x, ldx)) return rc;
    SG_CHECK_ARG(rows > 0, "batch statistics of an empty batch");
    if (feat == 0) return STEMGNN_OK;
    SG_CHECK_ARG(mean && var, "mean/var NULL");
    const BnPlan pl = bn_plan(rows, feat);
    const size_t need = (size_t)pl.blocks * 2 * (size_t)feat * sizeof(float);
    if (need > STEMGNN_BN_STATIC_SCRATCH_BYTES) {
        set_error("stemgnn_bn_stats_static: %d row blocks x 2 x %lld columns need %zu bytes of block partials, the "
                  "built-in scratch holds %u (feat <= %lld at this row count); use stemgnn_bn_stats with a caller "
                  "workspace", pl.blocks, (long long)feat, need, (unsigned)STEMGNN_BN_STATIC_SCRATCH_BYTES,
                  (long long)(STEMGNN_BN_STATIC_SCRATCH_BYTES / (8 * (size_t)pl.blocks) / 4 * 4));
        return STEMGNN_EINVAL;
    }
    float* part = nullptr;
    if (int rc = bn_static_scratch_address(&part)) return rc;
    // nothing in the scratch has to start at zero: col_reduce writes every partial col_final reads
    bn::Cols none{nullptr, nullptr, nullptr, nullptr};
    SG_LAUNCH(bn::col_reduce_kernel<0><<<pl.blocks, 256, 0, stream>>>(x, ldx, nullptr, 0, rows, (int)feat, pl.rows_per_block,
                                                                      none, 0, 0.f, 1.f, nullptr, part, l2_pingpong() ? 1 : 0));
    SG_LAUNCH(bn::col_final_kernel<0><<<(unsigned)((feat + 31) / 32), bn::CF_WARPS * 32, 0, stream>>>(part, pl.blocks, (int)feat, rows, x, mean, var, bn::TrainTail{}));
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}
