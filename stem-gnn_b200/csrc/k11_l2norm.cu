// k11_l2norm.cu -- K11: row-wise l2 normalisation of the VQ input and its backward.
//
// Replaces `l2norm(x)` = F.normalize(x, p=2, dim=-1) applied per head to the projected embeddings
// in front of the cosine codebook (reference STEM-GNN/model/vq.py:48-49,891; SURVEY.md 8f-1
// "project_in -> l2norm") and its autograd.  torch runs the forward as three kernels and the backward
// as eight elementwise / reduction passes over the [N, H*D] matrix (2.6 ms of a 17.7 ms pretrain
// mini-batch at 43k x 4 x 768); here each direction is one kernel, HBM-bound:
//   forward : y = x / max(||x||, eps), norm saved          2 * R*D*4 B (+ R*4)
//   backward: gx = (g - y * <g, y>) / max(||x||, eps)       3 * R*D*4 B   (gx = g / eps where clamped)
// One warp per row; the second sweep over the row re-reads lines the warp has just pulled into L1.
// Sums run per lane over float4 chunks and then through a fixed shuffle tree: deterministic.
#include "common.cuh"

namespace stemgnn {
namespace l2n {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__global__ void __launch_bounds__(256)
fwd_kernel(const float* __restrict__ x, int64_t rows, int dim4, float eps, float* __restrict__ y,
           float* __restrict__ norm) {
    const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (r >= rows) return;
    const float4* xr = reinterpret_cast<const float4*>(x) + r * dim4;
    float4* yr = reinterpret_cast<float4*>(y) + r * dim4;
    float s = 0.f;
    for (int c = lane; c < dim4; c += 32) {
        const float4 v = xr[c];
        s = fmaf(v.x, v.x, s); s = fmaf(v.y, v.y, s); s = fmaf(v.z, v.z, s); s = fmaf(v.w, v.w, s);
    }
    const float n = sqrtf(warp_sum(s));
    const float d = fmaxf(n, eps);
    if (lane == 0 && norm) norm[r] = n;
    for (int c = lane; c < dim4; c += 32) {
        float4 v = xr[c];
        v.x = __fdiv_rn(v.x, d); v.y = __fdiv_rn(v.y, d); v.z = __fdiv_rn(v.z, d); v.w = __fdiv_rn(v.w, d);
        yr[c] = v;
    }
}

__global__ void __launch_bounds__(256)
bwd_kernel(const float* __restrict__ g, const float* __restrict__ y, const float* __restrict__ norm,
           int64_t rows, int dim4, float eps, float* __restrict__ gx) {
    const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (r >= rows) return;
    const float4* gr = reinterpret_cast<const float4*>(g) + r * dim4;
    const float4* yr = reinterpret_cast<const float4*>(y) + r * dim4;
    float4* o = reinterpret_cast<float4*>(gx) + r * dim4;
    const float n = norm[r];
    const bool clamped = !(n > eps);          // d max(n, eps) / dn = 0: the norm does not depend on x
    float dot = 0.f;
    if (!clamped) {
        for (int c = lane; c < dim4; c += 32) {
            const float4 a = gr[c], b = yr[c];
            dot = fmaf(a.x, b.x, dot); dot = fmaf(a.y, b.y, dot); dot = fmaf(a.z, b.z, dot); dot = fmaf(a.w, b.w, dot);
        }
        dot = warp_sum(dot);
    }
    const float inv = 1.f / fmaxf(n, eps);
    for (int c = lane; c < dim4; c += 32) {
        const float4 a = gr[c], b = yr[c];
        float4 v;
        v.x = (a.x - b.x * dot) * inv; v.y = (a.y - b.y * dot) * inv;
        v.z = (a.z - b.z * dot) * inv; v.w = (a.w - b.w * dot) * inv;
        o[c] = v;
    }
}

}  // namespace l2n
}  // namespace stemgnn

using namespace stemgnn;

static int l2n_check(int64_t rows, int64_t dim, const void* a, const void* b) {
    SG_CHECK_ARG(rows >= 0 && dim >= 0, "negative size");
    SG_CHECK_ARG(dim % 4 == 0 && dim < ((int64_t)1 << 24), "dim must be a multiple of 4");
    SG_CHECK_ARG(rows == 0 || dim == 0 || (a && b), "NULL pointer");
    SG_CHECK_ARG((((uintptr_t)a | (uintptr_t)b) & 15) == 0, "rows must be 16-byte aligned");
    return STEMGNN_OK;
}

extern "C" int stemgnn_l2norm_forward(int64_t rows, int64_t dim, const float* x, float eps, float* y,
                                      float* norm_out, stemgnn_stream_t stream) {
    if (int rc = l2n_check(rows, dim, x, y)) return rc;
    if (rows == 0 || dim == 0) return STEMGNN_OK;
    SG_LAUNCH(l2n::fwd_kernel<<<(unsigned)((rows * 32 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        x, rows, (int)(dim / 4), eps, y, norm_out));
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}

extern "C" int stemgnn_l2norm_backward(int64_t rows, int64_t dim, const float* g, const float* y,
                                       const float* norm, float eps, float* gx, stemgnn_stream_t stream) {
    if (int rc = l2n_check(rows, dim, g, gx)) return rc;
    if (rows == 0 || dim == 0) return STEMGNN_OK;
    SG_CHECK_ARG(y && norm && (((uintptr_t)y) & 15) == 0, "y / norm NULL or misaligned");
    SG_LAUNCH(l2n::bwd_kernel<<<(unsigned)((rows * 32 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        g, y, norm, rows, (int)(dim / 4), eps, gx));
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}
