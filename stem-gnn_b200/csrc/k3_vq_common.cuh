// k3_vq_common.cuh -- shared view struct + the exact scoring formula of the reference.
#pragma once
#include <math.h>

#include "common.cuh"

namespace stemgnn {

struct VqView {
    const float* x; int64_t x_row_stride, x_head_stride;
    const float* embed;
    int64_t heads, embed_heads, rows;
    int K, D, metric;
    int64_t* ind; int64_t ind_row_stride, ind_head_stride;
    float* quantize; int64_t q_row_stride, q_head_stride;
};

// metric 0 (model/vq.py:31-35,472): -sqrt(clamp((x2 + e2) + (-2 * x.e), 0)); metric 1
// (model/vq.py:650): x.e.  `dot` is the fp32 inner product.
__device__ __forceinline__ float vq_score(int metric, float x2, float e2, float dot) {
    if (metric != 0) return dot;
    float t = __fadd_rn(__fadd_rn(x2, e2), __fmul_rn(dot, -2.0f));
    t = t > 0.f ? t : 0.f;
    return -__fsqrt_rn(t);
}

// kernels defined in k3_vq_simt.cu, used by both paths
__global__ void vq_sqnorm_kernel(const float* v, int64_t rows, int64_t heads, int D,
                                 int64_t row_stride, int64_t head_stride, float* out);
__global__ void vq_scan_kernel(VqView q, const float* x2, const float* e2, const int32_t* row_list,
                               const int32_t* row_count, int32_t* best_out);
__global__ void vq_gather_kernel(VqView q, int32_t* best, const int32_t* cand, const float* e2,
                                 double* partials, int32_t* open_list, int32_t* open_count,
                                 unsigned long long* stats, int rev);
__global__ void vq_rescan_kernel(VqView q, int32_t* best, const float* e2, double* partials, int nblk,
                                 const int32_t* open_list, const int32_t* open_count,
                                 unsigned long long* stats);
constexpr int VQ_GATHER_ROWS = 4;    // rows per gather WARP = per commitment partial / open-list entry
constexpr int VQ_SQ_CHUNKS = 128;    // first-stage CTAs of the commitment reduction
constexpr int VQ_CAND_CAP = 8;   // candidate slots per row written by the tcgen05 path
__global__ void vq_sqerr_chunk_kernel(const double* partials, int nblk, double* chunk_sums);
__global__ void vq_sqerr_reduce_kernel(const double* chunk_sums, int nchunks, double* sq_err);

// path 2 (k3_vq_tc.cu): returns 0 on success
bool vq_tc_supported(const VqView& q);
int vq_assign_tc_workspace(const VqView& q, size_t* bytes);
int vq_assign_tc(const VqView& q, const float* e2, int32_t* best, int32_t** cand_out, void* ws,
                 size_t ws_bytes, cudaStream_t stream);

}  // namespace stemgnn
