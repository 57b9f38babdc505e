// common.cuh -- shared helpers for the stemgnn_b200 CUDA sources (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/stemgnn_b200.h"

namespace stemgnn {

void set_error(const char* fmt, ...);
int sm_count();  // cached per process (device of first use)
// SMs the persistent one-CTA-per-SM kernels (K5, K6: ~200 KB of shared memory and all of TMEM per CTA) size their
// grids by: sm_count() minus what stemgnn_set_sm_reserve left for a collective running beside them.
int persistent_sms();
void count_launch();  // process-wide launch counter behind stemgnn_launch_count()
// Sweep direction of kernels that read a matrix the previous kernel of the step has just written / read front to
// back: walking it back to front finds its tail (up to the 126 MB L2's worth) still cached instead of fetching all
// of it from HBM again.  STEMGNN_L2_PINGPONG=0 switches the reversed sweeps off (A/B); results do not depend on it.
bool l2_pingpong();

// every kernel launch of the library goes through this so the count is a measurement
#define SG_LAUNCH(...) (stemgnn::count_launch(), __VA_ARGS__)

#define SG_CHECK_ARG(cond, msg)                                   \
    do {                                                          \
        if (!(cond)) {                                            \
            stemgnn::set_error("%s: %s", __func__, msg);          \
            return STEMGNN_EINVAL;                                \
        }                                                         \
    } while (0)

#define SG_CHECK_LAUNCH()                                                          \
    do {                                                                           \
        cudaError_t e__ = cudaGetLastError();                                      \
        if (e__ != cudaSuccess) {                                                  \
            stemgnn::set_error("%s: CUDA error %s", __func__, cudaGetErrorString(e__)); \
            return STEMGNN_ECUDA;                                                  \
        }                                                                          \
    } while (0)

#define SG_CHECK_CUDA(expr)                                                        \
    do {                                                                           \
        cudaError_t e__ = (expr);                                                  \
        if (e__ != cudaSuccess) {                                                  \
            stemgnn::set_error("%s: CUDA error %s", __func__, cudaGetErrorString(e__)); \
            return STEMGNN_ECUDA;                                                  \
        }                                                                          \
    } while (0)

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Bump allocator over the caller-provided workspace.
struct Workspace {
    char* base;
    size_t cap, off;
    Workspace(void* p, size_t c) : base((char*)p), cap(c), off(0) {}
    template <typename T>
    T* take(size_t n) {
        size_t bytes = align_up(n * sizeof(T), 256);
        T* r = (T*)(base + off);
        off += bytes;
        return r;
    }
    bool ok() const { return off <= cap && (base != nullptr || off == 0); }
};

// k0_graph.cu: reusable stable radix sort of (key, index) pairs and lower-bound offsets
size_t radix_sort_workspace(int64_t n);
const int32_t* radix_sort_pairs(const int64_t* k64, const int32_t* k32, int64_t n, int64_t key_range,
                                int32_t* vals_out, Workspace& ws, cudaStream_t stream);
void launch_lower_bound(const int32_t* sorted_keys, int64_t n, int64_t n_bins, int32_t* offsets,
                        cudaStream_t stream);

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

}  // namespace stemgnn
