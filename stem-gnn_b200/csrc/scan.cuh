// scan.cuh -- ordered device-wide compaction: item i asks for count(i) >= 0 consecutive output
// slots, emit(i, pos, c) is called with pos = sum of count(j) over j < i.  Three launches (tile
// sums, one-CTA scan of the tile sums, emit); the predicate is evaluated twice instead of being
// written to and read back from HBM.  Order-preserving, no atomics: the output is the one a
// sequential loop (torch boolean-mask indexing, a CPU sampler) produces.
//
// The item range `n` is a launch BOUND: functors may compare against a device-resident size and
// return 0 beyond it, which lets a caller chain stages without reading sizes back to the host.
#pragma once
#include "common.cuh"

namespace stemgnn {
namespace scan {

constexpr int THREADS = 256;
// ITEMS per thread (template parameter): 8 for edge-list sized streams, 1 when the item count is a
// few thousand and an item is heavy (the sampler's per-node subset draw: 5 CTAs of 2,048 nodes took
// 137 us), so that the work spreads over the SMs

// exclusive scan of one int per thread over the 256-thread CTA; *total = CTA sum
__device__ __forceinline__ int block_excl_scan(int v, int* total) {
    __shared__ int wsum[THREADS / 32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    __syncthreads();   // wsum may still be read by the previous call
    if (lane == 31) wsum[w] = inc;
    __syncthreads();
    int base = 0, tot = 0;
#pragma unroll
    for (int i = 0; i < THREADS / 32; ++i) {
        const int s = wsum[i];
        if (i < w) base += s;
        tot += s;
    }
    *total = tot;
    return base + inc - v;
}

template <int ITEMS, class Count>
__global__ void __launch_bounds__(THREADS)
tile_count_kernel(Count count, int64_t n, int32_t* __restrict__ tile_sums) {
    const int64_t base = (int64_t)blockIdx.x * (THREADS * ITEMS);
    int s = 0;
#pragma unroll
    for (int r = 0; r < ITEMS; ++r) {
        const int64_t i = base + (int64_t)r * THREADS + threadIdx.x;
        if (i < n) s += count(i);
    }
    int tot;
    block_excl_scan(s, &tot);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = tot;
}

// one CTA: tile_sums -> exclusive offsets in place; *total_out = grand total
static __global__ void __launch_bounds__(THREADS)
tile_scan_kernel(int32_t* __restrict__ tile_sums, int nt, int64_t* __restrict__ total_out) {
    int carry = 0;
    for (int b0 = 0; b0 < nt; b0 += THREADS) {
        const int b = b0 + threadIdx.x;
        const int v = b < nt ? tile_sums[b] : 0;
        int tot;
        const int ex = block_excl_scan(v, &tot);
        if (b < nt) tile_sums[b] = carry + ex;
        carry += tot;
    }
    if (threadIdx.x == 0 && total_out) *total_out = carry;
}

template <int ITEMS, class Count, class Emit>
__global__ void __launch_bounds__(THREADS)
tile_emit_kernel(Count count, Emit emit, int64_t n, const int32_t* __restrict__ tile_offs) {
    const int64_t base = (int64_t)blockIdx.x * (THREADS * ITEMS);
    int64_t running = tile_offs[blockIdx.x];
#pragma unroll 1
    for (int r = 0; r < ITEMS; ++r) {
        const int64_t i = base + (int64_t)r * THREADS + threadIdx.x;
        const int c = i < n ? count(i) : 0;
        int tot;
        const int ex = block_excl_scan(c, &tot);
        if (c > 0) emit(i, running + ex, c);
        running += tot;
    }
}

static inline int64_t num_tiles(int64_t n, int items) {
    const int64_t tile = (int64_t)THREADS * items;
    return n > 0 ? (n + tile - 1) / tile : 1;
}

// tile_offs: num_tiles(n, ITEMS) int32 of scratch; total_out: device int64 or NULL.  n may be 0.
template <int ITEMS, class Count, class Emit>
static inline void ordered_emit(Count count, Emit emit, int64_t n, int32_t* tile_offs,
                                int64_t* total_out, cudaStream_t stream) {
    const int nt = (int)num_tiles(n, ITEMS);
    SG_LAUNCH(tile_count_kernel<ITEMS><<<nt, THREADS, 0, stream>>>(count, n, tile_offs));
    SG_LAUNCH(tile_scan_kernel<<<1, THREADS, 0, stream>>>(tile_offs, nt, total_out));
    SG_LAUNCH(tile_emit_kernel<ITEMS><<<nt, THREADS, 0, stream>>>(count, emit, n, tile_offs));
}

}  // namespace scan
}  // namespace stemgnn
