// k0_graph.cu -- K0: COO -> CSR with a hand-written STABLE LSD radix sort (8-bit digits).
//
// Replaces the implicit scatter order of MessagePassing.propagate / scatter_mean
// (reference STEM-GNN/model/encoder.py:82,124): grouping edges by their aggregation row with a
// stable sort keeps "original edge order inside a row", the order scatter_add_ sums in, so the
// pull-based aggregation kernels can be bit-exact against the oracle.
//
// Layout in HBM: keys (int32 node id) and values (int32 edge id) ping-pong between two
// workspace buffers; the last pass writes the values straight into `perm`.  Each pass is
// histogram -> per-digit scan over tiles -> stable scatter.  Integer work, HBM-bound:
// per pass E*(4+4) B read + E*(4+4) B written (+ E*8 B in pass 0 for the int64 input row).
#include "common.cuh"
#include <cstdlib>

namespace stemgnn {

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_ITEMS = 16;                       // items per thread
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;     // 4096 items per CTA
constexpr int RS_WARP_ITEMS = 32 * RS_ITEMS;       // 512 items per warp, contiguous
// Digit width: 8 bits (256 bins), or 9 bits (512 bins) when that saves a pass -- 169,343 nodes are 18
// bits: two passes of 9 instead of 8 + 8 + 2; the 512 codes of the arxiv-shaped VQ: one pass.
constexpr int RS_MAX_BINS = 512;

// key of item i for this pass: either from the int64 edge_index row (pass 0) or the int32 buffer
__device__ __forceinline__ int load_key(const int64_t* k64, const int32_t* k32, int64_t i) {
    return k64 ? (int)k64[i] : k32[i];
}

template <int BINS>
__global__ void __launch_bounds__(RS_THREADS)
radix_hist_kernel(const int64_t* __restrict__ k64, const int32_t* __restrict__ k32, int64_t n,
                  int shift, int nblk, int32_t* __restrict__ block_hist /* [BINS][nblk] */) {
    __shared__ int hist[BINS];
    for (int b = threadIdx.x; b < BINS; b += RS_THREADS) hist[b] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * RS_TILE;
#pragma unroll 4
    for (int r = 0; r < RS_ITEMS; ++r) {
        int64_t i = base + (int64_t)r * RS_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&hist[(load_key(k64, k32, i) >> shift) & (BINS - 1)], 1);
    }
    __syncthreads();
    for (int b = threadIdx.x; b < BINS; b += RS_THREADS) block_hist[(int64_t)b * nblk + blockIdx.x] = hist[b];
}

__device__ __forceinline__ int block_excl_scan_256(int v, int* total_out) {
    // exclusive scan of one int per thread across a 256-thread CTA
    __shared__ int wsum[RS_WARPS];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    __syncthreads();  // protect wsum reuse across calls
    if (lane == 31) wsum[w] = inc;
    __syncthreads();
    int wbase = 0, tot = 0;
#pragma unroll
    for (int i = 0; i < RS_WARPS; ++i) {
        int s = wsum[i];
        if (i < w) wbase += s;
        tot += s;
    }
    *total_out = tot;
    return wbase + inc - v;
}

// One CTA per digit: exclusive scan of that digit's per-tile counts (in place) + digit total.
__global__ void __launch_bounds__(RS_THREADS)
radix_scan_kernel(int32_t* __restrict__ block_hist, int nblk, int32_t* __restrict__ digit_total) {
    int32_t* row = block_hist + (int64_t)blockIdx.x * nblk;
    int carry = 0;
    for (int b0 = 0; b0 < nblk; b0 += RS_THREADS) {
        int b = b0 + threadIdx.x;
        int v = b < nblk ? row[b] : 0;
        int tot;
        int ex = block_excl_scan_256(v, &tot);
        if (b < nblk) row[b] = carry + ex;
        carry += tot;
    }
    if (threadIdx.x == 0) digit_total[blockIdx.x] = carry;
}

// Stable scatter.  Item order inside a CTA is (warp, round, lane) = ascending index.
template <int BINS>
__global__ void __launch_bounds__(RS_THREADS)
radix_scatter_kernel(const int64_t* __restrict__ k64, const int32_t* __restrict__ k32,
                     const int32_t* __restrict__ vin /* NULL => identity */, int64_t n, int shift,
                     int nblk, const int32_t* __restrict__ block_hist,
                     const int32_t* __restrict__ digit_total, int32_t* __restrict__ kout,
                     int32_t* __restrict__ vout) {
    constexpr int PER = BINS / RS_THREADS;          // digits per thread in the digit-total scan (1 or 2)
    __shared__ int wcount[RS_WARPS][BINS];
    __shared__ int blockbase[BINS];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < RS_WARPS * BINS; i += RS_THREADS) (&wcount[0][0])[i] = 0;

    // global base of each digit = exclusive scan of digit totals, plus this tile's offset
    {
        int d[PER], sum = 0;
#pragma unroll
        for (int j = 0; j < PER; ++j) { d[j] = digit_total[threadIdx.x * PER + j]; sum += d[j]; }
        int tot;
        int ex = block_excl_scan_256(sum, &tot);
#pragma unroll
        for (int j = 0; j < PER; ++j) {
            const int b = threadIdx.x * PER + j;
            blockbase[b] = ex + block_hist[(int64_t)b * nblk + blockIdx.x];
            ex += d[j];
        }
    }
    __syncthreads();

    const int64_t wbase = (int64_t)blockIdx.x * RS_TILE + (int64_t)w * RS_WARP_ITEMS;
    int keys[RS_ITEMS];
    int vals[RS_ITEMS];
    unsigned short ranks[RS_ITEMS];
#pragma unroll
    for (int r = 0; r < RS_ITEMS; ++r) {
        int64_t i = wbase + r * 32 + lane;
        bool valid = i < n;
        keys[r] = valid ? load_key(k64, k32, i) : 0;
        vals[r] = valid ? (vin ? vin[i] : (int)i) : 0;
    }
    const unsigned lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < RS_ITEMS; ++r) {
        int64_t i = wbase + r * 32 + lane;
        bool valid = i < n;
        int d = valid ? ((keys[r] >> shift) & (BINS - 1)) : BINS + lane;  // invalid: unique, no peers
        unsigned peers = __match_any_sync(0xffffffffu, d);
        int leader = __ffs(peers) - 1;
        int old = 0;
        if (valid && lane == leader) {
            old = wcount[w][d];
            wcount[w][d] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        ranks[r] = (unsigned short)(old + __popc(peers & lt_mask));
        __syncwarp();
    }
    __syncthreads();
    for (int b = threadIdx.x; b < BINS; b += RS_THREADS) {   // exclusive prefix over warps per digit
        int run = 0;
#pragma unroll
        for (int i = 0; i < RS_WARPS; ++i) {
            int t = wcount[i][b];
            wcount[i][b] = run;
            run += t;
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < RS_ITEMS; ++r) {
        int64_t i = wbase + r * 32 + lane;
        if (i < n) {
            int d = (keys[r] >> shift) & (BINS - 1);
            int64_t pos = (int64_t)blockbase[d] + wcount[w][d] + ranks[r];
            kout[pos] = keys[r];
            vout[pos] = vals[r];
        }
    }
}

// rowptr[i] = lower_bound(sorted_keys, i); cnt[i] = float(max(deg_i, 1)).
__global__ void csr_rowptr_kernel(const int32_t* __restrict__ sorted_keys, int64_t n_edges,
                                  int64_t n_nodes, int32_t* __restrict__ rowptr) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n_nodes) return;
    int64_t lo = 0, hi = n_edges;
    while (lo < hi) {
        int64_t mid = (lo + hi) >> 1;
        if (sorted_keys[mid] < (int32_t)i) lo = mid + 1; else hi = mid;
    }
    rowptr[i] = (int32_t)lo;
}

__global__ void csr_cnt_kernel(const int32_t* __restrict__ rowptr, int64_t n_nodes,
                               float* __restrict__ cnt) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    int d = rowptr[i + 1] - rowptr[i];
    cnt[i] = (float)(d < 1 ? 1 : d);
}

__global__ void csr_col_kernel(const int64_t* __restrict__ other_row, const int32_t* __restrict__ perm,
                               int64_t n_edges, int32_t* __restrict__ col) {
    int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n_edges) col[p] = (int32_t)other_row[perm[p]];
}

// w[e] = dis[src] * 1.0f * dis[dst], dis = deg^-1/2 with inf -> 0 (gcn_norm).
__global__ void gcn_weight_kernel(const int64_t* __restrict__ ei, int64_t n_edges,
                                  const int32_t* __restrict__ rowptr_dst, float* __restrict__ w) {
    int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_edges) return;
    int64_t s = ei[e], d = ei[n_edges + e];
    float ds = (float)(rowptr_dst[s + 1] - rowptr_dst[s]);
    float dd = (float)(rowptr_dst[d + 1] - rowptr_dst[d]);
    // ATen pow(x, -0.5) on CPU floats is 1 / sqrt(x) (two IEEE roundings); deg==0 -> inf -> 0
    float is = ds > 0.f ? __fdiv_rn(1.0f, __fsqrt_rn(ds)) : 0.f;
    float id = dd > 0.f ? __fdiv_rn(1.0f, __fsqrt_rn(dd)) : 0.f;
    w[e] = __fmul_rn(__fmul_rn(is, 1.0f), id);
}

// ---------------------------------------------------------------------------------------------
// CSR build, round 2: "upper digits by LSD passes, lowest digit inside per-bucket CTAs".
//
// The plain LSD sort above costs 3 launches per digit plus rowptr / cnt / col kernels: 9 launches and ~100 us per
// build of the arxiv-shaped graph (1.17 M edges), 18 launches per bench step for the two groupings -- launch- and
// latency-bound at 0.05 of the HBM roofline.  Here the node id is cut into an UPPER part (sorted
// by stable LSD passes, one for 18-bit ids) and the lowest LB <= 9 bits.  After the upper passes the edges of the
// 2^LB nodes of one bucket are one contiguous segment, in original order; one CTA per bucket finishes the job:
// it counts its nodes' edges (-> rowptr and the clamped count of those nodes), ranks the segment stably by the
// low bits, and writes perm and col (the other endpoint, gathered through perm) -- the last radix pass, rowptr,
// cnt and col fused into one launch.  Both groupings of one edge_index (by destination, by source) travel
// through the same launches as grid.y = 2 (stemgnn_csr_build_pair): 4 launches per step instead of 18.
// Same result as the LSD sort: a stable sort by (upper, low) = a stable sort by node id.
struct K0Side {
    const int64_t* key_row;     // edge_index row grouped by
    const int64_t* other_row;   // the other endpoint
    int32_t* kbuf[2];           // keys ping-pong (upper passes)
    int32_t* vbuf[2];           // edge ids ping-pong
    int32_t* block_hist;        // [BINS][nblk]
    int32_t* digit_total;       // [BINS]
    int32_t* seg_start;         // [n_buckets + 1] (only with more than one upper pass)
    int32_t* rowptr; int32_t* col; int32_t* perm; float* cnt;
};
struct K0Params {
    K0Side s[2];
    int64_t E, N;
    int nblk;                   // tiles of the upper passes
    int low_bits;               // LB
    int n_buckets;
};

template <int BINS, int ITEMS>
__global__ void __launch_bounds__(RS_THREADS)
k0_hist_kernel(const K0Params P, const int pass, const int shift) {
    __shared__ int hist[BINS];
    const K0Side& sd = P.s[blockIdx.y];
    const int64_t* k64 = pass == 0 ? sd.key_row : nullptr;
    const int32_t* k32 = pass == 0 ? nullptr : sd.kbuf[(pass - 1) & 1];
    for (int b = threadIdx.x; b < BINS; b += RS_THREADS) hist[b] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * (RS_THREADS * ITEMS);
    int keys[ITEMS];
#pragma unroll
    for (int r = 0; r < ITEMS; ++r) {
        const int64_t i = base + (int64_t)r * RS_THREADS + threadIdx.x;
        keys[r] = i < P.E ? load_key(k64, k32, i) : -1;
    }
#pragma unroll
    for (int r = 0; r < ITEMS; ++r)
        if (keys[r] >= 0) atomicAdd(&hist[(keys[r] >> shift) & (BINS - 1)], 1);
    __syncthreads();
    for (int b = threadIdx.x; b < BINS; b += RS_THREADS) sd.block_hist[(int64_t)b * P.nblk + blockIdx.x] = hist[b];
}

__global__ void __launch_bounds__(RS_THREADS)
k0_scan_kernel(const K0Params P) {
    const K0Side& sd = P.s[blockIdx.y];
    int32_t* row = sd.block_hist + (int64_t)blockIdx.x * P.nblk;
    int carry = 0;
    for (int b0 = 0; b0 < P.nblk; b0 += RS_THREADS) {
        const int b = b0 + threadIdx.x;
        const int v = b < P.nblk ? row[b] : 0;
        int tot;
        const int ex = block_excl_scan_256(v, &tot);
        if (b < P.nblk) row[b] = carry + ex;
        carry += tot;
    }
    if (threadIdx.x == 0) sd.digit_total[blockIdx.x] = carry;
}

// stable ranks of the items a warp holds (ITEMS rounds of 32 consecutive items): rank among the warp's earlier
// items with the same digit, warp totals left in wc[digit]
template <int ITEMS>
__device__ __forceinline__ void warp_rank(const int (&dig)[ITEMS], int* wc, unsigned short (&ranks)[ITEMS]) {
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < ITEMS; ++r) {
        const bool valid = dig[r] >= 0;
        const int d = valid ? dig[r] : 0x10000 + lane;          // invalid: unique, no peers
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        const int leader = __ffs(peers) - 1;
        int old = 0;
        if (valid && lane == leader) {
            old = wc[d];
            wc[d] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        ranks[r] = (unsigned short)(old + __popc(peers & lt_mask));
        __syncwarp();
    }
}

template <int BINS, int ITEMS>
__global__ void __launch_bounds__(RS_THREADS)
k0_scatter_kernel(const K0Params P, const int pass, const int shift) {
    constexpr int PER = BINS / RS_THREADS;
    __shared__ int wcount[RS_WARPS][BINS];
    __shared__ int blockbase[BINS];
    const K0Side& sd = P.s[blockIdx.y];
    const int64_t* k64 = pass == 0 ? sd.key_row : nullptr;
    const int32_t* k32 = pass == 0 ? nullptr : sd.kbuf[(pass - 1) & 1];
    const int32_t* vin = pass == 0 ? nullptr : sd.vbuf[(pass - 1) & 1];
    int32_t* kout = sd.kbuf[pass & 1];
    int32_t* vout = sd.vbuf[pass & 1];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < RS_WARPS * BINS; i += RS_THREADS) (&wcount[0][0])[i] = 0;
    {
        int d[PER], sum = 0;
#pragma unroll
        for (int j = 0; j < PER; ++j) { d[j] = sd.digit_total[threadIdx.x * PER + j]; sum += d[j]; }
        int tot;
        int ex = block_excl_scan_256(sum, &tot);
#pragma unroll
        for (int j = 0; j < PER; ++j) {
            const int b = threadIdx.x * PER + j;
            blockbase[b] = ex + sd.block_hist[(int64_t)b * P.nblk + blockIdx.x];
            ex += d[j];
        }
    }
    __syncthreads();
    const int64_t wbase = (int64_t)blockIdx.x * (RS_THREADS * ITEMS) + (int64_t)w * (32 * ITEMS);
    int keys[ITEMS], vals[ITEMS], dig[ITEMS];
    unsigned short ranks[ITEMS];
#pragma unroll
    for (int r = 0; r < ITEMS; ++r) {
        const int64_t i = wbase + r * 32 + lane;
        const bool valid = i < P.E;
        keys[r] = valid ? load_key(k64, k32, i) : 0;
        vals[r] = valid ? (vin ? vin[i] : (int)i) : 0;
        dig[r] = valid ? ((keys[r] >> shift) & (BINS - 1)) : -1;
    }
    warp_rank<ITEMS>(dig, wcount[w], ranks);
    __syncthreads();
    for (int b = threadIdx.x; b < BINS; b += RS_THREADS) {
        int run = 0;
#pragma unroll
        for (int i = 0; i < RS_WARPS; ++i) {
            const int t = wcount[i][b];
            wcount[i][b] = run;
            run += t;
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < ITEMS; ++r) {
        if (dig[r] >= 0) {
            const int64_t pos = (int64_t)blockbase[dig[r]] + wcount[w][dig[r]] + ranks[r];
            kout[pos] = keys[r];
            vout[pos] = vals[r];
        }
    }
}

// seg_start[b] = lower_bound(sorted upper keys, b << low_bits), b in [0, n_buckets]
__global__ void k0_bounds_kernel(const K0Params P, const int final_buf) {
    const K0Side& sd = P.s[blockIdx.y];
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b > P.n_buckets) return;
    const int32_t* keys = sd.kbuf[final_buf];
    const int64_t want = (int64_t)b << P.low_bits;
    int64_t lo = 0, hi = P.E;
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if ((int64_t)keys[mid] < want) lo = mid + 1; else hi = mid;
    }
    sd.seg_start[b] = (int32_t)lo;
}

// One CTA per bucket of 2^low_bits consecutive node ids.  src_mode 0: the segment comes from the upper passes
// (kbuf / vbuf[final_buf]); its bounds from seg_start, or from the single upper pass's digit totals when
// bounds_from_totals.  src_mode 1 (no upper pass: N <= 2^low_bits): the whole edge list, keys from key_row.
constexpr int K0_BUCKET_ITEMS = 16;                           // what the plan sizes buckets for (one chunk per segment)
constexpr int K0_BUCKET_CHUNK = RS_THREADS * K0_BUCKET_ITEMS;
template <int BI>                                             // items per thread and chunk: 16, or 8 (A/B: fewer registers)
__global__ void __launch_bounds__(RS_THREADS, BI == 16 ? 2 : 3)
k0_bucket_kernel(const K0Params P, const int final_buf, const int src_mode, const int bounds_from_totals,
                 const int total_bins) {
    constexpr int NB = 512;
    __shared__ int ncount[NB];          // edges per node of the bucket, then exclusive offsets + running chunk offsets
    __shared__ int wcount[RS_WARPS][NB];
    __shared__ int seg[2];
    const K0Side& sd = P.s[blockIdx.y];
    const int b = blockIdx.x;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int nb = 1 << P.low_bits, mask = nb - 1;
    const int32_t* kin = src_mode == 0 ? sd.kbuf[final_buf] : nullptr;
    const int32_t* vin = src_mode == 0 ? sd.vbuf[final_buf] : nullptr;
    const int64_t* k64 = src_mode == 0 ? nullptr : sd.key_row;
    for (int i = threadIdx.x; i < NB; i += RS_THREADS) ncount[i] = 0;
    if (src_mode == 1) {
        if (threadIdx.x == 0) { seg[0] = 0; seg[1] = (int)P.E; }
    } else if (bounds_from_totals) {
        // bucket id == digit of the single upper pass: segment start = exclusive scan of the digit totals
        int part = 0, mine = 0;
        for (int i = threadIdx.x; i < total_bins; i += RS_THREADS) {
            const int t = sd.digit_total[i];
            if (i < b) part += t;
            if (i == b) mine = t;
        }
        // block sum of `part`; the thread that saw bucket b publishes the length
        int tot;
        block_excl_scan_256(part, &tot);
        if (threadIdx.x == 0) seg[0] = tot;
        __syncthreads();
        if (b % RS_THREADS == threadIdx.x) seg[1] = seg[0] + mine;
    } else if (threadIdx.x == 0) {
        seg[0] = sd.seg_start[b]; seg[1] = sd.seg_start[b + 1];
    }
    __syncthreads();
    const int s0 = seg[0], s1 = seg[1];
    // ---- phase A: edges per node (a chunk's 16 keys per thread are all in flight before the first atomic)
    for (int c0 = s0; c0 < s1; c0 += (RS_THREADS * BI)) {
        int kk[BI];
#pragma unroll
        for (int r = 0; r < BI; ++r) {
            const int i = c0 + r * RS_THREADS + threadIdx.x;
            kk[r] = i < s1 ? (load_key(k64, kin, i) & mask) : -1;
        }
#pragma unroll
        for (int r = 0; r < BI; ++r)
            if (kk[r] >= 0) atomicAdd(&ncount[kk[r]], 1);
    }
    __syncthreads();
    {
        // exclusive scan of the 512 counts (2 per thread), rowptr / cnt of this bucket's nodes
        const int c0 = ncount[2 * threadIdx.x], c1 = ncount[2 * threadIdx.x + 1];
        int tot;
        const int ex = block_excl_scan_256(c0 + c1, &tot);
        const int64_t node0 = ((int64_t)b << P.low_bits) + 2 * threadIdx.x;
        if (2 * threadIdx.x < nb) {
            if (node0 < P.N) { sd.rowptr[node0] = s0 + ex; if (sd.cnt) sd.cnt[node0] = (float)(c0 < 1 ? 1 : c0); }
            if (node0 + 1 < P.N && 2 * threadIdx.x + 1 < nb) {
                sd.rowptr[node0 + 1] = s0 + ex + c0;
                if (sd.cnt) sd.cnt[node0 + 1] = (float)(c1 < 1 ? 1 : c1);
            }
        }
        __syncthreads();
        ncount[2 * threadIdx.x] = ex;
        ncount[2 * threadIdx.x + 1] = ex + c0;
        if (b == (int)gridDim.x - 1 && threadIdx.x == 0) sd.rowptr[P.N] = (int32_t)P.E;
    }
    __syncthreads();
    // ---- phase B: stable placement, chunk after chunk in segment order
    for (int c0 = s0; c0 < s1; c0 += (RS_THREADS * BI)) {
        for (int i = threadIdx.x; i < RS_WARPS * NB; i += RS_THREADS) (&wcount[0][0])[i] = 0;
        __syncthreads();
        const int wbase = c0 + w * (32 * BI);
        int vals[BI], dig[BI];
        unsigned short ranks[BI];
#pragma unroll
        for (int r = 0; r < BI; ++r) {
            const int i = wbase + r * 32 + lane;
            const bool valid = i < s1;
            dig[r] = valid ? (load_key(k64, kin, i) & mask) : -1;
            vals[r] = valid ? (vin ? vin[i] : i) : 0;
        }
        int other[BI];                 // the other endpoint: independent gathers, issued before the ranking
#pragma unroll
        for (int r = 0; r < BI; ++r) other[r] = dig[r] >= 0 ? (int)sd.other_row[vals[r]] : 0;
        warp_rank<BI>(dig, wcount[w], ranks);
        __syncthreads();
        for (int d = threadIdx.x; d < nb; d += RS_THREADS) {
            int run = 0;
#pragma unroll
            for (int i = 0; i < RS_WARPS; ++i) {
                const int t = wcount[i][d];
                wcount[i][d] = run;
                run += t;
            }
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < BI; ++r) {
            if (dig[r] >= 0) {
                const int pos = s0 + ncount[dig[r]] + wcount[w][dig[r]] + ranks[r];
                sd.perm[pos] = vals[r];
                sd.col[pos] = other[r];
            }
        }
        if (c0 + (RS_THREADS * BI) < s1) {            // more chunks: advance the node offsets by this chunk's counts
            __syncthreads();                        // (every valid item adds one; the order no longer matters)
#pragma unroll
            for (int r = 0; r < BI; ++r)
                if (dig[r] >= 0) atomicAdd(&ncount[dig[r]], 1);
            __syncthreads();
        }
    }
}

static int key_bits(int64_t key_range) {
    int bits = 1;
    while (bits < 32 && ((int64_t)1 << bits) < key_range) ++bits;
    return bits;
}
// 9-bit digits when they need fewer passes than 8-bit ones
static int radix_digit_bits(int64_t key_range) {
    const int bits = key_bits(key_range);
    return (bits + 8) / 9 < (bits + 7) / 8 ? 9 : 8;
}
static int radix_passes(int64_t key_range) {
    const int bits = key_bits(key_range), db = radix_digit_bits(key_range);
    return (bits + db - 1) / db;
}

size_t radix_sort_workspace(int64_t n) {
    int64_t nblk = (n + RS_TILE - 1) / RS_TILE;
    size_t b = 0;
    b += 2 * align_up((size_t)n * 4, 256);       // keys ping-pong
    b += align_up((size_t)n * 4, 256);           // values (the other buffer is the caller's output)
    b += align_up((size_t)RS_MAX_BINS * (size_t)(nblk > 0 ? nblk : 1) * 4, 256);
    b += align_up(RS_MAX_BINS * 4, 256);
    return b;
}

// Stable sort of (key, index) pairs, keys in [0, key_range).  Keys come from k64 (int64, low 32
// bits used) or k32.  The sorted indices land in vals_out; returns the sorted keys buffer
// (inside the workspace).  n > 0.
const int32_t* radix_sort_pairs(const int64_t* k64, const int32_t* k32, int64_t n, int64_t key_range,
                                int32_t* vals_out, Workspace& ws, cudaStream_t stream) {
    int32_t* kbuf[2] = {ws.take<int32_t>(n), ws.take<int32_t>(n)};
    int32_t* vws = ws.take<int32_t>(n);
    const int nblk = (int)((n + RS_TILE - 1) / RS_TILE);
    int32_t* block_hist = ws.take<int32_t>((size_t)RS_MAX_BINS * nblk);
    int32_t* digit_total = ws.take<int32_t>(RS_MAX_BINS);
    const int P = radix_passes(key_range), DB = radix_digit_bits(key_range);
    const int32_t* kin = k32;
    const int32_t* vin = nullptr;
    for (int p = 0; p < P; ++p) {
        int32_t* kout = kbuf[p & 1];
        int32_t* vout = ((P - 1 - p) & 1) ? vws : vals_out;   // last pass writes vals_out
        const int64_t* kk64 = p == 0 ? k64 : nullptr;
        if (DB == 9) {
            SG_LAUNCH(radix_hist_kernel<512><<<nblk, RS_THREADS, 0, stream>>>(kk64, kin, n, 9 * p, nblk, block_hist));
            SG_LAUNCH(radix_scan_kernel<<<512, RS_THREADS, 0, stream>>>(block_hist, nblk, digit_total));
            SG_LAUNCH(radix_scatter_kernel<512><<<nblk, RS_THREADS, 0, stream>>>(kk64, kin, vin, n, 9 * p, nblk,
                                                                       block_hist, digit_total, kout, vout));
        } else {
            SG_LAUNCH(radix_hist_kernel<256><<<nblk, RS_THREADS, 0, stream>>>(kk64, kin, n, 8 * p, nblk, block_hist));
            SG_LAUNCH(radix_scan_kernel<<<256, RS_THREADS, 0, stream>>>(block_hist, nblk, digit_total));
            SG_LAUNCH(radix_scatter_kernel<256><<<nblk, RS_THREADS, 0, stream>>>(kk64, kin, vin, n, 8 * p, nblk,
                                                                       block_hist, digit_total, kout, vout));
        }
        kin = kout;
        vin = vout;
    }
    return kin;
}

void launch_lower_bound(const int32_t* sorted_keys, int64_t n, int64_t n_bins, int32_t* offsets,
                        cudaStream_t stream) {
    SG_LAUNCH(csr_rowptr_kernel<<<(unsigned)((n_bins + 1 + 255) / 256), 256, 0, stream>>>(sorted_keys, n,
                                                                                 n_bins, offsets));
}

}  // namespace stemgnn

using namespace stemgnn;

namespace {
struct K0Plan {
    int low_bits, passes, digit_bits, bins, items, nblk, n_buckets;
    size_t side_bytes;
};
K0Plan k0_plan(int64_t E, int64_t N) {
    K0Plan pl{};
    const int bits = key_bits(N > 1 ? N : 2);
    const int64_t avg = E / (N > 0 ? N : 1) + 1;
    int best_cost = 1 << 30;
    for (int lb = 9; lb >= 4; --lb) {               // ties go to the wider bucket (iterating downwards, strict <)
        const int ub = bits > lb ? bits - lb : 0;
        const int passes = (ub + 8) / 9;
        const int64_t chunks = (avg * ((int64_t)1 << lb) + K0_BUCKET_CHUNK - 1) / K0_BUCKET_CHUNK;
        const int cost = passes * 64 + (int)(chunks < 63 ? chunks : 63);
        if (cost < best_cost) { best_cost = cost; pl.low_bits = lb; pl.passes = passes; }
    }
    const int ub = bits > pl.low_bits ? bits - pl.low_bits : 0;
    pl.digit_bits = pl.passes ? (ub + pl.passes - 1) / pl.passes : 0;
    pl.bins = pl.digit_bits == 9 ? 512 : 256;
    const int64_t nblk16 = (E + RS_THREADS * 16 - 1) / (RS_THREADS * 16);
    pl.items = nblk16 >= 4 * (int64_t)sm_count() ? 16 : 8;     // small inputs: more, smaller tiles (the passes are latency-bound)
    pl.nblk = (int)((E + RS_THREADS * pl.items - 1) / (RS_THREADS * pl.items));
    pl.n_buckets = (int)(((N > 0 ? N : 1) + ((int64_t)1 << pl.low_bits) - 1) >> pl.low_bits);
    const int nbuf = pl.passes >= 2 ? 2 : pl.passes;
    pl.side_bytes = 2 * (size_t)nbuf * align_up((size_t)(E > 0 ? E : 1) * 4, 256)
                  + align_up((size_t)512 * (size_t)(pl.nblk > 0 ? pl.nblk : 1) * 4, 256) + align_up(512 * 4, 256)
                  + align_up((size_t)(pl.n_buckets + 1) * 4, 256);
    return pl;
}

template <int BINS, int ITEMS>
void k0_upper_pass(const K0Params& P, int sides, int pass, int shift, cudaStream_t stream) {
    const dim3 tiles((unsigned)P.nblk, (unsigned)sides);
    SG_LAUNCH(k0_hist_kernel<BINS, ITEMS><<<tiles, RS_THREADS, 0, stream>>>(P, pass, shift));
    SG_LAUNCH(k0_scan_kernel<<<dim3(BINS, (unsigned)sides), RS_THREADS, 0, stream>>>(P));
    SG_LAUNCH(k0_scatter_kernel<BINS, ITEMS><<<tiles, RS_THREADS, 0, stream>>>(P, pass, shift));
}

// `sides` groupings of one edge list through the same launches (grid.y = side)
int k0_build(const int64_t* edge_index, int64_t E, int64_t N, int sides, const int* group_rows, int32_t* const* rowptr,
             int32_t* const* col, int32_t* const* perm, float* const* cnt, void* workspace, size_t workspace_bytes,
             cudaStream_t stream) {
    const int T = 256;
    if (E == 0) {
        for (int i = 0; i < sides; ++i) {
            SG_CHECK_CUDA(cudaMemsetAsync(rowptr[i], 0, (size_t)(N + 1) * 4, stream));
            if (cnt[i] && N > 0) SG_LAUNCH(csr_cnt_kernel<<<(unsigned)((N + T - 1) / T), T, 0, stream>>>(rowptr[i], N, cnt[i]));
        }
        SG_CHECK_LAUNCH();
        return STEMGNN_OK;
    }
    const K0Plan pl = k0_plan(E, N);
    if (workspace == nullptr || workspace_bytes < pl.side_bytes * sides) {
        set_error("stemgnn_csr_build: workspace too small (%zu < %zu)", workspace_bytes, pl.side_bytes * sides);
        return STEMGNN_ENOMEM_WS;
    }
    Workspace ws(workspace, workspace_bytes);
    K0Params P{};
    P.E = E; P.N = N; P.nblk = pl.nblk; P.low_bits = pl.low_bits; P.n_buckets = pl.n_buckets;
    const int nbuf = pl.passes >= 2 ? 2 : pl.passes;
    for (int i = 0; i < sides; ++i) {
        K0Side& sd = P.s[i];
        sd.key_row = edge_index + (int64_t)group_rows[i] * E;
        sd.other_row = edge_index + (int64_t)(1 - group_rows[i]) * E;
        for (int j = 0; j < nbuf; ++j) { sd.kbuf[j] = ws.take<int32_t>(E); sd.vbuf[j] = ws.take<int32_t>(E); }
        sd.block_hist = ws.take<int32_t>((size_t)512 * pl.nblk);
        sd.digit_total = ws.take<int32_t>(512);
        sd.seg_start = ws.take<int32_t>(pl.n_buckets + 1);
        sd.rowptr = rowptr[i]; sd.col = col[i]; sd.perm = perm[i]; sd.cnt = cnt[i];
    }
    for (int p = 0; p < pl.passes; ++p) {
        const int shift = pl.low_bits + p * pl.digit_bits;
        if (pl.bins == 512) {
            if (pl.items == 16) k0_upper_pass<512, 16>(P, sides, p, shift, stream);
            else k0_upper_pass<512, 8>(P, sides, p, shift, stream);
        } else {
            if (pl.items == 16) k0_upper_pass<256, 16>(P, sides, p, shift, stream);
            else k0_upper_pass<256, 8>(P, sides, p, shift, stream);
        }
    }
    const int final_buf = pl.passes ? (pl.passes - 1) & 1 : 0;
    if (pl.passes >= 2)
        SG_LAUNCH(k0_bounds_kernel<<<dim3((unsigned)((pl.n_buckets + 1 + T - 1) / T), (unsigned)sides), T, 0, stream>>>(P, final_buf));
    // 8 items per thread (two chunks for a 3,500-edge segment, 80 registers, 3 CTAs per SM) against 16 (one chunk, 128
    // registers, 2 CTAs): 0.124 vs 0.131 ms per arxiv-shaped pair build on a B200; STEMGNN_K0_BUCKET_ITEMS=16 for the A/B
    static const int bucket_items = [] { const char* e = getenv("STEMGNN_K0_BUCKET_ITEMS"); return e ? atoi(e) : 8; }();
    const dim3 bgrid((unsigned)pl.n_buckets, (unsigned)sides);
    if (bucket_items != 16)
        SG_LAUNCH(k0_bucket_kernel<8><<<bgrid, RS_THREADS, 0, stream>>>(P, final_buf, pl.passes == 0 ? 1 : 0, pl.passes == 1 ? 1 : 0, pl.bins));
    else
        SG_LAUNCH(k0_bucket_kernel<16><<<bgrid, RS_THREADS, 0, stream>>>(P, final_buf, pl.passes == 0 ? 1 : 0, pl.passes == 1 ? 1 : 0, pl.bins));
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}

int k0_check_sizes(int64_t num_edges, int64_t num_nodes) {
    SG_CHECK_ARG(num_edges >= 0 && num_nodes >= 0, "negative size");
    SG_CHECK_ARG(num_edges < (int64_t)2147483647 && num_nodes < (int64_t)2147483647,
                 "int32 CSR: E and N must be < 2^31");
    return STEMGNN_OK;
}
}  // namespace

extern "C" int stemgnn_csr_build_workspace(int64_t num_edges, int64_t num_nodes, size_t* bytes) {
    SG_CHECK_ARG(bytes != nullptr, "bytes_host is NULL");
    if (int rc = k0_check_sizes(num_edges, num_nodes)) return rc;
    *bytes = k0_plan(num_edges, num_nodes).side_bytes;
    return STEMGNN_OK;
}

extern "C" int stemgnn_csr_build_pair_workspace(int64_t num_edges, int64_t num_nodes, size_t* bytes) {
    SG_CHECK_ARG(bytes != nullptr, "bytes_host is NULL");
    if (int rc = k0_check_sizes(num_edges, num_nodes)) return rc;
    *bytes = 2 * k0_plan(num_edges, num_nodes).side_bytes;
    return STEMGNN_OK;
}

extern "C" int stemgnn_csr_build(const int64_t* edge_index, int64_t E, int64_t N, int group_row,
                                 int32_t* rowptr, int32_t* col, int32_t* perm, float* cnt,
                                 void* workspace, size_t workspace_bytes, stemgnn_stream_t stream_) {
    SG_CHECK_ARG(group_row == 0 || group_row == 1, "group_row must be 0 or 1");
    SG_CHECK_ARG(rowptr && (E == 0 || (edge_index && col && perm)), "NULL pointer");
    if (int rc = k0_check_sizes(E, N)) return rc;
    return k0_build(edge_index, E, N, 1, &group_row, &rowptr, &col, &perm, &cnt, workspace, workspace_bytes,
                    (cudaStream_t)stream_);
}

extern "C" int stemgnn_csr_build_pair(const int64_t* edge_index, int64_t E, int64_t N, int32_t* const rowptr[2],
                                      int32_t* const col[2], int32_t* const perm[2], float* const cnt[2],
                                      void* workspace, size_t workspace_bytes, stemgnn_stream_t stream_) {
    SG_CHECK_ARG(rowptr && col && perm && cnt && rowptr[0] && rowptr[1], "NULL pointer");
    SG_CHECK_ARG(E == 0 || (edge_index && col[0] && col[1] && perm[0] && perm[1]), "NULL pointer");
    if (int rc = k0_check_sizes(E, N)) return rc;
    const int rows[2] = {0, 1};
    return k0_build(edge_index, E, N, 2, rows, rowptr, col, perm, cnt, workspace, workspace_bytes, (cudaStream_t)stream_);
}

extern "C" int stemgnn_gcn_edge_weight(const int64_t* edge_index_sl, int64_t E, int64_t N,
                                       const int32_t* rowptr_by_dst, float* w_out,
                                       stemgnn_stream_t stream) {
    SG_CHECK_ARG(E >= 0 && N >= 0, "negative size");
    if (E == 0) return STEMGNN_OK;
    SG_CHECK_ARG(edge_index_sl && rowptr_by_dst && w_out, "NULL pointer");
    SG_LAUNCH(gcn_weight_kernel<<<(unsigned)((E + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        edge_index_sl, E, rowptr_by_dst, w_out));
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}
