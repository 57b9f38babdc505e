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

extern "C" int stemgnn_csr_build_workspace(int64_t num_edges, int64_t num_nodes, size_t* bytes) {
    SG_CHECK_ARG(bytes != nullptr, "bytes_host is NULL");
    SG_CHECK_ARG(num_edges >= 0 && num_nodes >= 0, "negative size");
    SG_CHECK_ARG(num_edges < (int64_t)2147483647 && num_nodes < (int64_t)2147483647,
                 "int32 CSR: E and N must be < 2^31");
    size_t b = radix_sort_workspace(num_edges);
    *bytes = b;
    return STEMGNN_OK;
}

extern "C" int stemgnn_csr_build(const int64_t* edge_index, int64_t E, int64_t N, int group_row,
                                 int32_t* rowptr, int32_t* col, int32_t* perm, float* cnt,
                                 void* workspace, size_t workspace_bytes, stemgnn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SG_CHECK_ARG(group_row == 0 || group_row == 1, "group_row must be 0 or 1");
    SG_CHECK_ARG(rowptr && (E == 0 || (edge_index && col && perm)), "NULL pointer");
    size_t need = 0;
    int rc = stemgnn_csr_build_workspace(E, N, &need);
    if (rc) return rc;
    if (E > 0 && (workspace == nullptr || workspace_bytes < need)) {
        set_error("stemgnn_csr_build: workspace too small (%zu < %zu)", workspace_bytes, need);
        return STEMGNN_ENOMEM_WS;
    }
    const int T = 256;
    if (E == 0) {
        cudaMemsetAsync(rowptr, 0, (size_t)(N + 1) * 4, stream);
        if (cnt && N > 0) SG_LAUNCH(csr_cnt_kernel<<<(unsigned)((N + T - 1) / T), T, 0, stream>>>(rowptr, N, cnt));
        SG_CHECK_LAUNCH();
        return STEMGNN_OK;
    }
    Workspace ws(workspace, workspace_bytes);
    const int64_t* key_row = edge_index + (int64_t)group_row * E;
    const int64_t* other_row = edge_index + (int64_t)(1 - group_row) * E;
    const int32_t* kin = radix_sort_pairs(key_row, nullptr, E, N, perm, ws, stream);
    SG_LAUNCH(csr_rowptr_kernel<<<(unsigned)((N + 1 + T - 1) / T), T, 0, stream>>>(kin, E, N, rowptr));
    if (cnt && N > 0) SG_LAUNCH(csr_cnt_kernel<<<(unsigned)((N + T - 1) / T), T, 0, stream>>>(rowptr, N, cnt));
    SG_LAUNCH(csr_col_kernel<<<(unsigned)((E + T - 1) / T), T, 0, stream>>>(other_row, perm, E, col));
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
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
