// k1_aggregate.cu -- K1 (forward) / K2 (backward): pull-based CSR aggregation.
//
// Replaces index_select -> message -> scatter_add_ -> count/clamp/div and its autograd
// (reference STEM-GNN/model/encoder.py:82,94-97 MySAGEConv.propagate/message, :124 scatter_mean,
// :197/:199 GCN/GIN sums; PyG MeanAggregation / torch_scatter.scatter_sum underneath).
//
// One group of LANES threads owns one CSR row; lanes span the feature dimension with VEC-wide
// (128-bit when F%4==0) loads, CH chunks per lane.  The neighbours of a row are walked in
// stable (original edge) order and added ONE BY ONE with separately rounded adds, so the fp32
// result is bit-identical to the CPU scatter_add_ order; U edges are fetched ahead of the adds
// so U*CH independent 16 B loads per thread are in flight (HBM/L2-latency bound, guideline 7).
// No atomics, no E x F intermediates: per edge the kernel moves F*4 B (gathered row) + 4 B
// (col) [+ F*4 B dense edge feature + 4 B perm], per row F*4 B out + 4 B rowptr.
// Rows of >= 256 edges ("hubs") are handed to a second launch that spreads a row over many CTAs by
// column slices (aggregate_hub_kernel / aggregate_hub_ring_kernel): same order of additions per column.
#include "common.cuh"
#include <cstdlib>

namespace stemgnn {

enum AggMode { MODE_PLAIN = 0, MODE_XE = 1, MODE_W = 2 };

template <int VEC> struct Vec;
template <> struct Vec<4> {
    float v[4];
    __device__ __forceinline__ void load(const float* p) {
        float4 t = __ldg(reinterpret_cast<const float4*>(p));
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
    __device__ __forceinline__ void load_plain(const float* p) {      // coherent load (data this stream wrote before)
        float4 t = *reinterpret_cast<const float4*>(p);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
    __device__ __forceinline__ void store(float* p) const {
        *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    }
};
template <> struct Vec<1> {
    float v[1];
    __device__ __forceinline__ void load(const float* p) { v[0] = __ldg(p); }
    __device__ __forceinline__ void load_plain(const float* p) { v[0] = *p; }
    __device__ __forceinline__ void store(float* p) const { *p = v[0]; }
};

struct AggParams {
    const int32_t* rowptr; const int32_t* col; const int32_t* perm;
    int64_t n_rows, num_edges; int F;
    const float* gathered; const float* row_cnt; const float* gather_cnt;
    const float* edge_feat; const float* type_table; const int32_t* edge_type;
    const float* edge_weight; const float* row_x;
    int relu; float* out;
    int64_t out_ld, row_x_ld;   // row strides (floats) of out / row_x: column slices of wider matrices
    int* hub;                   // [0] big hubs, [1] small hubs, [2] next work item, [3] CTAs done, [4..] row ids; or NULL
    int hub_cap;                // slots of the row list: big hubs from the front, small ones from the back
    int accumulate;             // row sums start from out[r] instead of 0 (ABI 5)
    int hints;                  // L2 cache-policy hints on the streaming accesses (A/B switch)
    const float* row_add;       // added to the FINISHED row (after count division / relu mask), or NULL (ABI 6)
    int64_t row_add_ld;
};

// Rows with at least this many edges ("hubs": ogbn-arxiv's largest in-degree is 13,155, ogbn-products'
// 17,481) leave the row-per-group kernel: one group walks its row with U*CH loads in flight, so a
// 17,000-edge row alone takes milliseconds after every other row has finished.  They are appended to
// a work list and summed by aggregate_hub_kernel below (same order of additions, same bits).
constexpr int HUB_ROW_EDGES = 256;
constexpr int HUB_BIG_EDGES = 4096;    // from here on a hub is cut into 8-column instead of 32-column slices
constexpr int HUB_TILE_FLOATS = 8192;  // 32 KB staging buffer: 256 edges x 32 columns or 1024 edges x 8
constexpr int HUB_LOADERS = 256;
constexpr int HUB_THREADS = HUB_LOADERS + 32;

// One group of U edges of one row.  FULL (all U edges exist, every chunk of the lane inside F) drops
// the predicates; a separate predicate-free loop for full groups was measured and LOST 5-15 %
// (more registers and code for a latency-bound kernel), so the kernel instantiates FULL=false only.
template <int VEC, int CH, int U, int MODE, bool BWD, bool RELU, bool FULL>
__device__ __forceinline__ void edge_group(const AggParams& a, const int F, const int p, const int end,
                                           const int (&foff)[CH], const bool (&fok)[CH],
                                           const Vec<VEC> (&rx)[CH], Vec<VEC> (&acc)[CH]) {
    int cidx[U];
    int eid[U];
    float scal[U];  // bwd: gather_cnt divisor; MODE_W: edge weight
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const bool ok = FULL || p + u < end;
        cidx[u] = ok ? __ldg(a.col + p + u) : -1;
        eid[u] = (ok && MODE != MODE_PLAIN) ? __ldg(a.perm + p + u) : 0;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
        scal[u] = 1.f;
        if (FULL || cidx[u] >= 0) {
            if (MODE == MODE_W) scal[u] = __ldg(a.edge_weight + eid[u]);
            else if (BWD && a.gather_cnt) scal[u] = __ldg(a.gather_cnt + cidx[u]);
        }
    }
    Vec<VEC> g[U][CH];
    Vec<VEC> xe[U][CH];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        const float* gp = a.gathered + (int64_t)cidx[u] * F;
        const float* xp = nullptr;
        if (MODE == MODE_XE && (FULL || cidx[u] >= 0))
            xp = a.edge_feat ? a.edge_feat + (int64_t)eid[u] * F
                             : a.type_table + (int64_t)__ldg(a.edge_type + eid[u]) * F;
#pragma unroll
        for (int c = 0; c < CH; ++c) {
            if (FULL || (cidx[u] >= 0 && fok[c])) {
                g[u][c].load(gp + foff[c]);
                if (MODE == MODE_XE) xe[u][c].load(xp + foff[c]);
            }
        }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
        if (!FULL && cidx[u] < 0) continue;
#pragma unroll
        for (int c = 0; c < CH; ++c) {
            if (!FULL && !fok[c]) continue;
#pragma unroll
            for (int k = 0; k < VEC; ++k) {
                float m = g[u][c].v[k];
                if (MODE == MODE_W) {
                    m = __fmul_rn(scal[u], m);
                } else if (!BWD) {
                    if (MODE == MODE_XE) m = __fadd_rn(m, xe[u][c].v[k]);
                    if (RELU) m = m > 0.f ? m : 0.f;
                } else {
                    if (a.gather_cnt) m = __fdiv_rn(m, scal[u]);
                    if (MODE == MODE_XE && RELU) {
                        float pre = __fadd_rn(rx[c].v[k], xe[u][c].v[k]);
                        m = pre > 0.f ? m : 0.f;
                    }
                }
                acc[c].v[k] = __fadd_rn(acc[c].v[k], m);
            }
        }
    }
}

// Resident 256-thread blocks per SM the register allocation must allow.  The kernel is bound by
// memory latency x warps in flight: measured on B200 (scripts/agg_bench.py), 2 blocks lose 25-50 %
// against 3 everywhere, 4 blocks (64 registers) gain 10 % for one 16 B chunk per lane across a full
// warp (F = 100..128) and lose 5-10 % elsewhere (spills, L2 thrash), 5 blocks always lose.
template <int LANES, int CH>
struct AggOcc { static constexpr int min_blocks = (LANES == 32 && CH == 1) ? 4 : (CH <= 6 ? 3 : 2); };

template <int VEC, int LANES, int CH, int MODE, bool BWD, bool RELU>
__global__ void __launch_bounds__(256, AggOcc<LANES, CH>::min_blocks)
aggregate_kernel(const AggParams a) {
    constexpr int U = CH <= 1 ? 8 : (CH <= 2 ? 4 : (CH <= 4 ? 2 : 1));
    constexpr int ROWS_PER_WARP = 32 / LANES;
    constexpr int TILE_F = LANES * CH * VEC;
    const int lane = threadIdx.x & (LANES - 1);
    const int64_t row = ((int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * ROWS_PER_WARP
                        + ((threadIdx.x & 31) / LANES);
    if (row >= a.n_rows) return;
    const int F = a.F;
    const int beg = __ldg(a.rowptr + row), end = __ldg(a.rowptr + row + 1);
    if (a.hub != nullptr && end - beg >= HUB_ROW_EDGES) {
        if (lane == 0) {
            if (end - beg >= HUB_BIG_EDGES) a.hub[4 + atomicAdd(a.hub, 1)] = (int)row;
            else a.hub[4 + a.hub_cap - 1 - atomicAdd(a.hub + 1, 1)] = (int)row;
        }
        return;
    }

    for (int f0 = 0; f0 < F; f0 += TILE_F) {
        int foff[CH];
        bool fok[CH];
#pragma unroll
        for (int c = 0; c < CH; ++c) {
            foff[c] = f0 + (c * LANES + lane) * VEC;
            fok[c] = foff[c] < F;
        }
        Vec<VEC> acc[CH];
        Vec<VEC> rx[CH];   // backward relu mask source x[row] (MODE_XE per edge, MODE_PLAIN at the end)
#pragma unroll
        for (int c = 0; c < CH; ++c) {
#pragma unroll
            for (int k = 0; k < VEC; ++k) { acc[c].v[k] = 0.f; rx[c].v[k] = 0.f; }
            if (a.accumulate && fok[c]) acc[c].load_plain(a.out + row * a.out_ld + foff[c]);
            if (BWD && MODE != MODE_W && RELU && fok[c]) rx[c].load(a.row_x + row * a.row_x_ld + foff[c]);
        }

        for (int p = beg; p < end; p += U)
            edge_group<VEC, CH, U, MODE, BWD, RELU, false>(a, F, p, end, foff, fok, rx, acc);

        float rc = 1.f;
        if (!BWD && a.row_cnt) rc = __ldg(a.row_cnt + row);
#pragma unroll
        for (int c = 0; c < CH; ++c) {
            if (!fok[c]) continue;
#pragma unroll
            for (int k = 0; k < VEC; ++k) {
                float r = acc[c].v[k];
                if (!BWD && a.row_cnt) r = __fdiv_rn(r, rc);
                if (BWD && MODE == MODE_PLAIN && RELU) r = rx[c].v[k] > 0.f ? r : 0.f;
                acc[c].v[k] = r;
            }
            if (a.row_add) {
                Vec<VEC> ad;
                ad.load_plain(a.row_add + row * a.row_add_ld + foff[c]);
#pragma unroll
                for (int k = 0; k < VEC; ++k) acc[c].v[k] = __fadd_rn(acc[c].v[k], ad.v[k]);
            }
            acc[c].store(a.out + row * a.out_ld + foff[c]);
        }
    }
}


// ---------------------------------------------------------------------------------------------
// Plain mode (message = the gathered row, relu at most; no per-edge data), 64 < F <= 1024: the
// row-walking kernel with STAGED NEIGHBOUR INDICES.
//
// aggregate_kernel above gives every row to one warp that then leaves: rowptr -> col -> gathered
// rows are three dependent memory round trips per row, and with the 7 edges per row of the
// arxiv-shaped graph those round trips, not bytes, set the launch time (ncu round 1: issue active
// 42 %, long scoreboard, 91 instructions per edge-warp).  Here a warp owns STREAM_ROWS consecutive
// rows, i.e. one contiguous run of the CSR: its rowptr values come in one coalesced load (one per
// lane) and the col entries of the whole run travel through a two-register window per lane
// (2 x 32 indices, 128-byte aligned coalesced loads, the next 32 fetched while the current ones
// are used) and reach the lanes by shuffle.  Per row only the gather round trip is left.
// The streaming accesses (rowptr, col, out) carry an L2 evict-first policy and skip L1, so the
// 126 MB L2 keeps the gathered matrix instead of the output stream (round 1 measured 279 MB of DRAM
// traffic for 179 MB of compulsory bytes at F = 128 although X fits L2).
// Additions stay in edge order, one by one, separately rounded: same bits as aggregate_kernel.
constexpr int STREAM_ROWS = 8;

__device__ __forceinline__ uint64_t policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ int ld_stream_i32(const int32_t* p, uint64_t pol, bool hints) {
    int v;
    if (hints) asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.b32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol));
    else v = __ldg(p);
    return v;
}
__device__ __forceinline__ void st_stream_f4(float* p, const float (&v)[4], uint64_t pol, bool hints) {
    if (hints)
        asm volatile("st.global.L2::cache_hint.v4.f32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(p), "f"(v[0]), "f"(v[1]),
                     "f"(v[2]), "f"(v[3]), "l"(pol)
                     : "memory");
    else *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
}

template <int CH, bool RELU_MSG, bool MASK_END, int OCC>
__global__ void __launch_bounds__(256, OCC)
aggregate_stream_kernel(const AggParams a) {
    constexpr int U = CH <= 1 ? 8 : (CH <= 2 ? 4 : (CH <= 4 ? 2 : 1));
    const int lane = threadIdx.x & 31;
    const int64_t r0 = ((int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * STREAM_ROWS;
    if (r0 >= a.n_rows) return;
    const int nr = (int)(a.n_rows - r0 < STREAM_ROWS ? a.n_rows - r0 : STREAM_ROWS);
    const int F = a.F;
    const bool hints = a.hints != 0;
    const uint64_t pol = policy_evict_first();
    const int E = (int)a.num_edges;

    int rp = 0;
    if (lane <= nr) rp = ld_stream_i32(a.rowptr + r0 + lane, pol, hints);
    float cnt_l = 1.f;                          // this lane's row's clamped count: one coalesced load per warp
    if (a.row_cnt && lane < nr) cnt_l = __ldg(a.row_cnt + r0 + lane);
    int foff[CH];
    bool fok[CH];
#pragma unroll
    for (int c = 0; c < CH; ++c) {
        foff[c] = (c * 32 + lane) * 4;
        fok[c] = foff[c] < F;
    }
    // index window: w0 holds col[wb + lane], w1 holds col[wb + 32 + lane]
    int wb = __shfl_sync(0xffffffffu, rp, 0) & ~31;
    int w0 = wb + lane < E ? ld_stream_i32(a.col + wb + lane, pol, hints) : 0;
    int w1 = wb + 32 + lane < E ? ld_stream_i32(a.col + wb + 32 + lane, pol, hints) : 0;

    for (int r = 0; r < nr; ++r) {
        const int beg = __shfl_sync(0xffffffffu, rp, r), end = __shfl_sync(0xffffffffu, rp, r + 1);
        const int64_t row = r0 + r;
        if (a.hub != nullptr && end - beg >= HUB_ROW_EDGES) {
            if (lane == 0) {
                if (end - beg >= HUB_BIG_EDGES) a.hub[4 + atomicAdd(a.hub, 1)] = (int)row;
                else a.hub[4 + a.hub_cap - 1 - atomicAdd(a.hub + 1, 1)] = (int)row;
            }
            continue;
        }
        Vec<4> acc[CH];
        Vec<4> rx[CH];
#pragma unroll
        for (int c = 0; c < CH; ++c) {
#pragma unroll
            for (int k = 0; k < 4; ++k) { acc[c].v[k] = 0.f; rx[c].v[k] = 0.f; }
            if (a.accumulate && fok[c]) acc[c].load_plain(a.out + row * a.out_ld + foff[c]);
            if (MASK_END && fok[c]) rx[c].load(a.row_x + row * a.row_x_ld + foff[c]);
        }
        for (int p = beg; p < end; p += U) {
            if (p >= wb + 32) {                       // [p, p + U) must lie inside the window; move it early
                if (p >= wb + 64) {
                    wb = p & ~31;
                    w0 = wb + lane < E ? ld_stream_i32(a.col + wb + lane, pol, hints) : 0;
                } else {
                    w0 = w1;
                    wb += 32;
                }
                w1 = wb + 32 + lane < E ? ld_stream_i32(a.col + wb + 32 + lane, pol, hints) : 0;
            }
            int cidx[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int i = p + u - wb;             // warp-uniform
                const int v = __shfl_sync(0xffffffffu, i < 32 ? w0 : w1, i & 31);
                cidx[u] = p + u < end ? v : -1;
            }
            Vec<4> g[U][CH];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const float* gp = a.gathered + (int64_t)cidx[u] * F;
#pragma unroll
                for (int c = 0; c < CH; ++c)
                    if (cidx[u] >= 0 && fok[c]) g[u][c].load(gp + foff[c]);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (cidx[u] < 0) continue;
#pragma unroll
                for (int c = 0; c < CH; ++c) {
                    if (!fok[c]) continue;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        float m = g[u][c].v[k];
                        if (RELU_MSG) m = m > 0.f ? m : 0.f;
                        acc[c].v[k] = __fadd_rn(acc[c].v[k], m);
                    }
                }
            }
        }
        const float rc = __shfl_sync(0xffffffffu, cnt_l, r);
#pragma unroll
        for (int c = 0; c < CH; ++c) {
            if (!fok[c]) continue;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                float v = acc[c].v[k];
                if (a.row_cnt) v = __fdiv_rn(v, rc);
                if (MASK_END) v = rx[c].v[k] > 0.f ? v : 0.f;
                acc[c].v[k] = v;
            }
            if (a.row_add) {
                Vec<4> ad;
                ad.load_plain(a.row_add + row * a.row_add_ld + foff[c]);
#pragma unroll
                for (int k = 0; k < 4; ++k) acc[c].v[k] = __fadd_rn(acc[c].v[k], ad.v[k]);
            }
            st_stream_f4(a.out + row * a.out_ld + foff[c], acc[c].v, pol, hints);
        }
    }
}

// Hub rows.  A work item is (hub row, slice of feature columns): columns are independent, so slicing
// keeps every column's additions in edge order while a row spreads over several CTAs.  In a CTA 256
// loader threads (edge slots x lanes of 16 B) fetch a tile of edges, turn each fetched chunk into its
// message (the arithmetic of edge_group above) and stage it in shared memory; the ninth warp then adds
// the tile to its column sums one edge after the other while the loaders already have the next tile's
// loads in flight.  Two size classes share the 32 KB staging buffer:
//   hubs below HUB_BIG_EDGES: 32-column slices (one 128 B line per edge), 32 slots x U edges per tile --
//     a tile costs one memory round trip, which bounds the item;
//   bigger hubs: 8-column slices (one 32 B sector per edge), 128 slots x U edges per tile -- four times
//     the edges per round trip, so the item runs at the speed of its chain of dependent additions
//     (4 cycles per edge; 40 us for 19,000 edges), the floor for a sum in a prescribed order.
// Big hubs are listed from the front of the work list and small ones from its back, items are handed out
// front to back through a counter: longest first, and the order inside a class (atomics) does not matter.
// The last CTA to leave puts the four counters back to zero for the next call.
template <int MODE, bool BWD, bool RELU>
__global__ void __launch_bounds__(HUB_THREADS, 2)
aggregate_hub_kernel(const AggParams a) {
    constexpr int U = MODE == MODE_XE ? 4 : 8;
    static_assert((HUB_LOADERS / 2) * U * 8 <= HUB_TILE_FLOATS && (HUB_LOADERS / 8) * U * 32 <= HUB_TILE_FLOATS,
                  "tile does not fit the staging buffer");
    __shared__ float4 tile4[HUB_TILE_FLOATS / 4];
    __shared__ int s_item;
    const int tid = threadIdx.x;
    const bool loader = tid < HUB_LOADERS;
    const int F = a.F;
    // launched with programmatic stream serialisation: the CTAs may already be resident while the
    // row-per-group launch drains; its list and counters are visible from here on
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int n_big = a.hub[0], n_small = a.hub[1];
    if (n_big + n_small == 0) return;              // no hub in this structure: counters are still zero
    // 8-column slices only while 32-column ones would leave CTAs of this grid without an item: read by
    // sectors, a gathered row that has to come from DRAM costs more than its bytes
    const int slices_small = (F + 31) >> 5;
    const bool narrow = n_big * slices_small < (int)gridDim.x;
    const int slices_big = narrow ? (F + 7) >> 3 : slices_small;
    const int items_big = n_big * slices_big;
    const int n_items = items_big + n_small * slices_small;
    const int* list = a.hub + 4;

    int next_item = 0;                  // thread 0: fetched one item ahead, so the counter's round trip hides
    if (tid == 0) next_item = atomicAdd(a.hub + 2, 1);
    for (;;) {
        __syncthreads();                // everyone is done with s_item and the last tile of the previous item
        if (tid == 0) s_item = next_item;
        __syncthreads();
        const int item = s_item;
        if (item >= n_items) break;
        if (tid == 0) next_item = atomicAdd(a.hub + 2, 1);
        const bool big = item < items_big;
        const int j = big ? item : item - items_big;
        const int per = big ? slices_big : slices_small;
        const int64_t row = big ? list[j / per] : list[a.hub_cap - 1 - j / per];
        const int width = (big && narrow) ? 8 : 32; // columns of this slice
        const int lps = width >> 2;                 // 16 B lanes per edge slot
        const int slots = HUB_LOADERS / lps;        // edge slots per round
        const int tile_edges = slots * U;
        const int f0 = (j % per) * width;
        const int beg = __ldg(a.rowptr + row), end = __ldg(a.rowptr + row + 1);
        const int l = tid % lps, s = (tid / lps) % slots;
        const int c = tid - HUB_LOADERS;            // column of the slice a summing thread owns
        const int foff = f0 + l * 4;
        const bool fok = foff < F;          // F % 4 == 0 on this path: a 16 B chunk is inside or outside
        const bool col_ok = !loader && c < width && f0 + c < F;
        Vec<4> rx;
#pragma unroll
        for (int k = 0; k < 4; ++k) rx.v[k] = 0.f;
        if (loader && BWD && MODE == MODE_XE && RELU && fok) rx.load(a.row_x + row * a.row_x_ld + foff);
        float acc = (a.accumulate && col_ok) ? a.out[row * a.out_ld + f0 + c] : 0.f;
        // the slot indices of a tile are fetched while the rows of the tile before it are in flight, so a
        // tile costs one memory round trip instead of two dependent ones
        int cidx[U];
        int eid[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int e = beg + u * slots + s;
            const bool ok = loader && e < end && fok;
            cidx[u] = ok ? __ldg(a.col + e) : -1;
            eid[u] = (ok && MODE != MODE_PLAIN) ? __ldg(a.perm + e) : 0;
        }

        for (int p = beg; p < end; p += tile_edges) {
            Vec<4> m[U];
            if (loader) {
                float scal[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    scal[u] = 1.f;
                    if (cidx[u] >= 0) {
                        if (MODE == MODE_W) scal[u] = __ldg(a.edge_weight + eid[u]);
                        else if (BWD && a.gather_cnt) scal[u] = __ldg(a.gather_cnt + cidx[u]);
                    }
                }
                Vec<4> xe[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) { m[u].v[k] = 0.f; xe[u].v[k] = 0.f; }
                    if (cidx[u] >= 0) {
                        m[u].load(a.gathered + (int64_t)cidx[u] * F + foff);
                        if (MODE == MODE_XE) {
                            const float* xp = a.edge_feat ? a.edge_feat + (int64_t)eid[u] * F
                                                          : a.type_table + (int64_t)__ldg(a.edge_type + eid[u]) * F;
                            xe[u].load(xp + foff);
                        }
                    }
                }
#pragma unroll
                for (int u = 0; u < U; ++u) {           // next tile's slots; the loads above are already issued
                    const int e = p + tile_edges + u * slots + s;
                    const bool ok = e < end && fok;
                    cidx[u] = ok ? __ldg(a.col + e) : -1;
                    eid[u] = (ok && MODE != MODE_PLAIN) ? __ldg(a.perm + e) : 0;
                }
#pragma unroll
                for (int u = 0; u < U; ++u) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        float v = m[u].v[k];
                        if (MODE == MODE_W) {
                            v = __fmul_rn(scal[u], v);
                        } else if (!BWD) {
                            if (MODE == MODE_XE) v = __fadd_rn(v, xe[u].v[k]);
                            if (RELU) v = v > 0.f ? v : 0.f;
                        } else {
                            if (a.gather_cnt) v = __fdiv_rn(v, scal[u]);
                            if (MODE == MODE_XE && RELU) {
                                const float pre = __fadd_rn(rx.v[k], xe[u].v[k]);
                                v = pre > 0.f ? v : 0.f;
                            }
                        }
                        m[u].v[k] = v;
                    }
                }
            }
            __syncthreads();                    // the previous tile has been summed
            if (loader) {
#pragma unroll
                for (int u = 0; u < U; ++u)
                    tile4[(u * slots + s) * lps + l] = make_float4(m[u].v[0], m[u].v[1], m[u].v[2], m[u].v[3]);
            }
            __syncthreads();
            if (col_ok) {
                // edge i of the tile sits at float i * width + c; 16 values are fetched ahead of the 16
                // dependent additions of the batch before them
                const float* t = reinterpret_cast<const float*>(tile4) + c;
                const int n = min(tile_edges, end - p);
                const int batches = n >> 4;
                float cur[16];
                if (batches > 0) {
#pragma unroll
                    for (int q = 0; q < 16; ++q) cur[q] = t[q * width];
                }
                for (int b = 0; b < batches; ++b) {
                    float nxt[16];
                    const bool more = b + 1 < batches;
                    if (more) {
                        const float* tn = t + (b + 1) * 16 * width;
#pragma unroll
                        for (int q = 0; q < 16; ++q) nxt[q] = tn[q * width];
                    }
#pragma unroll
                    for (int q = 0; q < 16; ++q) acc = __fadd_rn(acc, cur[q]);
                    if (more) {
#pragma unroll
                        for (int q = 0; q < 16; ++q) cur[q] = nxt[q];
                    }
                }
                for (int i = batches << 4; i < n; ++i) acc = __fadd_rn(acc, t[i * width]);
            }
        }
        if (col_ok) {
            float r = acc;
            if (!BWD && a.row_cnt) r = __fdiv_rn(r, __ldg(a.row_cnt + row));
            if (BWD && MODE == MODE_PLAIN && RELU) r = __ldg(a.row_x + row * a.row_x_ld + f0 + c) > 0.f ? r : 0.f;
            if (a.row_add) r = __fadd_rn(r, a.row_add[row * a.row_add_ld + f0 + c]);
            a.out[row * a.out_ld + f0 + c] = r;
        }
    }
    if (tid == 0 && atomicAdd(a.hub + 3, 1) == (int)gridDim.x - 1) {
        // every CTA has read the counts and drawn its last item before it counted itself out
        a.hub[0] = 0; a.hub[1] = 0; a.hub[2] = 0; a.hub[3] = 0;
    }
}

// The same work items for the plain mode (message = the gathered row itself, relu at most): the loaders
// copy with cp.async straight into a ring of three 32 KB stages instead of staging in registers, so two
// tiles per CTA are in flight while a third is summed (the register version holds 32 registers of payload
// per loader and keeps one).  The relu moves into the summing warp; everything else is as above.
constexpr int HUB_RING_STAGES = 3;
constexpr int HUB_RING_BYTES = HUB_RING_STAGES * HUB_TILE_FLOATS * (int)sizeof(float);

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <bool BWD, bool RELU>
__global__ void __launch_bounds__(HUB_THREADS, 2)
aggregate_hub_ring_kernel(const AggParams a) {
    constexpr int U = 8;
    constexpr int NST = HUB_RING_STAGES;
    extern __shared__ float4 ring4[];
    __shared__ int s_item;
    const int tid = threadIdx.x;
    const bool loader = tid < HUB_LOADERS;
    const int F = a.F;
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int n_big = a.hub[0], n_small = a.hub[1];
    if (n_big + n_small == 0) return;
    const int slices_small = (F + 31) >> 5;
    const bool narrow = n_big * slices_small < (int)gridDim.x;
    const int slices_big = narrow ? (F + 7) >> 3 : slices_small;
    const int items_big = n_big * slices_big;
    const int n_items = items_big + n_small * slices_small;
    const int* list = a.hub + 4;

    int next_item = 0;
    if (tid == 0) next_item = atomicAdd(a.hub + 2, 1);
    for (;;) {
        __syncthreads();                // s_item and every stage of the previous item are free
        if (tid == 0) s_item = next_item;
        __syncthreads();
        const int item = s_item;
        if (item >= n_items) break;
        if (tid == 0) next_item = atomicAdd(a.hub + 2, 1);
        const bool big = item < items_big;
        const int j = big ? item : item - items_big;
        const int per = big ? slices_big : slices_small;
        const int64_t row = big ? list[j / per] : list[a.hub_cap - 1 - j / per];
        const int width = (big && narrow) ? 8 : 32;
        const int lps = width >> 2;
        const int slots = HUB_LOADERS / lps;
        const int tile_edges = slots * U;
        const int f0 = (j % per) * width;
        const int beg = __ldg(a.rowptr + row), end = __ldg(a.rowptr + row + 1);
        const int n_tiles = (end - beg + tile_edges - 1) / tile_edges;
        const int l = tid % lps, s = (tid / lps) % slots;
        const int c = tid - HUB_LOADERS;
        const int foff = f0 + l * 4;
        const bool fok = loader && foff < F;
        const bool col_ok = !loader && c < width && f0 + c < F;
        const float* src = a.gathered + foff;

        int cidx[U];                    // gathered rows of the tile to be issued next
        auto load_idx = [&](int tile) {
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int e = beg + tile * tile_edges + u * slots + s;
                cidx[u] = (fok && tile < n_tiles && e < end) ? __ldg(a.col + e) : -1;
            }
        };
        auto issue = [&](int tile) {
            float4* st = ring4 + (tile % NST) * (HUB_TILE_FLOATS / 4);
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (cidx[u] >= 0) cp_async16(st + (u * slots + s) * lps + l, src + (int64_t)cidx[u] * F);
            cp_async_commit();
        };
        // prologue: the slot indices of the first NST - 1 tiles in ONE round trip, then their copies
        static_assert(NST == 3, "the prologue below fills two stages");
        int first[U];
        load_idx(0);
#pragma unroll
        for (int u = 0; u < U; ++u) first[u] = cidx[u];
        load_idx(1);
        {
            float4* st = ring4;
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (first[u] >= 0) cp_async16(st + (u * slots + s) * lps + l, src + (int64_t)first[u] * F);
            cp_async_commit();
        }
        issue(1);
        load_idx(2);

        float acc = (a.accumulate && col_ok) ? a.out[row * a.out_ld + f0 + c] : 0.f;
        for (int t = 0; t < n_tiles; ++t) {
            cp_async_wait<NST - 2>();   // this thread's copies of tile t have landed
            __syncthreads();            // ... everyone's have, and tile t - 1 has been summed
            if (loader) {
                issue(t + NST - 1);     // into the stage tile t - 1 just left
                load_idx(t + NST);
            } else if (col_ok) {
                const float* tl = reinterpret_cast<const float*>(ring4 + (t % NST) * (HUB_TILE_FLOATS / 4)) + c;
                const int n = min(tile_edges, end - beg - t * tile_edges);
                const int batches = n >> 4;
                float cur[16];
                if (batches > 0) {
#pragma unroll
                    for (int q = 0; q < 16; ++q) cur[q] = tl[q * width];
                }
                for (int b = 0; b < batches; ++b) {
                    float nxt[16];
                    const bool more = b + 1 < batches;
                    if (more) {
                        const float* tn = tl + (b + 1) * 16 * width;
#pragma unroll
                        for (int q = 0; q < 16; ++q) nxt[q] = tn[q * width];
                    }
#pragma unroll
                    for (int q = 0; q < 16; ++q) {
                        float v = cur[q];
                        if (!BWD && RELU) v = v > 0.f ? v : 0.f;
                        acc = __fadd_rn(acc, v);
                    }
                    if (more) {
#pragma unroll
                        for (int q = 0; q < 16; ++q) cur[q] = nxt[q];
                    }
                }
                for (int i = batches << 4; i < n; ++i) {
                    float v = tl[i * width];
                    if (!BWD && RELU) v = v > 0.f ? v : 0.f;
                    acc = __fadd_rn(acc, v);
                }
            }
        }
        cp_async_wait<0>();             // only empty groups are left
        if (col_ok) {
            float r = acc;
            if (!BWD && a.row_cnt) r = __fdiv_rn(r, __ldg(a.row_cnt + row));
            if (BWD && RELU) r = __ldg(a.row_x + row * a.row_x_ld + f0 + c) > 0.f ? r : 0.f;
            if (a.row_add) r = __fadd_rn(r, a.row_add[row * a.row_add_ld + f0 + c]);
            a.out[row * a.out_ld + f0 + c] = r;
        }
    }
    if (tid == 0 && atomicAdd(a.hub + 3, 1) == (int)gridDim.x - 1) {
        a.hub[0] = 0; a.hub[1] = 0; a.hub[2] = 0; a.hub[3] = 0;
    }
}

// ---------------------------------------------------------------------------------------------
// Plain mode, one 16-byte chunk per lane (F <= 128): the EDGE-STREAM kernel with shared-memory staging.
//
// ncu on the row-per-warp kernel at F = 128 (arxiv-shaped layer 0): 2.0 TB/s of DRAM traffic, L2 hit 49 %, warps
// active 40 % -- bound by latency x bytes in flight: the gathered rows wait in registers (32 payload registers per
// thread, 64 in all -> 32 warps per SM, one 7-edge row each, and nothing in flight while a warp fetches rowptr / col).
// Here a warp owns 8 consecutive rows, i.e. ONE contiguous run of CSR slots, and streams that run in batches of 8
// edges whatever the row boundaries are: every lane copies its 16-byte column chunk of the 8 source rows with
// cp.async into a per-warp ring of two 4 KB slots, two batches always in flight, no payload registers (24 warps x
// 8 KB per SM).  A lane reads back only what it copied itself, so cp.async.wait_group is all the synchronisation
// there is.  Batches are consumed edge by edge in CSR order; when the edge counter reaches a row's end the row is
// finished (division, mask, row_add, store) and the next one starts -- the additions of a row are the same, in the
// same order, as in aggregate_kernel: same bits.  A run that contains a hub row (>= 256 edges, handed to the hub
// launch) is walked row by row instead.
constexpr int ES_BATCH = 8;                      // edges per batch
constexpr int ES_SMEM_BYTES = 64 * 1024;         // per CTA: (8 / CH) warps x 2 slots x 8 edges x CH x 512 B

// CH = 16-byte chunks per lane: 1 (F <= 128, 8 warps per CTA) or 2 (F <= 256, 4 warps per CTA, the same 64 KB)
template <int CH, bool RELU_MSG, bool MASK_END>
__global__ void __launch_bounds__(256 / CH, 3)
aggregate_edge_stream_kernel(const AggParams a) {
    extern __shared__ float4 es_ring[];
    constexpr int WPB = 8 / CH;                  // warps per CTA
    constexpr int SLOT4 = ES_BATCH * CH * 32;    // float4 per slot
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t r0 = ((int64_t)blockIdx.x * WPB + wib) * STREAM_ROWS;
    if (r0 >= a.n_rows) return;
    const int nr = (int)(a.n_rows - r0 < STREAM_ROWS ? a.n_rows - r0 : STREAM_ROWS);
    const int F = a.F;
    const bool hints = a.hints != 0;
    const uint64_t pol = policy_evict_first();
    const int E = (int)a.num_edges;
    int foff[CH];
    bool fok[CH];
#pragma unroll
    for (int c = 0; c < CH; ++c) { foff[c] = (c * 32 + lane) * 4; fok[c] = foff[c] < F; }
    float4* ring = es_ring + wib * (2 * SLOT4);

    int rp = 0;
    if (lane <= nr) rp = ld_stream_i32(a.rowptr + r0 + lane, pol, hints);
    float cnt_l = 1.f;
    if (a.row_cnt && lane < nr) cnt_l = __ldg(a.row_cnt + r0 + lane);
    const int e0 = __shfl_sync(0xffffffffu, rp, 0), e1 = __shfl_sync(0xffffffffu, rp, nr);

    // a hub row in the run: leave it to the hub launch and walk the other rows one by one (rare)
    bool has_hub = false;
    if (a.hub != nullptr) {
        const int nxt = __shfl_down_sync(0xffffffffu, rp, 1);
        has_hub = __any_sync(0xffffffffu, lane < nr && nxt - rp >= HUB_ROW_EDGES);
    }

    Vec<4> acc[CH], rx[CH];
    auto finish_row = [&](int r) {
        const int64_t row = r0 + r;
        const float rc = __shfl_sync(0xffffffffu, cnt_l, r);
#pragma unroll
        for (int c = 0; c < CH; ++c) {
            if (!fok[c]) continue;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                float v = acc[c].v[k];
                if (a.row_cnt) v = __fdiv_rn(v, rc);
                if (MASK_END) v = rx[c].v[k] > 0.f ? v : 0.f;
                acc[c].v[k] = v;
            }
            if (a.row_add) {
                Vec<4> ad;
                ad.load_plain(a.row_add + row * a.row_add_ld + foff[c]);
#pragma unroll
                for (int k = 0; k < 4; ++k) acc[c].v[k] = __fadd_rn(acc[c].v[k], ad.v[k]);
            }
            st_stream_f4(a.out + row * a.out_ld + foff[c], acc[c].v, pol, hints);
        }
    };
    auto start_row = [&](int r) {
        const int64_t row = r0 + r;
#pragma unroll
        for (int c = 0; c < CH; ++c) {
#pragma unroll
            for (int k = 0; k < 4; ++k) { acc[c].v[k] = 0.f; rx[c].v[k] = 0.f; }
            if (a.accumulate && fok[c]) acc[c].load_plain(a.out + row * a.out_ld + foff[c]);
            if (MASK_END && fok[c]) rx[c].load(a.row_x + row * a.row_x_ld + foff[c]);
        }
    };

    if (has_hub) {
        for (int r = 0; r < nr; ++r) {
            const int beg = __shfl_sync(0xffffffffu, rp, r), end = __shfl_sync(0xffffffffu, rp, r + 1);
            if (end - beg >= HUB_ROW_EDGES) {
                if (lane == 0) {
                    if (end - beg >= HUB_BIG_EDGES) a.hub[4 + atomicAdd(a.hub, 1)] = (int)(r0 + r);
                    else a.hub[4 + a.hub_cap - 1 - atomicAdd(a.hub + 1, 1)] = (int)(r0 + r);
                }
                continue;
            }
            start_row(r);
            for (int p = beg; p < end; ++p) {
                const int idx = __ldg(a.col + p);
#pragma unroll
                for (int c = 0; c < CH; ++c) {
                    if (!fok[c]) continue;
                    Vec<4> g;
                    g.load(a.gathered + (int64_t)idx * F + foff[c]);
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        float m = g.v[k];
                        if (RELU_MSG) m = m > 0.f ? m : 0.f;
                        acc[c].v[k] = __fadd_rn(acc[c].v[k], m);
                    }
                }
            }
            finish_row(r);
        }
        return;
    }

    // index window: w0 = col[wb + lane], w1 = col[wb + 32 + lane]
    int wb = e0 & ~31;
    int w0 = wb + lane < E ? ld_stream_i32(a.col + wb + lane, pol, hints) : 0;
    int w1 = wb + 32 + lane < E ? ld_stream_i32(a.col + wb + 32 + lane, pol, hints) : 0;
    const int nb = (e1 - e0 + ES_BATCH - 1) / ES_BATCH;
    auto issue = [&](int b) {                    // copies of batch b into slot b & 1 (an empty group past the end)
        if (b < nb) {
            const int p0 = e0 + b * ES_BATCH;
            if (p0 >= wb + 32) {                 // the first block is used up: slide (the new block is needed 3 batches on)
                w0 = w1;
                wb += 32;
                w1 = wb + 32 + lane < E ? ld_stream_i32(a.col + wb + 32 + lane, pol, hints) : 0;
            }
            float4* slot = ring + (b & 1) * SLOT4;
#pragma unroll
            for (int u = 0; u < ES_BATCH; ++u) {
                const int i = p0 + u - wb;       // warp-uniform, in [0, 64)
                const int idx = __shfl_sync(0xffffffffu, i < 32 ? w0 : w1, i & 31);
                const float* srow = a.gathered + (int64_t)idx * F;
#pragma unroll
                for (int c = 0; c < CH; ++c)
                    if (p0 + u < e1 && fok[c]) cp_async16(slot + (u * CH + c) * 32 + lane, srow + foff[c]);
            }
        }
        cp_async_commit();
    };
    issue(0);
    issue(1);
    int r = 0;
    int row_end = __shfl_sync(0xffffffffu, rp, 1);
    start_row(0);
    for (int b = 0; b < nb; ++b) {
        cp_async_wait<1>();                      // this lane's copies of batch b have landed
        const float4* slot = ring + (b & 1) * SLOT4;
        const int p0 = e0 + b * ES_BATCH;
        float4 v[ES_BATCH][CH];
#pragma unroll
        for (int u = 0; u < ES_BATCH; ++u)
#pragma unroll
            for (int c = 0; c < CH; ++c) v[u][c] = slot[(u * CH + c) * 32 + lane];
        issue(b + 2);                            // into the slot just read (its copies land hundreds of cycles later)
#pragma unroll
        for (int u = 0; u < ES_BATCH; ++u) {
            const int p = p0 + u;
            if (p >= e1) break;
            while (p == row_end) {               // rows that ended (empty ones included)
                finish_row(r);
                ++r;
                row_end = __shfl_sync(0xffffffffu, rp, r + 1);
                start_row(r);
            }
#pragma unroll
            for (int c = 0; c < CH; ++c) {
                const float m[4] = {v[u][c].x, v[u][c].y, v[u][c].z, v[u][c].w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    float t = m[k];
                    if (RELU_MSG) t = t > 0.f ? t : 0.f;
                    acc[c].v[k] = __fadd_rn(acc[c].v[k], t);
                }
            }
        }
    }
    cp_async_wait<0>();
    for (;;) {                                   // the last row with edges and the empty rows behind it
        finish_row(r);
        if (++r >= nr) break;
        start_row(r);
    }
}

// Second launch of an aggregation whose caller lent a hub scratch.  The number of hubs is only known on
// the device: a fixed grid (two CTAs per SM, fewer when the edge count bounds the work items below that)
// drains the list, and leaves at once if it is empty.  It is a programmatic dependent launch: the grid's
// launch latency overlaps the first launch's tail (it waits at griddepcontrol.wait), which is what the
// second launch costs a structure without hubs.
template <int MODE, bool BWD, bool RELU>
static void launch_hub(const AggParams& p, cudaStream_t s) {
    static const bool pdl = [] { const char* e = getenv("STEMGNN_AGG_PDL"); return e ? atoi(e) != 0 : true; }();
    static const bool ring_on = [] { const char* e = getenv("STEMGNN_AGG_HUB_RING"); return e ? atoi(e) != 0 : true; }();
    const int64_t items = (p.num_edges / HUB_ROW_EDGES) * ((p.F + 7) / 8);
    const int64_t full = 2 * (int64_t)sm_count();
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(items < full ? items : full));
    cfg.blockDim = dim3(HUB_THREADS);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    if constexpr (MODE == MODE_PLAIN) {
        static const bool ring_ok = cudaFuncSetAttribute(aggregate_hub_ring_kernel<BWD, RELU>,
                                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                         HUB_RING_BYTES) == cudaSuccess;
        if (ring_on && ring_ok && !(BWD && p.gather_cnt)) {
            cfg.dynamicSmemBytes = HUB_RING_BYTES;
            SG_LAUNCH(cudaLaunchKernelEx(&cfg, aggregate_hub_ring_kernel<BWD, RELU>, p));
            return;
        }
    }
    SG_LAUNCH(cudaLaunchKernelEx(&cfg, aggregate_hub_kernel<MODE, BWD, RELU>, p));
}

template <int VEC, int LANES, int CH, int MODE, bool BWD, bool RELU>
static void launch(const AggParams& p, cudaStream_t s) {
    constexpr int ROWS_PER_BLOCK = 8 * (32 / LANES);
    unsigned grid = (unsigned)((p.n_rows + ROWS_PER_BLOCK - 1) / ROWS_PER_BLOCK);
    SG_LAUNCH(aggregate_kernel<VEC, LANES, CH, MODE, BWD, RELU><<<grid, 256, 0, s>>>(p));
    if (VEC == 4 && p.hub != nullptr) launch_hub<MODE, BWD, RELU>(p, s);
}

template <int CH, bool BWD, bool RELU>
static void launch_stream(const AggParams& p, cudaStream_t s) {
    constexpr int ROWS_PER_BLOCK = 8 * STREAM_ROWS;
    unsigned grid = (unsigned)((p.n_rows + ROWS_PER_BLOCK - 1) / ROWS_PER_BLOCK);
    // resident 256-thread blocks the register allocation must allow; one 16 B chunk per lane (F <= 128) is
    // measured at 4 (64 registers, a few spilled words) against 3 -- STEMGNN_AGG_STREAM_OCC picks for the A/B
    static const int occ1 = [] { const char* e = getenv("STEMGNN_AGG_STREAM_OCC"); return e ? atoi(e) : 3; }();
    bool done = false;
    if constexpr (CH == 1) {
        if (occ1 == 4) {
            SG_LAUNCH(aggregate_stream_kernel<CH, !BWD && RELU, BWD && RELU, 4><<<grid, 256, 0, s>>>(p));
            done = true;
        }
    }
    if (!done) SG_LAUNCH(aggregate_stream_kernel<CH, !BWD && RELU, BWD && RELU, (CH <= 6 ? 3 : 2)><<<grid, 256, 0, s>>>(p));
    if (p.hub != nullptr) launch_hub<MODE_PLAIN, BWD, RELU>(p, s);
}

// Widths above this take aggregate_stream_kernel.  Measured on B200 (profiles/r02_agg_stream_ab.log): F = 256
// 210 -> 179 us forward / 255 -> 224 backward on the arxiv-shaped graph, F = 512 839 -> 750; one 16 B chunk per
// lane (F <= 128) gains nothing (124 us either way) and loses 19 % on 25-edge rows at F = 100 (2,122 -> 2,528 us,
// products-sized half graph), so those widths stay on aggregate_kernel.  STEMGNN_AGG_STREAM=0 keeps every shape
// there, =1 moves the threshold down to 64, =<n> to n (A/B switch for scripts/agg_bench.py).
static int stream_min_f() {
    static const int f = [] {
        const char* e = getenv("STEMGNN_AGG_STREAM");
        if (!e) return 128;
        const int v = atoi(e);
        return v == 0 ? (1 << 30) : (v == 1 ? 64 : v);
    }();
    return f;
}

template <int CH, bool BWD, bool RELU>
static bool launch_edge_stream(const AggParams& p, cudaStream_t s) {
    static const bool ok = cudaFuncSetAttribute(aggregate_edge_stream_kernel<CH, !BWD && RELU, BWD && RELU>,
                                                cudaFuncAttributeMaxDynamicSharedMemorySize, ES_SMEM_BYTES) == cudaSuccess;
    if (!ok) return false;
    constexpr int ROWS_PER_BLOCK = (8 / CH) * STREAM_ROWS;
    const unsigned grid = (unsigned)((p.n_rows + ROWS_PER_BLOCK - 1) / ROWS_PER_BLOCK);
    SG_LAUNCH(aggregate_edge_stream_kernel<CH, !BWD && RELU, BWD && RELU><<<grid, 256 / CH, ES_SMEM_BYTES, s>>>(p));
    if (p.hub != nullptr) launch_hub<MODE_PLAIN, BWD, RELU>(p, s);
    return true;
}

// plain messages: the cp.async edge-stream kernel for 64 < F <= 128 (STEMGNN_AGG_EDGE_STREAM=0 keeps aggregate_kernel)
// and, with STEMGNN_AGG_EDGE_STREAM=2, for 128 < F <= 256 as well (A/B against aggregate_stream_kernel)
static int edge_stream_mode() {
    static const int m = [] { const char* e = getenv("STEMGNN_AGG_EDGE_STREAM"); return e ? atoi(e) : 1; }();
    return m;
}

template <int MODE, bool BWD, bool RELU>
static void dispatch_shape(const AggParams& p, bool vec4, cudaStream_t s) {
    if (!vec4) { launch<1, 32, 8, MODE, BWD, RELU>(p, s); return; }
    if constexpr (MODE == MODE_PLAIN) {
        if (!(BWD && p.gather_cnt)) {
            if (edge_stream_mode() >= 1 && p.F > 64 && p.F <= 128 && launch_edge_stream<1, BWD, RELU>(p, s)) return;
            if (edge_stream_mode() >= 2 && p.F > 128 && p.F <= 256 && launch_edge_stream<2, BWD, RELU>(p, s)) return;
        }
        if (p.F > stream_min_f() && p.F <= 1024 && !(BWD && p.gather_cnt)) {
            const int ch = (p.F + 127) / 128;
            if (ch <= 1) launch_stream<1, BWD, RELU>(p, s);
            else if (ch <= 2) launch_stream<2, BWD, RELU>(p, s);
            else if (ch <= 3) launch_stream<3, BWD, RELU>(p, s);
            else if (ch <= 4) launch_stream<4, BWD, RELU>(p, s);
            else if (ch <= 6) launch_stream<6, BWD, RELU>(p, s);
            else launch_stream<8, BWD, RELU>(p, s);
            return;
        }
    }
    const int f4 = p.F / 4;
    if (f4 <= 8) launch<4, 8, 1, MODE, BWD, RELU>(p, s);
    else if (f4 <= 16) launch<4, 16, 1, MODE, BWD, RELU>(p, s);
    else if (f4 <= 32) launch<4, 32, 1, MODE, BWD, RELU>(p, s);
    else if (f4 <= 64) launch<4, 32, 2, MODE, BWD, RELU>(p, s);
    else if (f4 <= 96) launch<4, 32, 3, MODE, BWD, RELU>(p, s);
    else if (f4 <= 128) launch<4, 32, 4, MODE, BWD, RELU>(p, s);
    else if (f4 <= 192) launch<4, 32, 6, MODE, BWD, RELU>(p, s);
    else launch<4, 32, 8, MODE, BWD, RELU>(p, s);
}

template <int MODE, bool BWD>
static void dispatch_act(const AggParams& p, bool vec4, cudaStream_t s) {
    if (p.relu) dispatch_shape<MODE, BWD, true>(p, vec4, s);
    else dispatch_shape<MODE, BWD, false>(p, vec4, s);
}

// out[r, :] = x[r, :] / cnt[r]; one float4 (or one float) per thread, grid-stride.
template <int VEC>
__global__ void __launch_bounds__(256)
row_scale_kernel(const float* __restrict__ x, const float* __restrict__ cnt, float* __restrict__ out,
                 const int64_t n_vec, const int vec_per_row) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec; i += (int64_t)gridDim.x * blockDim.x) {
        const float c = __ldg(cnt + i / vec_per_row);
        Vec<VEC> v;
        v.load(x + i * VEC);
#pragma unroll
        for (int k = 0; k < VEC; ++k) v.v[k] = __fdiv_rn(v.v[k], c);
        v.store(out + i * VEC);
    }
}

}  // namespace stemgnn

using namespace stemgnn;

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// four counters + one slot per possible hub (disjoint rows of >= HUB_ROW_EDGES edges each)
static size_t hub_ws_bytes(int64_t num_edges) { return sizeof(int) * (size_t)(4 + num_edges / HUB_ROW_EDGES + 1); }

extern "C" int stemgnn_aggregate_hub_workspace(int64_t num_edges, size_t* bytes) {
    SG_CHECK_ARG(bytes != nullptr && num_edges >= 0, "bytes is NULL or negative size");
    *bytes = hub_ws_bytes(num_edges);
    return STEMGNN_OK;
}

extern "C" int stemgnn_aggregate(const stemgnn_aggregate_args* a, stemgnn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SG_CHECK_ARG(a != nullptr, "args is NULL");
    SG_CHECK_ARG(a->n_rows >= 0 && a->feat >= 0 && a->num_edges >= 0, "negative size");
    SG_CHECK_ARG(a->feat < (1 << 30), "feat too large");
    if (a->n_rows == 0 || a->feat == 0) return STEMGNN_OK;
    SG_CHECK_ARG(a->rowptr && a->out, "rowptr/out is NULL");
    SG_CHECK_ARG(a->num_edges == 0 || (a->col && a->gathered), "col/gathered is NULL");
    const bool has_xe = a->edge_feat || (a->type_table && a->edge_type);
    const bool has_w = a->edge_weight != nullptr;
    SG_CHECK_ARG(!(has_xe && has_w), "edge features and edge weights cannot be combined");
    SG_CHECK_ARG(!(has_w && a->act != STEMGNN_ACT_NONE), "weighted mode has no activation");
    SG_CHECK_ARG(!((has_xe || has_w) && a->num_edges > 0 && !a->perm), "perm required with per-edge data");
    SG_CHECK_ARG(!(a->backward && a->act == STEMGNN_ACT_RELU && !a->row_x), "backward relu needs row_x");
    SG_CHECK_ARG(!(a->backward && a->row_cnt), "row_cnt is forward-only");
    SG_CHECK_ARG(!(!a->backward && a->gather_cnt), "gather_cnt is backward-only");

    AggParams p;
    p.rowptr = a->rowptr; p.col = a->col; p.perm = a->perm;
    p.n_rows = a->n_rows; p.num_edges = a->num_edges; p.F = (int)a->feat;
    p.gathered = a->gathered; p.row_cnt = a->row_cnt; p.gather_cnt = a->gather_cnt;
    p.edge_feat = a->edge_feat; p.type_table = a->type_table; p.edge_type = a->edge_type;
    p.edge_weight = a->edge_weight; p.row_x = a->row_x;
    p.relu = a->act == STEMGNN_ACT_RELU; p.out = a->out;
    p.out_ld = a->out_ld ? a->out_ld : a->feat;
    p.row_x_ld = a->row_x_ld ? a->row_x_ld : a->feat;
    SG_CHECK_ARG(p.out_ld >= a->feat && p.row_x_ld >= a->feat, "out_ld / row_x_ld smaller than feat");

    bool vec4 = (a->feat % 4 == 0) && aligned16(a->gathered) && aligned16(a->out) &&
                aligned16(a->edge_feat) && aligned16(a->type_table) && aligned16(a->row_x) && aligned16(a->row_add) &&
                p.out_ld % 4 == 0 && p.row_x_ld % 4 == 0 && p.row_add_ld % 4 == 0;
    // hub rows (>= HUB_ROW_EDGES edges) go through the work list when the caller lends one; without it,
    // or on the unaligned scalar path, every row stays in the row-per-group kernel
    p.hub = nullptr; p.hub_cap = 0;
    if (a->hub_ws != nullptr && vec4 && a->num_edges >= HUB_ROW_EDGES) {
        SG_CHECK_ARG(a->hub_ws_bytes >= hub_ws_bytes(a->num_edges), "hub_ws smaller than stemgnn_aggregate_hub_workspace");
        p.hub = static_cast<int*>(a->hub_ws);
        p.hub_cap = (int)(a->num_edges / HUB_ROW_EDGES + 1);
    }
    p.accumulate = a->accumulate != 0;
    p.row_add = a->row_add;
    p.row_add_ld = a->row_add_ld ? a->row_add_ld : a->feat;
    SG_CHECK_ARG(!a->row_add || p.row_add_ld >= a->feat, "row_add_ld smaller than feat");
    static const bool hints_on = [] { const char* e = getenv("STEMGNN_AGG_HINTS"); return e ? atoi(e) != 0 : true; }();
    p.hints = hints_on;
    const int mode = has_w ? MODE_W : (has_xe ? MODE_XE : MODE_PLAIN);
    // a backward pass with neither a per-edge division nor a relu mask is the plain sum over the
    // transposed structure: the forward kernel
    const bool bwd = a->backward && (a->gather_cnt || p.relu);
    if (mode == MODE_W) dispatch_shape<MODE_W, false, false>(p, vec4, stream);
    else if (mode == MODE_XE) {
        if (bwd) dispatch_act<MODE_XE, true>(p, vec4, stream);
        else dispatch_act<MODE_XE, false>(p, vec4, stream);
    } else {
        if (bwd) dispatch_act<MODE_PLAIN, true>(p, vec4, stream);
        else dispatch_act<MODE_PLAIN, false>(p, vec4, stream);
    }
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}

extern "C" int stemgnn_row_scale(int64_t n_rows, int64_t feat, const float* x, const float* cnt, float* out,
                                 stemgnn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SG_CHECK_ARG(n_rows >= 0 && feat >= 0, "negative size");
    if (n_rows == 0 || feat == 0) return STEMGNN_OK;
    SG_CHECK_ARG(x && cnt && out, "x/cnt/out is NULL");
    const bool vec4 = (feat % 4 == 0) && aligned16(x) && aligned16(out);
    const int64_t n_vec = vec4 ? n_rows * feat / 4 : n_rows * feat;
    const int64_t want = (n_vec + 255) / 256;
    const unsigned grid = (unsigned)(want < 148 * 16 ? want : 148 * 16);
    if (vec4) SG_LAUNCH(row_scale_kernel<4><<<grid, 256, 0, stream>>>(x, cnt, out, n_vec, (int)(feat / 4)));
    else SG_LAUNCH(row_scale_kernel<1><<<grid, 256, 0, stream>>>(x, cnt, out, n_vec, (int)feat));
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}
