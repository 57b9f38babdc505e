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
#include "common.cuh"

namespace stemgnn {

enum AggMode { MODE_PLAIN = 0, MODE_XE = 1, MODE_W = 2 };

template <int VEC> struct Vec;
template <> struct Vec<4> {
    float v[4];
    __device__ __forceinline__ void load(const float* p) {
        float4 t = __ldg(reinterpret_cast<const float4*>(p));
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
    __device__ __forceinline__ void store(float* p) const {
        *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    }
};
template <> struct Vec<1> {
    float v[1];
    __device__ __forceinline__ void load(const float* p) { v[0] = __ldg(p); }
    __device__ __forceinline__ void store(float* p) const { *p = v[0]; }
};

struct AggParams {
    const int32_t* rowptr; const int32_t* col; const int32_t* perm;
    int64_t n_rows; int F;
    const float* gathered; const float* row_cnt; const float* gather_cnt;
    const float* edge_feat; const float* type_table; const int32_t* edge_type;
    const float* edge_weight; const float* row_x;
    int relu; float* out;
    int64_t out_ld, row_x_ld;   // row strides (floats) of out / row_x: column slices of wider matrices
};

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
            acc[c].store(a.out + row * a.out_ld + foff[c]);
        }
    }
}

template <int VEC, int LANES, int CH, int MODE, bool BWD, bool RELU>
static void launch(const AggParams& p, cudaStream_t s) {
    constexpr int ROWS_PER_BLOCK = 8 * (32 / LANES);
    unsigned grid = (unsigned)((p.n_rows + ROWS_PER_BLOCK - 1) / ROWS_PER_BLOCK);
    SG_LAUNCH(aggregate_kernel<VEC, LANES, CH, MODE, BWD, RELU><<<grid, 256, 0, s>>>(p));
}

template <int MODE, bool BWD, bool RELU>
static void dispatch_shape(const AggParams& p, bool vec4, cudaStream_t s) {
    if (!vec4) { launch<1, 32, 8, MODE, BWD, RELU>(p, s); return; }
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
    p.n_rows = a->n_rows; p.F = (int)a->feat;
    p.gathered = a->gathered; p.row_cnt = a->row_cnt; p.gather_cnt = a->gather_cnt;
    p.edge_feat = a->edge_feat; p.type_table = a->type_table; p.edge_type = a->edge_type;
    p.edge_weight = a->edge_weight; p.row_x = a->row_x;
    p.relu = a->act == STEMGNN_ACT_RELU; p.out = a->out;
    p.out_ld = a->out_ld ? a->out_ld : a->feat;
    p.row_x_ld = a->row_x_ld ? a->row_x_ld : a->feat;
    SG_CHECK_ARG(p.out_ld >= a->feat && p.row_x_ld >= a->feat, "out_ld / row_x_ld smaller than feat");

    bool vec4 = (a->feat % 4 == 0) && aligned16(a->gathered) && aligned16(a->out) &&
                aligned16(a->edge_feat) && aligned16(a->type_table) && aligned16(a->row_x) &&
                p.out_ld % 4 == 0 && p.row_x_ld % 4 == 0;
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
