// k5_linear_tc.cu -- K5: the dense layers of the conv on the tcgen05 tensor cores at fp32 accuracy.
//
// Replaces, for MySAGEConv (reference STEM-GNN/model/encoder.py:84-88)
//     out = lin_l(aggregated) + lin_r(x)            -> ONE call, two (A, W) sources, one bias
// and its backward-data products g @ W, plus the single Linear of the GCN / GIN layers
// (encoder.py:196-199).  torch runs these as SIMT fp32 SGEMMs (no tensor cores for fp32 inputs):
// 55 % of the ogbn-arxiv step.  SURVEY.md 8f-1.
//
//   out[m, n] = sum_s sum_k A_s[m, k] * W_s[n, k]  (+ bias[n])          fp32 in, fp32 out
//
// Accuracy: "3xTF32".  Every operand is split x = hi + lo with hi = tf32(x) (round to nearest),
// lo = x - hi (exact in fp32; the MMA reads its tf32 part), and three kind::tf32 MMAs accumulate
// hi*hi + lo*hi + hi*lo in fp32 TMEM accumulators.  The dropped lo*lo term and the rounding of lo
// are 2^-22 relative: the result sits within ~1e-6 of the fp32 product (tests hold 1e-5, the
// north_star bound) with the full fp32 exponent range -- no scaling pass, no overflow cases.
//
// A persistent CTA owns a 128-row tile with a double-buffered 2 x 256-column TMEM accumulator (the
// epilogue of tile t overlaps the MMAs of tile t+1), or -- for wide layers whose split weight
// image is several MB -- a 256-row tile whose two M=128 accumulators share every W stage, which
// halves the L2 traffic of the weight stream (Shape<MH>, chosen per call).  Eight
// producer warps read the fp32 rows of A once from HBM, split them and store the hi / lo images
// K-major with the 64-byte swizzle (16 tf32 per row: a 48 or 64 KB stage holds A hi/lo + W hi/lo,
// four or three stages fit).  W is split once per call into the same swizzled stage images in global memory and
// streamed by one elected thread with 1-D bulk async copies.  One elected thread issues the MMAs;
// eight epilogue warps drain TMEM through a padded shared-memory transpose so that every global
// store instruction writes full 128-byte lines.
// FLOPs: 3 * 2*M*N*K on the tensor pipe (tf32); HBM: M*K*4 (A) + M*N*4 (out).
#include <cstdlib>

#include "common.cuh"
#include "tc_ptx.cuh"

namespace stemgnn {
namespace lin {

using namespace tcptx;

constexpr int BN = 256;                  // output columns per accumulator = max UMMA N
constexpr int BK = 16;                   // tf32 per smem row (64 B, SWIZZLE_64B)
constexpr int IMG_BYTES = 256 * 64;      // one 256-row operand image of a stage: 16 KB
constexpr int W_STAGE_BYTES = 2 * IMG_BYTES;
// Tile shape: MH = UMMA M=128 halves per tile.
//   MH = 2: 256-row tiles, both halves share every W stage (half the L2 traffic of the weight
//           stream), accumulators use all 512 TMEM columns -> the epilogue is NOT overlapped;
//   MH = 1: 128-row tiles, 2 x 256-column accumulators double-buffered -> the epilogue of tile t
//           overlaps the MMAs of tile t+1, twice as many tiles (finer tail), W streamed twice as often.
template <int MH> struct Shape {
    static constexpr int BM = 128 * MH;
    static constexpr int A_IMG = BM * 64;                        // hi or lo image of the A rows of a stage
    static constexpr int STAGE_BYTES = 2 * A_IMG + W_STAGE_BYTES;   // A hi, A lo, W hi, W lo
    static constexpr int NS = MH == 2 ? 3 : 4;                   // stages
    static constexpr int NBUF = MH == 2 ? 1 : 2;                 // accumulator buffers
    static constexpr int SMEM_BYTES = NS * STAGE_BYTES + 8 * (32 * 33) * 4 + 24 * 8 + 16;
};
constexpr int EPI_WARPS = 8, PROD_WARP0 = 8, LOAD_WARP = 16, MMA_WARP = 17;
constexpr int NUM_THREADS = 576;
constexpr int STAGING_WORDS = 32 * 33;   // per epilogue warp: 32 rows x 32 columns, padded

struct LinParams {
    const float* a[2];
    int64_t lda[2];
    int k[2];
    int kc[2];              // 16-column chunks per source
    int kct;                // chunks per tile
    int64_t m;
    int n, n_tiles;
    const float* wimg;      // [n_tiles][kct][W hi image | W lo image]
    const float* bias;
    float* out;
    int64_t ldo;
    int accumulate;         // out += instead of out =
    const float* row_div;   // out[m, :] = (product + bias) / row_div[m] (IEEE division), or NULL.  ABI 6
};

// byte offset of the 16 B unit j of row r in a 256-row SWIZZLE_64B image
__device__ __host__ __forceinline__ uint32_t sw64_off(int r, int j) {
    return (uint32_t)((r >> 3) * 512 + (r & 7) * 64 + ((j ^ ((r >> 1) & 3)) << 4));
}

// ---- W -> split, swizzled stage images (once per call; W is at most a few hundred KB) -----------
struct WPrep {
    const float* w[2];
    int64_t ldw[2];
    int k[2], kc[2], kct;
    int n, n_tiles, transposed;
    float* img;
};

__global__ void __launch_bounds__(256) lin_w_image_kernel(const WPrep p) {
    // one thread per (n_tile, chunk, row, 16 B unit)
    const int64_t total = (int64_t)p.n_tiles * p.kct * 256 * 4;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int j = (int)(i & 3), r = (int)((i >> 2) & 255);
        const int64_t sc = i >> 10;
        const int c = (int)(sc % p.kct), nt = (int)(sc / p.kct);
        const int s = c < p.kc[0] ? 0 : 1;
        const int kbase = (s == 0 ? c : c - p.kc[0]) * BK + 4 * j;
        const int n = nt * BN + r;
        float hi[4], lo[4];
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            float v = 0.f;
            const int k = kbase + t;
            if (n < p.n && k < (s ? p.k[1] : p.k[0]))
                v = p.transposed ? __ldg((s ? p.w[1] : p.w[0]) + (int64_t)k * (s ? p.ldw[1] : p.ldw[0]) + n) : __ldg((s ? p.w[1] : p.w[0]) + (int64_t)n * (s ? p.ldw[1] : p.ldw[0]) + k);
            tf32_split(v, hi[t], lo[t]);
        }
        uint8_t* st = reinterpret_cast<uint8_t*>(p.img) + ((int64_t)nt * p.kct + c) * W_STAGE_BYTES + sw64_off(r, j);
        *reinterpret_cast<float4*>(st) = make_float4(hi[0], hi[1], hi[2], hi[3]);
        *reinterpret_cast<float4*>(st + IMG_BYTES) = make_float4(lo[0], lo[1], lo[2], lo[3]);
    }
}

// ---- the GEMM ------------------------------------------------------------------------------------
template <int MH>
__global__ void __launch_bounds__(NUM_THREADS, 1) lin_tc_kernel(const LinParams p) {
    using SH = Shape<MH>;
    constexpr int BM = SH::BM, NS = SH::NS, NBUF = SH::NBUF, STAGE_BYTES = SH::STAGE_BYTES, A_IMG = SH::A_IMG;
    constexpr int W_OFF = 2 * A_IMG;      // W hi image inside a stage; W lo follows
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* stages = smem;
    float* s_stage = reinterpret_cast<float*>(smem + NS * STAGE_BYTES);      // [8][32 x 33]
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_stage + EPI_WARPS * STAGING_WORDS);
    uint64_t* a_full = bars;              // [NS] 128 producer arrivals
    uint64_t* w_full = bars + NS;         // [NS] loader expect_tx
    uint64_t* empty = bars + 2 * NS;      // [NS] tcgen05.commit
    uint64_t* acc_full = bars + 3 * NS;   // [NBUF] tcgen05.commit after the last chunk of a tile
    uint64_t* acc_empty = acc_full + 2;   // [NBUF] 256 epilogue arrivals
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 24);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t row_tiles = (p.m + BM - 1) / BM;
    const int64_t n_items = row_tiles * p.n_tiles;

    if (threadIdx.x == 0) {
        for (int i = 0; i < NS; ++i) {
            mbar_init(smem_u32(a_full + i), 128);
            mbar_init(smem_u32(w_full + i), 1);
            mbar_init(smem_u32(empty + i), 1);
        }
        for (int i = 0; i < NBUF; ++i) {
            mbar_init(smem_u32(acc_full + i), 1);
            mbar_init(smem_u32(acc_empty + i), EPI_WARPS * 32);
        }
        fence_barrier_init();
    }
    if (warp == MMA_WARP) tmem_alloc(smem_u32(tmem_slot), 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == LOAD_WARP) {
        // ================= W loader: one bulk copy per stage =================
        if (lane == 0) {
            uint32_t seq = 0;
            for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x) {
                const int nt = (int)(item % p.n_tiles);
                const uint8_t* src = reinterpret_cast<const uint8_t*>(p.wimg) + (int64_t)nt * p.kct * W_STAGE_BYTES;
                for (int c = 0; c < p.kct; ++c, ++seq) {
                    const uint32_t slot = seq % NS;
                    mbar_wait(smem_u32(empty + slot), ((seq / NS) & 1u) ^ 1u);
                    mbar_expect_tx(smem_u32(w_full + slot), W_STAGE_BYTES);
                    bulk_g2s(smem_u32(stages + slot * STAGE_BYTES + W_OFF), src + (int64_t)c * W_STAGE_BYTES,
                             W_STAGE_BYTES, smem_u32(w_full + slot));
                }
            }
        }
    } else if (warp == MMA_WARP) {
        // ================= MMA issue =================
        // The whole warp walks the loop (warp-uniform control flow: descriptors and barrier
        // addresses stay in uniform registers, no per-instruction predicate loops); one elected
        // lane issues the MMAs and commits.
        {
            uint32_t seq = 0, item_it = 0;
            for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x, ++item_it) {
                const int nt = (int)(item % p.n_tiles);
                int ncols = p.n - nt * BN;
                ncols = ncols > BN ? BN : ncols;
                const int n_mma = ncols < 16 ? 16 : ((ncols + 15) & ~15);
                const uint32_t idesc = make_idesc_tf32(128, n_mma, 0, 0);
                const uint32_t buf = item_it % NBUF;
                mbar_wait(smem_u32(acc_empty + buf), ((item_it / NBUF) & 1u) ^ 1u);
                for (int c = 0; c < p.kct; ++c, ++seq) {
                    const uint32_t slot = seq % NS, ph = (seq / NS) & 1u;
                    mbar_wait(smem_u32(a_full + slot), ph);
                    mbar_wait(smem_u32(w_full + slot), ph);
                    tc_fence_after();
                    const uint32_t base = smem_u32(stages + slot * STAGE_BYTES);
                    if (elect_one()) {
#pragma unroll
                        for (int h = 0; h < MH; ++h) {
#pragma unroll
                            for (int ks = 0; ks < 2; ++ks) {
                                const uint64_t a_hi = smem_desc_sw64(base + h * 8192 + ks * 32);
                                const uint64_t a_lo = smem_desc_sw64(base + A_IMG + h * 8192 + ks * 32);
                                const uint64_t w_hi = smem_desc_sw64(base + W_OFF + ks * 32);
                                const uint64_t w_lo = smem_desc_sw64(base + W_OFF + IMG_BYTES + ks * 32);
                                const uint32_t d = tmem_base + (uint32_t)(MH == 2 ? h : buf) * BN;
                                umma_tf32(d, a_lo, w_hi, idesc, (c > 0 || ks > 0) ? 1u : 0u);
                                umma_tf32(d, a_hi, w_lo, idesc, 1u);
                                umma_tf32(d, a_hi, w_hi, idesc, 1u);
                            }
                        }
                        umma_commit(smem_u32(empty + slot));
                        if (c == p.kct - 1) umma_commit(smem_u32(acc_full + buf));
                    }
                    __syncwarp();
                }
            }
        }
    } else if (warp >= PROD_WARP0 && warp < LOAD_WARP) {
        // ================= producers: fp32 rows -> tf32 hi / lo images =================
        // Two groups of 128 threads take alternate chunks.  Thread (j, rr) of a group owns the
        // 16 B unit j of rows rr, rr+32, ... of the tile: a quad reads 64 contiguous bytes per row.
        constexpr int RPT = 4 * MH;       // rows per thread and chunk
        const int pt = (threadIdx.x - PROD_WARP0 * 32) & 127;
        const int grp = (threadIdx.x - PROD_WARP0 * 32) >> 7;
        const int j = pt & 3, rr = pt >> 2;
        const uint32_t soff = sw64_off(rr, j);          // rows rr + 32 i: + i * 2048
        uint32_t item_it = 0;
        for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x, ++item_it) {
            const int64_t row0 = (item / p.n_tiles) * BM;
#pragma unroll 1
            for (int c = grp; c < p.kct; c += 2) {
                const uint32_t seq = item_it * (uint32_t)p.kct + (uint32_t)c;
                const uint32_t slot = seq % NS;
                const int s = c < p.kc[0] ? 0 : 1;
                const int kcol = (s == 0 ? c : c - p.kc[0]) * BK + 4 * j;
                const bool col_ok = kcol < (s ? p.k[1] : p.k[0]);
                const float* ap = (s ? p.a[1] : p.a[0]) + kcol;
                const int64_t lda = (s ? p.lda[1] : p.lda[0]);
                float4 v[RPT];
#pragma unroll
                for (int i = 0; i < RPT; ++i) {
                    const int64_t r = row0 + rr + 32 * i;
                    v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (col_ok && r < p.m) v[i] = __ldg(reinterpret_cast<const float4*>(ap + r * lda));
                }
                // pull this group's next chunk towards L2 while this one is converted
                if (c + 2 < p.kct) {
                    const int c2 = c + 2;
                    const int s2 = c2 < p.kc[0] ? 0 : 1;
                    const int kcol2 = (s2 == 0 ? c2 : c2 - p.kc[0]) * BK + 4 * j;
                    if (kcol2 < (s2 ? p.k[1] : p.k[0])) {
#pragma unroll
                        for (int i = 0; i < RPT; ++i) {
                            const int64_t r = row0 + rr + 32 * i;
                            if (r < p.m) prefetch_l2((s2 ? p.a[1] : p.a[0]) + kcol2 + r * (s2 ? p.lda[1] : p.lda[0]));
                        }
                    }
                }
                mbar_wait(smem_u32(empty + slot), ((seq / NS) & 1u) ^ 1u);
                uint8_t* dst = stages + slot * STAGE_BYTES + soff;
#pragma unroll
                for (int i = 0; i < RPT; ++i) {
                    float4 hi, lo;
                    tf32_split4(v[i], hi, lo);
                    *reinterpret_cast<float4*>(dst + i * 2048) = hi;
                    *reinterpret_cast<float4*>(dst + A_IMG + i * 2048) = lo;
                }
                fence_async_smem();
                mbar_arrive(smem_u32(a_full + slot));
            }
        }
    } else {
        // ================= epilogue: TMEM -> padded smem transpose -> full-line global stores =====
        // MH = 2: warp (q, h) drains lane quarter q of M half h, all column chunks.
        // MH = 1: warps w and w+4 share lane quarter q and take alternate 32-column chunks.
        const int q = warp & 3, h = warp >> 2;
        float* st = s_stage + warp * STAGING_WORDS;
        uint32_t item_it = 0;
        for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x, ++item_it) {
            const int64_t rt = item / p.n_tiles;
            const int nt = (int)(item - rt * p.n_tiles);
            int ncols = p.n - nt * BN;
            ncols = ncols > BN ? BN : ncols;
            const uint32_t buf = item_it % NBUF;
            const int64_t row_base = rt * BM + (MH == 2 ? h * 128 : 0) + q * 32;
            mbar_wait(smem_u32(acc_full + buf), (item_it / NBUF) & 1u);
            tc_fence_after();
            const uint32_t t_addr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(MH == 2 ? h : buf) * BN;
            if (row_base < p.m) {
#pragma unroll 1
                for (int c0 = (MH == 2 ? 0 : h * 32); c0 < ncols; c0 += (MH == 2 ? 32 : 64)) {
                    float v[32];
                    tmem_ld32(t_addr + (uint32_t)c0, v);
                    tmem_ld_wait();
                    __syncwarp();
#pragma unroll
                    for (int t = 0; t < 32; ++t) st[lane * 33 + t] = v[t];
                    __syncwarp();
                    const int col = nt * BN + c0 + lane;
                    const bool col_ok = c0 + lane < ncols;
                    const float b = (p.bias && col_ok) ? __ldg(p.bias + col) : 0.f;
                    float* op = p.out + row_base * p.ldo + col;
                    const int nrow = (int)(p.m - row_base < 32 ? p.m - row_base : 32);
                    if (col_ok) {
                        if (p.row_div) {
                            // MeanAggregation's backward wants g_agg[m, :] / max(deg(m), 1): the quotient of the value
                            // this epilogue holds, so the K2a pass over the matrix is saved (same bits: one
                            // round-to-nearest sum, one IEEE division)
#pragma unroll 4
                            for (int i = 0; i < nrow; ++i)
                                op[(int64_t)i * p.ldo] = __fdiv_rn(st[i * 33 + lane] + b, __ldg(p.row_div + row_base + i));
                        } else if (!p.accumulate) {
#pragma unroll 4
                            for (int i = 0; i < nrow; ++i) op[(int64_t)i * p.ldo] = st[i * 33 + lane] + b;
                        } else {
                            // 16 reads of the chunk in flight before the first add: as a per-row
                            // read-modify-write chain (4 in flight) an accumulating call cost 3x a
                            // plain one (ogbn shapes never chain; the 768-wide mix layers do)
#pragma unroll 1
                            for (int i0 = 0; i0 < 32; i0 += 16) {
                                float old[16];
#pragma unroll
                                for (int i = 0; i < 16; ++i) old[i] = i0 + i < nrow ? op[(int64_t)(i0 + i) * p.ldo] : 0.f;
#pragma unroll
                                for (int i = 0; i < 16; ++i)
                                    if (i0 + i < nrow) op[(int64_t)(i0 + i) * p.ldo] = old[i] + (st[(i0 + i) * 33 + lane] + b);
                            }
                        }
                    }
                }
            }
            tc_fence_before();
            mbar_arrive(smem_u32(acc_empty + buf));
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == MMA_WARP) tmem_dealloc(tmem_base, 512);
}

}  // namespace lin
}  // namespace stemgnn

using namespace stemgnn;

static bool lin_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

static int lin_validate(const stemgnn_linear_args* a) {
    SG_CHECK_ARG(a != nullptr, "args is NULL");
    SG_CHECK_ARG(a->m >= 0 && a->n >= 0, "negative size");
    SG_CHECK_ARG(a->n_src == 1 || a->n_src == 2, "n_src must be 1 or 2");
    SG_CHECK_ARG(a->n < (1 << 24), "n too large");
    for (int s = 0; s < a->n_src; ++s) {
        SG_CHECK_ARG(a->k[s] > 0 && a->k[s] < (1 << 24), "k out of range");
        SG_CHECK_ARG(a->k[s] % 4 == 0 && a->lda[s] % 4 == 0 && a->lda[s] >= a->k[s],
                     "k and lda must be multiples of 4 floats (16-byte row segments)");
    }
    return STEMGNN_OK;
}

static void lin_chunks(const stemgnn_linear_args* a, int kc[2], int& kct, int& n_tiles) {
    kc[0] = (a->k[0] + lin::BK - 1) / lin::BK;
    kc[1] = a->n_src == 2 ? (a->k[1] + lin::BK - 1) / lin::BK : 0;
    kct = kc[0] + kc[1];
    n_tiles = (int)((a->n + lin::BN - 1) / lin::BN);
}

extern "C" int stemgnn_linear_workspace(const stemgnn_linear_args* a, size_t* bytes) {
    SG_CHECK_ARG(bytes != nullptr, "bytes is NULL");
    if (int rc = lin_validate(a)) return rc;
    int kc[2], kct, n_tiles;
    lin_chunks(a, kc, kct, n_tiles);
    *bytes = (size_t)n_tiles * kct * lin::W_STAGE_BYTES + 256;
    return STEMGNN_OK;
}

extern "C" int stemgnn_linear(const stemgnn_linear_args* a, void* workspace, size_t workspace_bytes,
                              stemgnn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (int rc = lin_validate(a)) return rc;
    if (a->m == 0 || a->n == 0) return STEMGNN_OK;
    int kc[2], kct, n_tiles;
    lin_chunks(a, kc, kct, n_tiles);
    const size_t need = (size_t)n_tiles * kct * lin::W_STAGE_BYTES + 256;
    SG_CHECK_ARG(workspace && workspace_bytes >= need, "workspace too small (stemgnn_linear_workspace)");
    SG_CHECK_ARG(a->out && a->ldo >= a->n, "out is NULL or ldo < n");
    SG_CHECK_ARG(!(a->row_div && a->accumulate), "row_div divides the finished product: not with accumulate = 1");
    for (int s = 0; s < a->n_src; ++s) {
        SG_CHECK_ARG(a->a[s] && a->w[s], "a/w is NULL");
        SG_CHECK_ARG(lin_aligned16(a->a[s]), "a must be 16-byte aligned");
        SG_CHECK_ARG(a->ldw[s] >= (a->w_transposed ? a->n : a->k[s]), "ldw too small");
    }
    float* img = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~(uintptr_t)255);

    lin::WPrep w;
    lin::LinParams p;
    for (int s = 0; s < 2; ++s) {
        const bool on = s < a->n_src;
        w.w[s] = on ? a->w[s] : nullptr; w.ldw[s] = on ? a->ldw[s] : 0; w.k[s] = on ? a->k[s] : 0; w.kc[s] = kc[s];
        p.a[s] = on ? a->a[s] : nullptr; p.lda[s] = on ? a->lda[s] : 0; p.k[s] = on ? a->k[s] : 0; p.kc[s] = kc[s];
    }
    w.kct = kct; w.n = (int)a->n; w.n_tiles = n_tiles; w.transposed = a->w_transposed; w.img = img;
    const int64_t units = (int64_t)n_tiles * kct * 1024;
    SG_LAUNCH(lin::lin_w_image_kernel<<<(unsigned)((units + 255) / 256), 256, 0, stream>>>(w));
    SG_CHECK_LAUNCH();

    p.kct = kct; p.m = a->m; p.n = (int)a->n; p.n_tiles = n_tiles; p.wimg = img; p.bias = a->bias;
    p.out = a->out; p.ldo = a->ldo; p.accumulate = a->accumulate != 0;
    p.row_div = a->row_div;
    // Tile shape (measured, scripts/lin_bench.py, M = 169,343): 128-row tiles with the double-buffered
    // accumulator win 10-17 % while the split weight image is small (N=256, K<=512: 114 / 196 us
    // against 136 / 218 us); once it reaches a few MB (N=512, K=1024: 4 MB) streaming it twice as
    // often costs more than the exposed epilogue (730 vs 687 us).  STEMGNN_LIN_BM=128|256 forces one.
    static const int forced = [] { const char* e = getenv("STEMGNN_LIN_BM"); return e ? atoi(e) : 0; }();
    // The 256-row shape is ~6 % faster per tile there, but has half as many work items: when those
    // do not fill whole rounds of the grid the tail eats the gain (M = 43,008, N = 768: 504 items =
    // 3.4 rounds of 148 CTAs -> 557 us against 455 us with 1,008 items of 128 rows).
    const int sms = persistent_sms();
    auto fill = [&](int mhh) {
        const int64_t it = ((a->m + 128 * mhh - 1) / (128 * mhh)) * n_tiles;
        const int64_t rounds = (it + sms - 1) / sms;
        return (double)it / (double)(rounds * sms);
    };
    const bool wide = (size_t)n_tiles * kct * lin::W_STAGE_BYTES >= (2u << 20);
    const int mh = forced == 256 ? 2 : (forced == 128 ? 1 : ((wide && fill(2) >= 0.93 * fill(1)) ? 2 : 1));
    const int smem_bytes = mh == 2 ? lin::Shape<2>::SMEM_BYTES : lin::Shape<1>::SMEM_BYTES;
    cudaError_t e = mh == 2
        ? cudaFuncSetAttribute(lin::lin_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes)
        : cudaFuncSetAttribute(lin::lin_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    if (e != cudaSuccess) {
        set_error("stemgnn_linear: cudaFuncSetAttribute(%d B smem): %s", smem_bytes, cudaGetErrorString(e));
        return STEMGNN_ECUDA;
    }
    const int bm = 128 * mh;
    const int64_t items = ((a->m + bm - 1) / bm) * n_tiles;
    const unsigned grid = (unsigned)(items < sms ? items : sms);
    if (mh == 2) SG_LAUNCH(lin::lin_tc_kernel<2><<<grid, lin::NUM_THREADS, smem_bytes, stream>>>(p));
    else SG_LAUNCH(lin::lin_tc_kernel<1><<<grid, lin::NUM_THREADS, smem_bytes, stream>>>(p));
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}
