// k6_wgrad_tc.cu -- K6: weight gradients of the dense layers on the tcgen05 tensor cores.
//
//   gW_s[n, k] = sum_m G[m, n] * A_s[m, k]        (autograd of lin_l / lin_r, reference
//                                                  STEM-GNN/model/encoder.py:84-88)
//
// The reduction runs over the ROWS (nodes), the slow index of both operands in memory.  tf32 MMAs
// do take such "MN-major" operands (one legal layout, SWIZZLE_128B_BASE32B) but were measured at
// 1/5 of the K-major rate, so the producers TRANSPOSE instead: a thread loads a 4-row x 4-column
// block of G or A_s (four float4), and stores its four columns as four 16-byte units of the
// K-major / 64B-swizzle image (row = feature, K = node), split hi + lo like K5 (3xTF32:
// hi*hi + lo*hi + hi*lo, fp32 accumulate).  Lanes of a quarter-warp take different 4-row groups
// and swap column pairs so that every 128-bit store wavefront is bank-conflict free.
//
// The tensor-core accumulator truncates on every accumulate (measured: -2.7e-8 relative per MMA),
// so a long reduction is cut into SLABS of 512 rows: a work item is (slab, n block, k block), its
// 256 x 256 fp32 tile is accumulated in TMEM (two M=128 halves) and then added, round-to-nearest,
// to the CTA's private partial tile in global memory (L2-resident).  A second kernel sums the
// per-CTA partials in CTA order: deterministic, no atomics.
// FLOPs: 3 * 2*M*N*K (tf32); HBM: M*N*4 per k block + M*K*4 per n block.
#include "common.cuh"
#include "tc_ptx.cuh"

namespace stemgnn {
namespace wgrad {

using namespace tcptx;

constexpr int BT = 256;                   // tile edge: n block (two UMMA M=128 halves) and k block (UMMA N)
constexpr int SLAB = 512;                 // rows per work item
constexpr int KM = 16;                    // rows per stage (two K=8 MMA steps)
constexpr int IMG_BYTES = KM * BT * 4;    // 16 KB: one operand image of a stage
constexpr int STAGE_BYTES = 4 * IMG_BYTES;    // G hi, G lo, A hi, A lo
constexpr int NS = 3;
constexpr int EPI_WARPS = 8, PROD_WARP0 = 8, MMA_WARP = 0;
// warps 0-7 epilogue, 8-15 producers.  Lane 0 of warp 0 also issues the MMAs of an item before the
// warp joins its epilogue: the accumulators are single-buffered, so the two never overlap, and 512
// threads leave 128 registers per thread (544 spilled, and local-memory loads behind the streaming
// global loads were the measured bottleneck).
constexpr int NUM_THREADS = 512;
constexpr int STAGING_WORDS = 32 * 33;
constexpr int SMEM_BYTES = NS * STAGE_BYTES + EPI_WARPS * STAGING_WORDS * 4 + 16 * 8 + 16;
constexpr int MAX_COMBOS = 64;

struct WgParams {
    const float* g; int64_t ldg;
    const float* a[2]; int64_t lda[2];
    int k[2];
    int64_t m;
    int n, nb, kbt;          // n blocks, k blocks over the CONCATENATED sources [A_0 | A_1]
    int slots;               // partial tiles per CTA (the grid is a multiple of nb*kbt: one tile id per CTA)
    int64_t slabs;
    int slab;                // rows per work item (<= SLAB, multiple of KM)
    float* partial;          // [grid][slots][256*256]: local item i of a CTA lands in slot i % slots
};

// byte offset of the 16 B unit `ku` (four consecutive nodes) of feature row c in a 256-row
// K-major SWIZZLE_64B image
__device__ __forceinline__ uint32_t sw64_off(int c, int ku) {
    return (uint32_t)((c >> 3) * 512 + (c & 7) * 64 + ((ku ^ ((c >> 1) & 3)) << 4));
}

__device__ __forceinline__ void split_store(uint8_t* hi_p, uint8_t* lo_p, float x, float y, float z, float w) {
    float4 hi, lo;
    tf32_split4(make_float4(x, y, z, w), hi, lo);
    *reinterpret_cast<float4*>(hi_p) = hi;
    *reinterpret_cast<float4*>(lo_p) = lo;
}

// Transposed store of a 4-node x 4-feature block (v[i] = the 4 features of node 4*mg + i) into the
// hi / lo images at `img`: feature column j goes to row c0 + j, K unit mg.  `flip` swaps the column
// pairs (0,1) and (2,3) in time so the two halves of a quarter-warp hit opposite row parities.
__device__ __forceinline__ void store_block_t(uint8_t* img, const float4 (&v)[4], int c0, int mg, bool flip) {
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        // column handled at step t: t ^ flip
        float e[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float lo2 = (t & 2) ? v[i].z : v[i].x, hi2 = (t & 2) ? v[i].w : v[i].y;
            e[i] = ((t & 1) != (int)flip) ? hi2 : lo2;
        }
        const int c = c0 + ((t & 2) | ((t & 1) ^ (int)flip));
        const uint32_t off = sw64_off(c, mg);
        split_store(img + off, img + IMG_BYTES + off, e[0], e[1], e[2], e[3]);
    }
}

__global__ void __launch_bounds__(NUM_THREADS, 1) wgrad_tc_kernel(const WgParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* stages = smem;
    float* s_stage = reinterpret_cast<float*>(smem + NS * STAGE_BYTES);
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_stage + EPI_WARPS * STAGING_WORDS);
    uint64_t* full = bars;                // [NS] 128 producer arrivals
    uint64_t* empty = bars + NS;          // [NS] tcgen05.commit
    uint64_t* acc_full = bars + 2 * NS;
    uint64_t* acc_empty = acc_full + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int combos = p.nb * p.kbt;
    const int64_t n_items = p.slabs * combos;

    if (threadIdx.x == 0) {
        for (int i = 0; i < NS; ++i) {
            mbar_init(smem_u32(full + i), 128);
            mbar_init(smem_u32(empty + i), 1);
        }
        mbar_init(smem_u32(acc_full), 1);
        mbar_init(smem_u32(acc_empty), EPI_WARPS * 32);
        fence_barrier_init();
    }
    if (warp == MMA_WARP) tmem_alloc(smem_u32(tmem_slot), 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp >= PROD_WARP0) {
        // ================= producers: 16 nodes x (256 G features + 256 A features) per stage =====
        // Two groups of 128 threads take alternate stages.  Lane (mg, uo) of warp w owns the 4-node
        // group mg and the 16 B column units 16 w + uo and 16 w + uo + 8 of both operands: one load
        // instruction of a warp reads four full 128-byte lines.
        const int pt = (threadIdx.x - PROD_WARP0 * 32) & 127;
        const int grp = (threadIdx.x - PROD_WARP0 * 32) >> 7;
        const int mg = pt & 3, uo = (pt >> 2) & 7, wq = pt >> 5;
        const bool flip = uo & 1;
        uint32_t seq0 = 0;
        for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int64_t slab = item / combos;
            const int combo = (int)(item - slab * combos);
            const int nbi = combo / p.kbt, kbi = combo - nbi * p.kbt;
            const int kblk = kbi * BT, nblk = nbi * BT;      // kblk: column of the concatenated [A_0 | A_1]
            const int64_t m0 = slab * p.slab;
            int64_t rows = p.m - m0;
            rows = rows > p.slab ? p.slab : rows;
            const int n_st = (int)((rows + KM - 1) / KM);
            const int64_t ldg = p.ldg;
            const float* gbase = p.g + nblk;
            // this thread's two column units of the k block: which source, which column
            const float* acol[2];
            int64_t lda2[2];
            bool a_ok2[2];
#pragma unroll
            for (int b = 0; b < 2; ++b) {
                const int kc = kblk + 4 * (16 * wq + uo + 8 * b);
                const int s = kc < p.k[0] ? 0 : 1;
                const int cs = s == 0 ? kc : kc - p.k[0];
                a_ok2[b] = cs < (s ? p.k[1] : p.k[0]);
                acol[b] = a_ok2[b] ? (s ? p.a[1] : p.a[0]) + cs : nullptr;
                lda2[b] = (s ? p.lda[1] : p.lda[0]);
            }
#pragma unroll 1
            for (int c = grp; c < n_st; c += 2) {
                const uint32_t seq = seq0 + (uint32_t)c;
                const uint32_t slot = seq % NS;
                float4 vg[2][4], va[2][4];
#pragma unroll
                for (int b = 0; b < 2; ++b) {
                    const int col = 4 * (16 * wq + uo + 8 * b);
                    const bool g_ok = nblk + col < p.n, a_ok = a_ok2[b];
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int64_t r = m0 + c * KM + 4 * mg + i;
                        vg[b][i] = make_float4(0.f, 0.f, 0.f, 0.f);
                        va[b][i] = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (r < m0 + rows) {
                            if (g_ok) vg[b][i] = __ldg(reinterpret_cast<const float4*>(gbase + r * ldg + col));
                            if (a_ok) va[b][i] = __ldg(reinterpret_cast<const float4*>(acol[b] + r * lda2[b]));
                        }
                    }
                }
                if (c + 2 < n_st) {      // pull this group's next stage towards L2
#pragma unroll
                    for (int b = 0; b < 2; ++b) {
                        const int col = 4 * (16 * wq + uo + 8 * b);
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const int64_t r = m0 + (c + 2) * KM + 4 * mg + i;
                            if (r < m0 + rows) {
                                if (nblk + col < p.n) prefetch_l2(gbase + r * ldg + col);
                                if (a_ok2[b]) prefetch_l2(acol[b] + r * lda2[b]);
                            }
                        }
                    }
                }
                mbar_wait(smem_u32(empty + slot), ((seq / NS) & 1u) ^ 1u);
                uint8_t* st = stages + slot * STAGE_BYTES;
#pragma unroll
                for (int b = 0; b < 2; ++b) {
                    const int col = 4 * (16 * wq + uo + 8 * b);
                    store_block_t(st, vg[b], col, mg, flip);
                    store_block_t(st + 2 * IMG_BYTES, va[b], col, mg, flip);
                }
                fence_async_smem();
                mbar_arrive(smem_u32(full + slot));
            }
            seq0 += (uint32_t)n_st;
        }
    } else {
        // ================= epilogue: partial[cta][combo] (+)= accumulator =================
        const int q = warp & 3, h = warp >> 2;
        float* st = s_stage + warp * STAGING_WORDS;
        const uint32_t idesc = make_idesc_tf32(128, BT, 0, 0);
        uint32_t item_it = 0, mma_seq = 0;
        for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x, ++item_it) {
            const bool first = item_it < (uint32_t)p.slots;
            float* tile = p.partial + ((int64_t)blockIdx.x * p.slots + item_it % (uint32_t)p.slots) * (BT * BT);
            const int row_base = h * 128 + q * 32;
            if (!first) {       // the tile is read back below: pull its lines towards L2 while the MMAs run
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(tile + (int64_t)(row_base + lane) * BT + i * 32));
            }
            if (warp == MMA_WARP) {
                {
                    // ---- MMA issue for this item (the previous epilogue has drained TMEM): the
                    // whole warp walks the loop, one elected lane issues (uniform registers, no
                    // per-instruction predicate loops)
                    const int64_t slab = item / combos;
                    int64_t rows = p.m - slab * p.slab;
                    rows = rows > p.slab ? p.slab : rows;
                    const int n_st = (int)((rows + KM - 1) / KM);
                    mbar_wait(smem_u32(acc_empty), (item_it & 1u) ^ 1u);
                    tc_fence_after();
                    for (int c = 0; c < n_st; ++c, ++mma_seq) {
                        const uint32_t slot = mma_seq % NS;
                        mbar_wait(smem_u32(full + slot), (mma_seq / NS) & 1u);
                        tc_fence_after();
                        const uint32_t base = smem_u32(stages + slot * STAGE_BYTES);
                        if (elect_one()) {
#pragma unroll
                        for (int hh = 0; hh < 2; ++hh) {
#pragma unroll
                            for (int ks = 0; ks < 2; ++ks) {
                                const uint32_t go = base + hh * (IMG_BYTES / 2) + ks * 32;
                                const uint64_t g_hi = smem_desc_sw64(go);
                                const uint64_t g_lo = smem_desc_sw64(go + IMG_BYTES);
                                const uint64_t a_hi = smem_desc_sw64(base + 2 * IMG_BYTES + ks * 32);
                                const uint64_t a_lo = smem_desc_sw64(base + 3 * IMG_BYTES + ks * 32);
                                const uint32_t d = tmem_base + (uint32_t)hh * BT;
                                umma_tf32(d, g_lo, a_hi, idesc, (c > 0 || ks > 0) ? 1u : 0u);
                                umma_tf32(d, g_hi, a_lo, idesc, 1u);
                                umma_tf32(d, g_hi, a_hi, idesc, 1u);
                            }
                        }
                        umma_commit(smem_u32(empty + slot));
                        if (c == n_st - 1) umma_commit(smem_u32(acc_full));
                        }
                        __syncwarp();
                    }
                }
            }
            mbar_wait(smem_u32(acc_full), item_it & 1u);
            tc_fence_after();
            const uint32_t t_addr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)h * BT;
#pragma unroll 1
            for (int c0 = 0; c0 < BT; c0 += 32) {
                float v[32];
                tmem_ld32(t_addr + (uint32_t)c0, v);
                tmem_ld_wait();
                __syncwarp();
#pragma unroll
                for (int t = 0; t < 32; ++t) st[lane * 33 + t] = v[t];
                __syncwarp();
                float* op = tile + (int64_t)row_base * BT + c0 + lane;
                if (first) {
#pragma unroll 4
                    for (int i = 0; i < 32; ++i) op[i * BT] = st[i * 33 + lane];
                } else {
                    float o[32];
#pragma unroll
                    for (int i = 0; i < 32; ++i) o[i] = op[i * BT];
#pragma unroll
                    for (int i = 0; i < 32; ++i) op[i * BT] = o[i] + st[i * 33 + lane];
                }
            }
            tc_fence_before();
            mbar_arrive(smem_u32(acc_empty));
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == MMA_WARP) tmem_dealloc(tmem_base, 512);
}

// gw_s[n, k] = sum, in (slot, CTA) order, of the partial tiles of that (n block, k block).  Slot i of
// CTA c holds the items c + (i + t*slots)*grid, t = 0, 1, ..: all the same combo because the grid is
// a multiple of the combo count.
struct WgReduce {
    const float* partial;
    int grid, nb, kbt, slots;
    int64_t n_items;
    int n, k[2];
    float* gw[2]; int64_t ldgw[2];
};

__global__ void __launch_bounds__(256) wgrad_reduce_kernel(const WgReduce p) {
    // A block owns 32 float4 (128 consecutive elements of one tile); its 8 warps split the list of
    // partial tiles (warp w takes every 8th), each summing in list order with several 512-byte loads
    // in flight; the 8 warp sums are added in warp order.  Fixed order: deterministic.
    __shared__ float4 sh[8][32];
    const int combos = p.nb * p.kbt;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t blocks = (int64_t)combos * (BT * BT / 128);
    for (int64_t blk = blockIdx.x; blk < blocks; blk += gridDim.x) {
        const int combo = (int)(blk / (BT * BT / 128));
        const int e = ((int)(blk - (int64_t)combo * (BT * BT / 128)) * 32 + lane) * 4;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        int idx = 0;      // running index over this combo's partial tiles
        for (int sl = 0; sl < p.slots; ++sl) {
            const int64_t first_item = (int64_t)sl * p.grid;
            if (first_item >= p.n_items) break;
            const int base = (int)(first_item % combos);
            const int c_end = (int)(p.n_items - first_item < p.grid ? p.n_items - first_item : p.grid);
            const float* src = p.partial + (int64_t)sl * (BT * BT) + e;
            const int64_t cstride = (int64_t)p.slots * (BT * BT);
            const int c_first = (combo - base + combos) % combos;
            const int cnt = c_first < c_end ? (c_end - c_first + combos - 1) / combos : 0;
            // tiles idx .. idx+cnt-1 of the list; warp w takes list positions == w (mod 8)
            int t = ((w - idx) % 8 + 8) % 8;
#pragma unroll 4
            for (; t < cnt; t += 8) {
                const float4 v = __ldg(reinterpret_cast<const float4*>(src + (int64_t)(c_first + t * combos) * cstride));
                acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
            }
            idx += cnt;
        }
        sh[w][lane] = acc;
        __syncthreads();
        if (w == 0) {
            float4 r = sh[0][lane];
#pragma unroll
            for (int i = 1; i < 8; ++i) { const float4 v = sh[i][lane]; r.x += v.x; r.y += v.y; r.z += v.z; r.w += v.w; }
            const int nbi = combo / p.kbt, kbi = combo - nbi * p.kbt;
            const int row = nbi * BT + e / BT, kc = kbi * BT + (e % BT);
            const int s = kc < p.k[0] ? 0 : 1;
            const int col = s == 0 ? kc : kc - p.k[0];
            if (row < p.n && col < (s ? p.k[1] : p.k[0])) {      // k % 4 == 0: the 4 columns are in range together
                float* o = (s ? p.gw[1] : p.gw[0]) + (int64_t)row * (s ? p.ldgw[1] : p.ldgw[0]) + col;
                o[0] = r.x; o[1] = r.y; o[2] = r.z; o[3] = r.w;
            }
        }
        __syncthreads();
    }
}

// column sums of g (the bias gradient): per-block partial sums over a row range, then a fixed-order
// sum of the block partials
__global__ void __launch_bounds__(256) colsum_partial_kernel(const float* __restrict__ g, int64_t ldg, int64_t m, int n,
                                                             int64_t rows_per_block, float* __restrict__ part) {
    // thread (cu, rl): 16 B column unit cu + 64 t, rows r0 + rl, r0 + rl + 4, ...; the four row lanes of a
    // unit are summed through shared memory in a fixed order
    __shared__ float4 sh[4][64];
    const int cu = threadIdx.x & 63, rl = threadIdx.x >> 6;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
    const int64_t r1 = r0 + rows_per_block < m ? r0 + rows_per_block : m;
    for (int c0 = 0; c0 < n; c0 += 256) {
        const int c = c0 + 4 * cu;
        float4 s0 = make_float4(0.f, 0.f, 0.f, 0.f), s1 = s0;
        if (c < n) {
            int64_t r = r0 + rl;
            for (; r + 28 < r1; r += 32) {          // 8 independent 16 B loads in flight per thread
                float4 v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) v[u] = __ldg(reinterpret_cast<const float4*>(g + (r + 4 * u) * ldg + c));
#pragma unroll
                for (int u = 0; u < 8; u += 2) {
                    s0.x += v[u].x; s0.y += v[u].y; s0.z += v[u].z; s0.w += v[u].w;
                    s1.x += v[u + 1].x; s1.y += v[u + 1].y; s1.z += v[u + 1].z; s1.w += v[u + 1].w;
                }
            }
            for (; r < r1; r += 4) {
                const float4 a = __ldg(reinterpret_cast<const float4*>(g + r * ldg + c));
                s0.x += a.x; s0.y += a.y; s0.z += a.z; s0.w += a.w;
            }
        }
        sh[rl][cu] = make_float4(s0.x + s1.x, s0.y + s1.y, s0.z + s1.z, s0.w + s1.w);
        __syncthreads();
        if (rl == 0 && c < n) {
            const float4 a = sh[0][cu], b = sh[1][cu], d = sh[2][cu], e = sh[3][cu];
            float* o = part + (int64_t)blockIdx.x * n + c;
            o[0] = (a.x + b.x) + (d.x + e.x); o[1] = (a.y + b.y) + (d.y + e.y);
            o[2] = (a.z + b.z) + (d.z + e.z); o[3] = (a.w + b.w) + (d.w + e.w);
        }
        __syncthreads();
    }
}
constexpr int CSF_WARPS = 32;
__global__ void __launch_bounds__(CSF_WARPS * 32) colsum_final_kernel(const float* __restrict__ part, int blocks, int n,
                                                                      float* __restrict__ out) {
    // one block per 32 columns; the 32 warps take every 32nd partial row (coalesced 128 B reads, all of a warp's
    // loads in flight at once), then their sums are added in warp order
    __shared__ float sh[CSF_WARPS][32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + lane;
    float s0 = 0.f, s1 = 0.f;
    if (c < n) {
        int b = w;
#pragma unroll 8
        for (; b + CSF_WARPS < blocks; b += 2 * CSF_WARPS) { s0 += part[(int64_t)b * n + c]; s1 += part[(int64_t)(b + CSF_WARPS) * n + c]; }
        if (b < blocks) s0 += part[(int64_t)b * n + c];
    }
    sh[w][lane] = s0 + s1;
    __syncthreads();
    if (w == 0 && c < n) {
        float r = sh[0][lane];
#pragma unroll
        for (int i = 1; i < CSF_WARPS; ++i) r += sh[i][lane];
        out[c] = r;
    }
}

}  // namespace wgrad
}  // namespace stemgnn

using namespace stemgnn;

namespace {
struct WgPlan { int nb, kbt, combos, grid, slots, slab, cs_blocks; int64_t slabs, items; size_t partial_bytes, cs_bytes; };

int wg_plan(const stemgnn_linear_wgrad_args* a, WgPlan& pl) {
    SG_CHECK_ARG(a != nullptr, "args is NULL");
    SG_CHECK_ARG(a->m >= 0 && a->n >= 0, "negative size");
    SG_CHECK_ARG(a->n_src == 1 || a->n_src == 2, "n_src must be 1 or 2");
    SG_CHECK_ARG(a->n % 4 == 0 && a->ldg % 4 == 0 && a->ldg >= a->n, "n and ldg must be multiples of 4 floats");
    int64_t kcat = 0;
    for (int s = 0; s < a->n_src; ++s) {
        SG_CHECK_ARG(a->k[s] > 0 && a->k[s] % 4 == 0 && a->lda[s] % 4 == 0 && a->lda[s] >= a->k[s],
                     "k and lda must be positive multiples of 4 floats");
        kcat += a->k[s];
    }
    pl.nb = (int)((a->n + wgrad::BT - 1) / wgrad::BT);
    SG_CHECK_ARG(kcat < (1 << 24) && a->n < (1 << 24), "layer too wide");
    pl.kbt = (int)((kcat + wgrad::BT - 1) / wgrad::BT);
    pl.combos = pl.nb * pl.kbt;
    SG_CHECK_ARG(pl.combos <= wgrad::MAX_COMBOS, "weight matrix too large: more than 64 tiles of 256 x 256");
    // Rows per item: at most SLAB (accuracy), shrunk so that the items fill whole rounds of the grid
    // (169,343 rows: 331 slabs of 512 x 2 tiles = 4.47 rounds of 148 CTAs -> 365 slabs of 464 = 4.93)
    const int sms = persistent_sms();
    {
        const int64_t slabs0 = (a->m + wgrad::SLAB - 1) / wgrad::SLAB;
        const int64_t geff = (int64_t)(sms / pl.combos) * pl.combos;      // the grid (see below)
        const int64_t rounds = (slabs0 * pl.combos + geff - 1) / geff;
        int64_t slabs_t = rounds * geff / pl.combos;
        if (slabs_t < slabs0) slabs_t = slabs0;
        int64_t slab = a->m > 0 ? (a->m + slabs_t - 1) / slabs_t : wgrad::SLAB;
        slab = (slab + wgrad::KM - 1) / wgrad::KM * wgrad::KM;
        if (slab > wgrad::SLAB) slab = wgrad::SLAB;
        if (slab < wgrad::KM) slab = wgrad::KM;
        pl.slab = (int)slab;
    }
    pl.slabs = (a->m + pl.slab - 1) / pl.slab;
    pl.items = pl.slabs * pl.combos;
    // The grid is a multiple of the tile count, so CTA c only ever sees tile c % combos (items are
    // dealt round-robin, slab-major: the CTAs of one round share their G / A slabs through L2) and any
    // number of partial slots per CTA is legal.
    pl.grid = (int)(pl.items < sms ? (pl.items > 0 ? pl.items : 1) : (sms / pl.combos) * pl.combos);
    // One partial tile per local item while all of them stay under 256 MB (written once, read once by
    // the reduce).  Beyond that (wide layers: 3072 x 768 weights over 43k rows were 3,024 tiles =
    // 790 MB written and read back, the reduce alone 430 us) the CTA keeps as few tiles as fit in L2
    // next to the operands and adds every further item into them (read-modify-write through L2).
    const int64_t per_cta = (pl.items + pl.grid - 1) / pl.grid;
    const int64_t tile_bytes = (int64_t)wgrad::BT * wgrad::BT * 4;
    int64_t slots = per_cta;
    if ((int64_t)pl.grid * per_cta * tile_bytes > (256ll << 20)) slots = (64ll << 20) / ((int64_t)pl.grid * tile_bytes);
    if (slots < 1) slots = 1;
    pl.slots = (int)slots;
    pl.partial_bytes = (size_t)pl.grid * pl.slots * wgrad::BT * wgrad::BT * 4;
    pl.cs_blocks = sms * 4;
    pl.cs_bytes = a->gbias ? (size_t)pl.cs_blocks * a->n * 4 : 0;
    return STEMGNN_OK;
}
}  // namespace

extern "C" int stemgnn_linear_wgrad_workspace(const stemgnn_linear_wgrad_args* a, size_t* bytes) {
    SG_CHECK_ARG(bytes != nullptr, "bytes is NULL");
    WgPlan pl;
    if (int rc = wg_plan(a, pl)) return rc;
    *bytes = 256 + pl.partial_bytes + pl.cs_bytes;
    return STEMGNN_OK;
}

extern "C" int stemgnn_linear_wgrad(const stemgnn_linear_wgrad_args* a, void* workspace, size_t workspace_bytes,
                                    stemgnn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    WgPlan pl;
    if (int rc = wg_plan(a, pl)) return rc;
    if (a->n == 0) return STEMGNN_OK;
    SG_CHECK_ARG(workspace && workspace_bytes >= 256 + pl.partial_bytes + pl.cs_bytes,
                 "workspace too small (stemgnn_linear_wgrad_workspace)");
    SG_CHECK_ARG(a->m == 0 || (a->g && (reinterpret_cast<uintptr_t>(a->g) & 15) == 0), "g is NULL or not 16-byte aligned");
    for (int s = 0; s < a->n_src; ++s) {
        SG_CHECK_ARG(a->gw[s] && a->ldgw[s] >= a->k[s], "gw is NULL or ldgw < k");
        SG_CHECK_ARG(a->m == 0 || (a->a[s] && (reinterpret_cast<uintptr_t>(a->a[s]) & 15) == 0), "a is NULL or not 16-byte aligned");
    }
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~(uintptr_t)255);
    float* partial = reinterpret_cast<float*>(base);
    float* cs_part = reinterpret_cast<float*>(base + pl.partial_bytes);

    if (a->m == 0) {      // empty reduction: zero gradients
        for (int s = 0; s < a->n_src; ++s)
            SG_CHECK_CUDA(cudaMemset2DAsync(a->gw[s], a->ldgw[s] * 4, 0, a->k[s] * 4, a->n, stream));
        if (a->gbias) SG_CHECK_CUDA(cudaMemsetAsync(a->gbias, 0, a->n * 4, stream));
        return STEMGNN_OK;
    }

    wgrad::WgParams p;
    p.g = a->g; p.ldg = a->ldg;
    for (int s = 0; s < 2; ++s) {
        const bool on = s < a->n_src;
        p.a[s] = on ? a->a[s] : nullptr; p.lda[s] = on ? a->lda[s] : 0; p.k[s] = on ? (int)a->k[s] : 0;
    }
    p.m = a->m; p.n = (int)a->n; p.nb = pl.nb; p.kbt = pl.kbt; p.slabs = pl.slabs; p.slab = pl.slab; p.partial = partial; p.slots = pl.slots;
    cudaError_t e = cudaFuncSetAttribute(wgrad::wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, wgrad::SMEM_BYTES);
    if (e != cudaSuccess) {
        set_error("stemgnn_linear_wgrad: cudaFuncSetAttribute(%d B smem): %s", wgrad::SMEM_BYTES, cudaGetErrorString(e));
        return STEMGNN_ECUDA;
    }
    SG_LAUNCH(wgrad::wgrad_tc_kernel<<<pl.grid, wgrad::NUM_THREADS, wgrad::SMEM_BYTES, stream>>>(p));
    SG_CHECK_LAUNCH();

    wgrad::WgReduce r;
    r.partial = partial; r.grid = pl.grid; r.nb = pl.nb; r.kbt = pl.kbt; r.slots = pl.slots; r.n_items = pl.items;
    r.n = (int)a->n;
    for (int s = 0; s < 2; ++s) {
        const bool on = s < a->n_src;
        r.k[s] = on ? (int)a->k[s] : 0; r.gw[s] = on ? a->gw[s] : nullptr; r.ldgw[s] = on ? a->ldgw[s] : 0;
    }
    const int64_t rblocks = (int64_t)pl.combos * (wgrad::BT * wgrad::BT / 128);
    SG_LAUNCH(wgrad::wgrad_reduce_kernel<<<(unsigned)rblocks, 256, 0, stream>>>(r));
    SG_CHECK_LAUNCH();

    if (a->gbias) {
        const int64_t rpb = (a->m + pl.cs_blocks - 1) / pl.cs_blocks;
        const int blocks = (int)((a->m + rpb - 1) / rpb);
        SG_LAUNCH(wgrad::colsum_partial_kernel<<<blocks, 256, 0, stream>>>(a->g, a->ldg, a->m, (int)a->n, rpb, cs_part));
        SG_LAUNCH(wgrad::colsum_final_kernel<<<(unsigned)((a->n + 31) / 32), wgrad::CSF_WARPS * 32, 0, stream>>>(cs_part, blocks, (int)a->n, a->gbias));
        SG_CHECK_LAUNCH();
    }
    return STEMGNN_OK;
}
