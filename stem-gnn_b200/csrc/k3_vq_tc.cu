// k3_vq_tc.cu -- K3 path 2: the codebook search as a tcgen05 / TMEM GEMM fused with the row
// argmax, for sm_100a.
//
// Replaces dist = einsum('h n d, h c d -> h n c') / -cdist and argmax (reference
// STEM-GNN/model/vq.py:650,652 cosine; :472,474 Euclidean; gumbel_sample :79) without ever
// materialising the N x K score matrix.
//
//   scores~ = S * fp16(X) . fp16(E / S)^T  [-  e2/2 for the Euclidean metric]   (tensor cores)
//
// fp16 operands (11-bit significand, same kind::f16 rate as bf16) keep the error band 8x
// tighter than bf16 would; S is a per-codebook power of two that maps max|e| to [2^13, 2^14).
//
// A persistent CTA owns a 128-row tile of X.  Eight producer warps read the fp32 rows with
// 128-bit loads, round to fp16 and store them K-major / 128B-swizzled in shared memory (the
// canonical UMMA operand layout) -- X is read from HBM exactly once, as fp32.  The codebook is
// pre-converted once per call into the same swizzled tile image in global memory, so one
// elected thread streams it with 1-D bulk async copies (cp.async.bulk -> mbarrier complete_tx)
// through a ring of 32 KB stages.  One elected thread issues tcgen05.mma (M=128, N<=256, K=16,
// f16 -> fp32) into a double-buffered 2 x 256-column TMEM accumulator; eight epilogue warps read
// the accumulator with tcgen05.ld (one row per thread, two warps per lane quarter), add -e2/2 and keep the running best
// score plus every code within a rigorous error band of it.
//
// Parity: with x~ = fp16(x), e~ = S fp16(e/S) the tensor-core dot differs from the exact one by
// at most |x - x~||e| + |x~||e - e~| (+ fp32 accumulation slop); both residual norms are
// MEASURED (producers / prep kernel), not bounded a priori, so the band is rigorous and tight
// (~2^-10 |x||e|).  The fp32 winner is always inside [best~ - band, best~].  Rows whose band
// holds a single code are final; the others get best = -1 and their candidate list, and the
// gather kernel (k3_vq_simt.cu) re-scores those few codes with the reference's fp32 formula,
// first index winning ties.  A row whose list overflows (or with |x| >= 2^15, outside the fp16
// range) gets best = -2 and is re-scanned exactly there.
// FLOPs: 2*N*K*D on the tensor pipe; HBM: N*D*4 B (x) + N*4 B (best).
#include <cuda_fp16.h>

#include <cstdlib>

#include "k3_vq_common.cuh"
#include "tc_ptx.cuh"

namespace stemgnn {
namespace tc {

constexpr int BM = 128;                    // rows per tile = UMMA M
constexpr int BN = 256;                    // codes per accumulator tile = max UMMA N
constexpr int BK = 64;                     // bf16 columns per smem chunk (one 128 B swizzle row)
constexpr int A_CHUNK_BYTES = BM * 128;    // 16 KB
constexpr int B_STAGE_BYTES = BN * 128;    // 32 KB
constexpr int CAP = VQ_CAND_CAP;           // candidates kept per row
constexpr int EPI_WARPS = 8;               // two per TMEM lane quarter: they split the columns
constexpr int NUM_THREADS = 576;           // warps 0-7 epilogue, 8-15 producers, 16 loader, 17 MMA
constexpr int PROD_WARP0 = 8, LOAD_WARP = 16, MMA_WARP = 17;
constexpr int MAX_AS = 20;                 // A chunk slots (resident: kcn * 1|2 buffers; ring: 4)
constexpr int MAX_SB = 6;
constexpr int MERGE_WORDS = 8;             // per-row state handed from the odd-chunk warp to its partner
// ne2 ring [4][256] f32, band ring [4][128] f32, group-1 partials [128] float2, merge [128][12],
// barriers, tmem slot
constexpr int SMEM_EXTRA = 4 * BN * 4 + 4 * BM * 4 + BM * 8 + BM * MERGE_WORDS * 4 + 80 * 8 + 16;
constexpr int SMEM_MAX = 232448;           // 227 KB opt-in limit per CTA

using namespace tcptx;

// ---------------------------------------------------------------------------------- params
struct TcParams {
    const float* x; int64_t x_row_stride, x_head_stride;
    int64_t heads, embed_heads, rows;
    int K, D;
    int n_tiles, kcn;                // code tiles of 256, 64-column chunks
    int sb;                          // B ring stages
    int a_slots;                     // A chunk slots: kcn << a_dbl (resident) or a ring of 4
    int a_dbl;                       // resident mode: 1 = two A buffers (row tiles alternate)
    const __half* img;               // [embed_heads][n_tiles][kcn][256 x 64] swizzled tile images
    const float* ne2;                // [embed_heads][n_tiles*256]: -e2/2 | 0, padded with -inf
    const uint32_t* head_stats;      // [embed_heads][4] float bits: max|e|, max e2, max |e - e~|^2, S
    float c_acc, c2;                 // band = 2(rx*emax + 1.001|x|*remax + c_acc*|x|*emax) + c2*(x2+e2max)
    int32_t* best;                   // [heads*rows]: code, -1 (see cand) or -2 (exact re-scan)
    int32_t* cand;                   // [heads*rows][CAP], -1 padded
};

__device__ __forceinline__ int tile_codes32(const TcParams& p, int n) {
    int rem = p.K - n * BN;
    rem = rem > BN ? BN : rem;
    return (rem + 31) & ~31;
}

// Sum 4 per-lane values across the 4 lanes of a quad so that lane j ends with the total of
// value index j (3 shuffles instead of 8).
__device__ __forceinline__ float transpose_reduce4(const float (&a)[4], int j) {
    const float k0 = (j & 2) ? a[2] : a[0], s0 = (j & 2) ? a[0] : a[2];
    const float k1 = (j & 2) ? a[3] : a[1], s1 = (j & 2) ? a[1] : a[3];
    const float b0 = k0 + __shfl_xor_sync(0xffffffffu, s0, 2);
    const float b1 = k1 + __shfl_xor_sync(0xffffffffu, s1, 2);
    const float keep = (j & 1) ? b1 : b0, send = (j & 1) ? b0 : b1;
    return keep + __shfl_xor_sync(0xffffffffu, send, 1);
}

// ---------------------------------------------------------------------------------- kernel
// Slot and mbarrier phase of an A chunk.  Resident mode fills every chunk once per row tile
// (slots kcn << a_dbl, row tiles alternate between the two buffers); ring mode once per (row tile,
// code tile) through 4 slots, ring_seq counting this CTA's chunks.  No divisions: these run once
// per chunk in the single-thread MMA loop.
template <bool RESIDENT>
__device__ __forceinline__ void a_slot(const TcParams& p, uint32_t tile_it, uint32_t ring_seq, int kc,
                                       uint32_t& slot, uint32_t& phase) {
    if (RESIDENT) {
        slot = (tile_it & (uint32_t)p.a_dbl) * (uint32_t)p.kcn + (uint32_t)kc;
        phase = (tile_it >> p.a_dbl) & 1u;
    } else {
        slot = ring_seq & 3u;
        phase = (ring_seq >> 2) & 1u;
    }
}

template <bool RESIDENT>
__global__ void __launch_bounds__(NUM_THREADS, 1) vq_tc_kernel(const TcParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sA = smem;
    uint8_t* sB = sA + (size_t)p.a_slots * A_CHUNK_BYTES;
    float* s_ne2 = reinterpret_cast<float*>(sB + (size_t)p.sb * B_STAGE_BYTES);   // [4][256]
    float* s_band = s_ne2 + 4 * BN;                                        // [4][128]
    float2* s_part = reinterpret_cast<float2*>(s_band + 4 * BM);                   // [128]
    float* s_merge = reinterpret_cast<float*>(s_part + BM);                        // [128][12]
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_merge + BM * MERGE_WORDS);
    uint64_t* a_full = bars;                 // [MAX_AS]
    uint64_t* a_empty = bars + MAX_AS;       // [MAX_AS]
    uint64_t* b_full = bars + 2 * MAX_AS;    // [MAX_SB]
    uint64_t* b_empty = b_full + MAX_SB;     // [MAX_SB]
    uint64_t* t_full = b_empty + MAX_SB;     // [2]
    uint64_t* t_empty = t_full + 2;          // [2]
    uint64_t* band_full = t_empty + 2;       // [4]
    uint64_t* part_full = band_full + 4;     // [1]
    uint64_t* part_empty = part_full + 1;    // [1]
    uint64_t* merge_full = part_empty + 1;   // [1]
    uint64_t* merge_empty = merge_full + 1;  // [1]
    uint64_t* ne2_full = merge_empty + 1;    // [4]
    uint64_t* ne2_empty = ne2_full + 4;      // [4]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(ne2_empty + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t row_tiles = (p.rows + BM - 1) / BM;
    const int64_t n_items = row_tiles * p.heads;

    if (warp == LOAD_WARP && lane == 0) {
        if (smem_u32(smem) & 1023u) __trap();   // the swizzled operand tiles need 1024 B alignment
        for (int i = 0; i < MAX_AS; ++i) { mbar_init(smem_u32(a_full + i), BM); mbar_init(smem_u32(a_empty + i), 1); }
        for (int i = 0; i < MAX_SB; ++i) { mbar_init(smem_u32(b_full + i), 1); mbar_init(smem_u32(b_empty + i), 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(smem_u32(t_full + i), 1); mbar_init(smem_u32(t_empty + i), EPI_WARPS * 32); }
        for (int i = 0; i < 4; ++i) mbar_init(smem_u32(band_full + i), BM);
        for (int i = 0; i < 4; ++i) { mbar_init(smem_u32(ne2_full + i), 1); mbar_init(smem_u32(ne2_empty + i), EPI_WARPS * 32); }
        mbar_init(smem_u32(part_full), BM);
        mbar_init(smem_u32(part_empty), BM);
        mbar_init(smem_u32(merge_full), BM);
        mbar_init(smem_u32(merge_empty), BM);
        fence_barrier_init();
    }
    if (warp == MMA_WARP) tmem_alloc(smem_u32(tmem_slot), 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == LOAD_WARP) {
        // ================= codebook loader: bulk async copies into the B ring =================
        if (lane == 0) {
            uint32_t s = 0, ph = 0, acc = 0;      // B ring stage / phase carried, not recomputed
            for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x) {
                const int64_t h = item / row_tiles;
                const int64_t eh = p.embed_heads == 1 ? 0 : h;
                for (int n = 0; n < p.n_tiles; ++n, ++acc) {
                    // the tile's -e2/2 slice for the epilogue: a ring of 4, two tiles ahead of TMEM
                    const uint32_t ns = acc & 3u;
                    mbar_wait(smem_u32(ne2_empty + ns), ((acc >> 2) & 1u) ^ 1u);
                    mbar_expect_tx(smem_u32(ne2_full + ns), BN * 4);
                    bulk_g2s(smem_u32(s_ne2 + ns * BN), p.ne2 + (eh * p.n_tiles + n) * (int64_t)BN, BN * 4,
                             smem_u32(ne2_full + ns));
                    const uint32_t bytes = (uint32_t)tile_codes32(p, n) * 128u;
                    const uint8_t* src = reinterpret_cast<const uint8_t*>(p.img) +
                                         ((eh * p.n_tiles + n) * (int64_t)p.kcn) * B_STAGE_BYTES;
                    for (int kc = 0; kc < p.kcn; ++kc) {
                        mbar_wait(smem_u32(b_empty + s), ph ^ 1u);
                        mbar_expect_tx(smem_u32(b_full + s), bytes);
                        bulk_g2s(smem_u32(sB + (size_t)s * B_STAGE_BYTES), src + (size_t)kc * B_STAGE_BYTES,
                                 bytes, smem_u32(b_full + s));
                        if (++s == (uint32_t)p.sb) { s = 0; ph ^= 1u; }
                    }
                }
            }
        }
    } else if (warp == MMA_WARP) {
        // ================= MMA issuer: one thread drives the tensor core =================
        // The whole warp walks the loop -- warp-uniform control flow keeps descriptors and barrier
        // addresses in uniform registers -- and one elected lane issues.  Ring positions are
        // carried, not recomputed.  (The first version ran the loop in lane 0 only, with three
        // integer divisions per chunk and a predicate loop around every UTCHMMA: 430 ns per
        // 64-column chunk with the MMAs taken out, more than the 276 ns the four MMAs need, so the
        // tensor pipe idled 59 % of the time whatever the ring depth or the epilogue did.)
        uint32_t sb = 0, sb_ph = 0, acc_it = 0, tile_it = 0, ring_seq = 0;
        const int last_ksteps = ((p.D - (p.kcn - 1) * BK) + 15) / 16;
        const uint64_t a_desc0 = smem_desc_sw128(smem_u32(sA)), b_desc0 = smem_desc_sw128(smem_u32(sB));
        for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x, ++tile_it) {
            for (int n = 0; n < p.n_tiles; ++n, ++acc_it) {
                const uint32_t idesc = make_idesc(BM, tile_codes32(p, n));
                const uint32_t st = acc_it & 1u;
                mbar_wait(smem_u32(t_empty + st), ((acc_it >> 1) & 1u) ^ 1u);
                const uint32_t d_tmem = tmem_base + st * BN;
                const bool last_n = n == p.n_tiles - 1;
                for (int kc = 0; kc < p.kcn; ++kc, ++ring_seq) {
                    uint32_t sa, a_ph;
                    a_slot<RESIDENT>(p, tile_it, ring_seq, kc, sa, a_ph);
                    if (!RESIDENT || n == 0) mbar_wait(smem_u32(a_full + sa), a_ph);
                    mbar_wait(smem_u32(b_full + sb), sb_ph);
                    tc_fence_after();
                    // the start-address field holds addr >> 4 (14 bits, no carry below 256 KB)
                    const uint64_t a_desc = a_desc0 + (uint64_t)(sa * (uint32_t)(A_CHUNK_BYTES >> 4));
                    const uint64_t b_desc = b_desc0 + (uint64_t)(sb * (uint32_t)(B_STAGE_BYTES >> 4));
                    const int ksteps = kc == p.kcn - 1 ? last_ksteps : BK / 16;
                    if (elect_one()) {
#pragma unroll
                        for (int k = 0; k < BK / 16; ++k)   // +32 B along K inside the swizzle row
                            if (k < ksteps)
                                umma_f16(d_tmem, a_desc + (uint64_t)(2 * k), b_desc + (uint64_t)(2 * k), idesc,
                                         (uint32_t)((kc | k) != 0));
                        umma_commit(smem_u32(b_empty + sb));
                        // the A chunk is free once the last code tile that reads it has consumed it
                        if (!RESIDENT || last_n) umma_commit(smem_u32(a_empty + sa));
                        if (kc == p.kcn - 1) umma_commit(smem_u32(t_full + st));
                    }
                    __syncwarp();
                    if (++sb == (uint32_t)p.sb) { sb = 0; sb_ph ^= 1u; }
                }
            }
        }
    } else if (warp >= PROD_WARP0 && warp < LOAD_WARP) {
        // ================= producers: fp32 rows -> fp16, swizzled K-major A chunks =================
        // Two groups of 128 threads take alternate 64-column chunks, so two chunks (2 x 32 KB of
        // fp32) are in flight per SM.  Thread (j, rr) of a group owns rows rr, rr+32, rr+64, rr+96
        // and the 16 B slots j, j+4, j+8, j+12 of each: a quad reads 64 contiguous bytes per load.
        const int pt = (threadIdx.x - PROD_WARP0 * 32) & 127;
        const int grp = (threadIdx.x - PROD_WARP0 * 32) >> 7;
        const int j = pt & 3, rr = pt >> 2;
        // byte offset of slot f = j + 4k of row rr in the swizzled chunk: 16 B unit (f>>1)^(rr&7), half f&1
        uint32_t soff[4];
#pragma unroll
        for (int k = 0; k < 4; ++k)
            soff[k] = (uint32_t)((rr >> 3) * 1024 + (rr & 7) * 128 + (((((j + 4 * k) >> 1)) ^ (rr & 7)) << 4) + (j & 1) * 8);
        uint32_t tile_it = 0, ring_base = 0;
        for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x, ++tile_it) {
            const int64_t h = item / row_tiles, mt = item - h * row_tiles;
            const int64_t eh = p.embed_heads == 1 ? 0 : h;
            const int64_t row0 = mt * BM;
            const float* xb = p.x + h * p.x_head_stride;
            float sq[4], rq[4];         // |x|^2 and |x - fp16(x)|^2 of this thread's slices of 4 rows
#pragma unroll
            for (int i = 0; i < 4; ++i) { sq[i] = 0.f; rq[i] = 0.f; }
            const int n_pass = RESIDENT ? 1 : p.n_tiles;
#pragma unroll 1
            for (int n = 0; n < n_pass; ++n) {
#pragma unroll 1
                for (int kc = grp; kc < p.kcn; kc += 2) {
                    uint32_t slot, a_ph;
                    a_slot<RESIDENT>(p, tile_it, ring_base + (uint32_t)kc, kc, slot, a_ph);
                    mbar_wait(smem_u32(a_empty + slot), a_ph ^ 1u);
                    float4 v[4][4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int64_t r = row0 + rr + 32 * i;
                        const float* rp = xb + r * p.x_row_stride + kc * BK + 4 * j;
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            v[i][k] = make_float4(0.f, 0.f, 0.f, 0.f);
                            if (r < p.rows && kc * BK + 4 * (j + 4 * k) < p.D)
                                v[i][k] = __ldg(reinterpret_cast<const float4*>(rp + 16 * k));
                        }
                    }
                    uint8_t* dst = sA + (size_t)slot * A_CHUNK_BYTES;
                    const bool first = n == 0;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const float4 t = v[i][k];
                            const __half2 lo = __floats2half2_rn(t.x, t.y);
                            const __half2 hi = __floats2half2_rn(t.z, t.w);
                            uint2 pk;
                            pk.x = *reinterpret_cast<const uint32_t*>(&lo);
                            pk.y = *reinterpret_cast<const uint32_t*>(&hi);
                            *reinterpret_cast<uint2*>(dst + i * 4096 + soff[k]) = pk;
                            if (first) {
                                const float2 fl = __half22float2(lo), fh = __half22float2(hi);
                                float d;
                                sq[i] = fmaf(t.x, t.x, sq[i]); d = t.x - fl.x; rq[i] = fmaf(d, d, rq[i]);
                                sq[i] = fmaf(t.y, t.y, sq[i]); d = t.y - fl.y; rq[i] = fmaf(d, d, rq[i]);
                                sq[i] = fmaf(t.z, t.z, sq[i]); d = t.z - fh.x; rq[i] = fmaf(d, d, rq[i]);
                                sq[i] = fmaf(t.w, t.w, sq[i]); d = t.w - fh.y; rq[i] = fmaf(d, d, rq[i]);
                            }
                        }
                    }
                    fence_async_smem();
                    mbar_arrive(smem_u32(a_full + slot));
                }
                ring_base += (uint32_t)p.kcn;
                if (n == 0) {
                    // lane j of the quad gets the totals of row rr + 32j over this group's chunks
                    const float s_g = transpose_reduce4(sq, j), r_g = transpose_reduce4(rq, j);
                    const int row = rr + 32 * j;
                    if (grp == 1) {
                        mbar_wait(smem_u32(part_empty), (tile_it & 1u) ^ 1u);
                        s_part[row] = make_float2(s_g, r_g);
                        mbar_arrive(smem_u32(part_full));
                    } else {
                        mbar_wait(smem_u32(part_full), tile_it & 1u);
                        const float2 o = s_part[row];
                        mbar_arrive(smem_u32(part_empty));
                        const float s = s_g + o.x, r = r_g + o.y;
                        const float e2max = __uint_as_float(__ldg(p.head_stats + eh * 4 + 1));
                        const float emax = sqrtf(e2max);
                        const float remax = sqrtf(__uint_as_float(__ldg(p.head_stats + eh * 4 + 2)));
                        const float xn = sqrtf(s);
                        float b = 2.f * (1.001f * sqrtf(r) * emax + 1.001f * xn * remax + p.c_acc * xn * emax) +
                                  p.c2 * (s + e2max);
                        // |x| >= 2^15 could overflow fp16: an infinite band sends the row to the
                        // exact re-scan (NaN/Inf rows end there as well).  The ring is 4 deep: the
                        // epilogue of tile t-4 is done before the producers of tile t can run.
                        if (!(s < 1073741824.f)) b = INFINITY;
                        s_band[(tile_it & 3u) * BM + row] = b;
                        mbar_arrive(smem_u32(band_full + (tile_it & 3u)));
                    }
                }
            }
        }
    } else if (warp < EPI_WARPS) {
        // ================= epilogue: TMEM -> registers, running best + candidate band =================
        // One row per thread; warps w and w+4 share TMEM lane quarter w&3 and take alternate
        // 32-column chunks (two warps per scheduler hide each other's latencies).  Every chunk is
        // read from TMEM once: FFMA + FMNMX tree for its max, then -- only when some row of the warp
        // has a score within the band of its running best -- a branch-free pass that records the
        // codes with score >= best - band in three index slots + a count.  No values are kept:
        // when a later chunk beats the old best by more than the band the list is simply cleared,
        // otherwise it stays a superset of the in-band codes, which is all the fp32 re-scoring
        // needs.  More than three recorded codes -> exact re-scan of the row.  (The first version
        // swept every tile twice with the TMEM load latency exposed: 3.6 us per 256-code tile
        // whatever D was, i.e. the epilogue, not the tensor pipe, set the rate.)
        const int quarter = warp & 3, half = warp >> 2;
        const int row_in_tile = quarter * 32 + lane;
        uint32_t acc_it = 0, tile_it = 0;
        for (int64_t item = blockIdx.x; item < n_items; item += gridDim.x, ++tile_it) {
            const int64_t h = item / row_tiles, mt = item - h * row_tiles;
            const int64_t eh = p.embed_heads == 1 ? 0 : h;
            const int64_t m = mt * BM + row_in_tile;
            const float scale = __uint_as_float(__ldg(p.head_stats + eh * 4 + 3));
            float rb = -INFINITY, band = 0.f;
            int cnt = 0, p0 = -1, p1 = -1, p2 = -1;
            // one 32-column chunk, already in registers: running best, and the codes within the
            // band of it.  thr uses the best INCLUDING this chunk, so a code inside the final band
            // is always recorded (the final best is >= every running best); the list is cleared
            // when a chunk beats everything before it by more than the band.
            auto sweep = [&](float (&v)[32], uint32_t ne2s, int c0, int n) {
                float cm = -INFINITY;
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    const float4 na = lds128(ne2s + (uint32_t)(c0 + g * 8) * 4u);
                    const float4 nb = lds128(ne2s + (uint32_t)(c0 + g * 8 + 4) * 4u);
                    float* w = v + g * 8;
                    w[0] = fmaf(w[0], scale, na.x); w[1] = fmaf(w[1], scale, na.y);
                    w[2] = fmaf(w[2], scale, na.z); w[3] = fmaf(w[3], scale, na.w);
                    w[4] = fmaf(w[4], scale, nb.x); w[5] = fmaf(w[5], scale, nb.y);
                    w[6] = fmaf(w[6], scale, nb.z); w[7] = fmaf(w[7], scale, nb.w);
                    cm = fmaxf(cm, fmaxf(fmaxf(fmaxf(w[0], w[1]), fmaxf(w[2], w[3])),
                                         fmaxf(fmaxf(w[4], w[5]), fmaxf(w[6], w[7]))));
                }
                if (cm - band > rb) { cnt = 0; p0 = -1; p1 = -1; p2 = -1; }
                rb = fmaxf(rb, cm);
                const float thr = rb - band;
                if (__any_sync(0xffffffffu, cm >= thr)) {
                    const int base = n * BN + c0;
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const bool hit = v[i] >= thr;
                        p2 = hit ? p1 : p2;
                        p1 = hit ? p0 : p1;
                        p0 = hit ? base + i : p0;
                        cnt += hit ? 1 : 0;
                    }
                }
            };
#pragma unroll 1
            for (int n = 0; n < p.n_tiles; ++n, ++acc_it) {
                const int ncols = tile_codes32(p, n);
                const uint32_t st = acc_it & 1u, ns = acc_it & 3u;
                if (n == 0) {
                    mbar_wait(smem_u32(band_full + (tile_it & 3u)), (tile_it >> 2) & 1u);
                    band = s_band[(tile_it & 3u) * BM + row_in_tile];
                }
                mbar_wait(smem_u32(ne2_full + ns), (acc_it >> 2) & 1u);
                const uint32_t ne2s = smem_u32(s_ne2 + ns * BN);
                mbar_wait(smem_u32(t_full + st), (acc_it >> 1) & 1u);
                tc_fence_after();
                const uint32_t t_addr = tmem_base + ((uint32_t)(quarter * 32) << 16) + st * BN + (uint32_t)(half * 32);
                // this warp's chunks are columns half*32 + 64 i; the TMEM load of chunk i+1 is in
                // flight while chunk i is swept (two register buffers)
                const int nch = (ncols - half * 32 + 63) >> 6;
                float va[32], vb[32];
                if (nch > 0) tmem_ld32(t_addr, va);
#pragma unroll 1
                for (int i = 0; i < nch; i += 2) {
                    tmem_ld_wait();
                    if (i + 1 < nch) tmem_ld32(t_addr + (uint32_t)((i + 1) * 64), vb);
                    sweep(va, ne2s, half * 32 + i * 64, n);
                    if (i + 1 < nch) {
                        tmem_ld_wait();
                        if (i + 2 < nch) tmem_ld32(t_addr + (uint32_t)((i + 2) * 64), va);
                        sweep(vb, ne2s, half * 32 + (i + 1) * 64, n);
                    }
                }
                tc_fence_before();
                mbar_arrive(smem_u32(t_empty + st));
                mbar_arrive(smem_u32(ne2_empty + ns));
            }
            // ---- merge the two column halves of the row and publish
            float* mg = s_merge + row_in_tile * MERGE_WORDS;
            if (half == 1) {
                mbar_wait(smem_u32(merge_empty), (tile_it & 1u) ^ 1u);
                mg[0] = rb;
                mg[1] = __int_as_float(cnt); mg[2] = __int_as_float(p0);
                mg[3] = __int_as_float(p1); mg[4] = __int_as_float(p2);
                mbar_arrive(smem_u32(merge_full));
            } else {
                mbar_wait(smem_u32(merge_full), tile_it & 1u);
                const float rb1 = mg[0];
                const int cnt1 = __float_as_int(mg[1]);
                const int o0 = __float_as_int(mg[2]), o1 = __float_as_int(mg[3]), o2 = __float_as_int(mg[4]);
                mbar_arrive(smem_u32(merge_empty));
                if (m < p.rows) {
                    const int64_t flat = h * p.rows + m;
                    const float t = fmaxf(rb, rb1) - band;
                    // a half whose own best is below the row's band contributes nothing
                    const bool use0 = rb >= t, use1 = rb1 >= t;
                    const bool lost = (use0 && cnt > 3) || (use1 && cnt1 > 3);
                    int keep = 0, only = -1;
                    int32_t* cd = p.cand + flat * CAP;
                    if (use0) {
                        if (p2 >= 0) { only = p2; cd[keep++] = p2; }
                        if (p1 >= 0) { only = p1; cd[keep++] = p1; }
                        if (p0 >= 0) { only = p0; cd[keep++] = p0; }
                    }
                    if (use1) {
                        if (o2 >= 0) { only = o2; cd[keep++] = o2; }
                        if (o1 >= 0) { only = o1; cd[keep++] = o1; }
                        if (o0 >= 0) { only = o0; cd[keep++] = o0; }
                    }
                    // NaN rows / infinite band (|x| outside the fp16 range) / lost candidates
                    if (lost || keep == 0 || !(t > -INFINITY)) {
                        p.best[flat] = -2;
                    } else if (keep == 1) {
                        p.best[flat] = only;
                    } else {
                        for (int j = keep; j < CAP; ++j) cd[j] = -1;
                        p.best[flat] = -1;
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == MMA_WARP) tmem_dealloc(tmem_base, 512);
}

// -------------------------------------------------------------------------- codebook prep
// head_stats[eh] = {max|e|, max e2, max |e - e~|^2, S} as float bits (all >= 0: uint order ==
// float order, so atomicMax works).  S = 2^(floor(log2 max|e|) - 13) puts max|e|/S in
// [2^13, 2^14): no fp16 overflow, and elements down to 2^-28 max|e| stay normal.
__global__ void vq_tc_maxabs_kernel(const float* __restrict__ embed, int64_t embed_heads, int K, int D,
                                    uint32_t* __restrict__ head_stats) {
    const int64_t gw = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (gw >= embed_heads * K) return;
    const float* e = embed + gw * D;
    float m = 0.f;
    for (int d = lane_id(); d < D; d += 32) m = fmaxf(m, fabsf(__ldg(e + d)));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane_id() == 0) atomicMax(head_stats + (gw / K) * 4 + 0, __float_as_uint(m));
}

__device__ __forceinline__ float head_scale(const uint32_t* head_stats, int64_t eh) {
    const float m = __uint_as_float(head_stats[eh * 4 + 0]);
    if (!(m > 0.f) || !(m < INFINITY)) return 1.f;
    int ex = ilogbf(m) - 13;
    ex = ex < -100 ? -100 : (ex > 100 ? 100 : ex);
    return ldexpf(1.f, ex);
}

// One thread per 16-byte chunk (8 fp16) of the swizzled tile image; padding is zero.
__global__ void vq_tc_image_kernel(const float* __restrict__ embed, int64_t embed_heads, int K, int D,
                                   int n_tiles, int kcn, const uint32_t* __restrict__ head_stats,
                                   __half* __restrict__ img) {
    const int64_t total = embed_heads * n_tiles * (int64_t)kcn * BN * 8;
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= total) return;
    const int j = (int)(t & 7);                 // logical 16-byte chunk in the 128 B row
    const int r = (int)((t >> 3) % BN);         // row (code) in the tile
    int64_t tile = (t >> 3) / BN;               // (eh, n, kc) flattened
    const int kc = (int)(tile % kcn);
    const int n = (int)((tile / kcn) % n_tiles);
    const int64_t eh = tile / kcn / n_tiles;
    const float inv = 1.f / head_scale(head_stats, eh);     // power of two: exact
    const int code = n * BN + r, col = kc * BK + j * 8;
    float f[8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
        f[i] = (code < K && col + i < D) ? __ldg(embed + (eh * K + code) * (int64_t)D + col + i) * inv : 0.f;
    uint4 pk;
    const __half2 b0 = __floats2half2_rn(f[0], f[1]), b1 = __floats2half2_rn(f[2], f[3]);
    const __half2 b2 = __floats2half2_rn(f[4], f[5]), b3 = __floats2half2_rn(f[6], f[7]);
    pk.x = *reinterpret_cast<const uint32_t*>(&b0); pk.y = *reinterpret_cast<const uint32_t*>(&b1);
    pk.z = *reinterpret_cast<const uint32_t*>(&b2); pk.w = *reinterpret_cast<const uint32_t*>(&b3);
    const int64_t off = tile * B_STAGE_BYTES + (int64_t)(r >> 3) * 1024 + (r & 7) * 128 + ((j ^ (r & 7)) << 4);
    *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(img) + off) = pk;
}

// One warp per (padded) code: ne2 = -e2/2 (Euclid) | 0 (dot), -inf for padding; max e2 and the
// max squared fp16 residual |e - S fp16(e/S)|^2 go to head_stats; head_stats[3] = S.
__global__ void vq_tc_ne2_kernel(const float* __restrict__ embed, const float* __restrict__ e2,
                                 int64_t embed_heads, int K, int D, int n_tiles, int metric,
                                 float* __restrict__ ne2, uint32_t* __restrict__ head_stats) {
    const int64_t kp = (int64_t)n_tiles * BN;
    const int64_t gw = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (gw >= embed_heads * kp) return;
    const int64_t eh = gw / kp;
    const int c = (int)(gw - eh * kp);
    if (c >= K) {
        if (lane_id() == 0) ne2[gw] = -INFINITY;
        return;
    }
    const float S = head_scale(head_stats, eh), inv = 1.f / S;
    const float* e = embed + (eh * K + c) * (int64_t)D;
    float r = 0.f;
    for (int d = lane_id(); d < D; d += 32) {
        const float v = __ldg(e + d);
        const float df = v - __half2float(__float2half_rn(v * inv)) * S;
        r = fmaf(df, df, r);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
    if (lane_id() == 0) {
        const float ee = e2[eh * K + c];
        ne2[gw] = metric == 0 ? -0.5f * ee : 0.f;
        atomicMax(head_stats + eh * 4 + 1, __float_as_uint(ee));
        atomicMax(head_stats + eh * 4 + 2, __float_as_uint(r * 1.0001f));
        if (c == 0) head_stats[eh * 4 + 3] = __float_as_uint(S);
    }
}

struct Plan {
    int n_tiles, kcn, sb, a_slots, a_dbl;
    bool resident;
    size_t smem, img_bytes, ne2_bytes;
};

static bool make_plan(const VqView& q, Plan& pl) {
    pl.n_tiles = (q.K + BN - 1) / BN;
    pl.kcn = (q.D + BK - 1) / BK;
    const size_t avail = SMEM_MAX - SMEM_EXTRA;
    const size_t a_res = (size_t)pl.kcn * A_CHUNK_BYTES;
    // The codebook stream is latency-bound: every 32 KB stage is a bulk copy out of L2 (~1-2 us
    // under load) that feeds 4 MMAs (~0.3 us), so the ring wants as many stages as fit.  Measured at
    // 1M x 16384: D=128 with 2 stages 568 TFLOP/s, D=256 with 2 stages + double-buffered A 735.
    pl.resident = a_res + 2 * (size_t)B_STAGE_BYTES <= avail;
    if (pl.resident) {
        const size_t sb1 = (avail - a_res) / B_STAGE_BYTES;
        const size_t sb2 = 2 * a_res + 2 * (size_t)B_STAGE_BYTES <= avail && 2 * pl.kcn <= MAX_AS
                               ? (avail - 2 * a_res) / B_STAGE_BYTES : 0;
        // second A buffer (next row tile loads under this tile's MMAs): always when it still leaves
        // 4 stages; with few code tiles per row tile (the reload is a large share) down to 2 stages
        bool dbl = sb2 >= 4 || (sb2 >= 2 && pl.n_tiles < 8);
        if (const char* ev = getenv("STEMGNN_VQ_DOUBLE_A")) dbl = sb2 >= 2 && atoi(ev) != 0;   // tuning knob
        pl.a_slots = pl.kcn * (dbl ? 2 : 1);
        pl.a_dbl = dbl ? 1 : 0;
        const size_t sb = dbl ? sb2 : sb1;
        pl.sb = (int)(sb > (size_t)MAX_SB ? (size_t)MAX_SB : sb);
        if (const char* ev = getenv("STEMGNN_VQ_STAGES")) pl.sb = atoi(ev) >= 2 && atoi(ev) < pl.sb ? atoi(ev) : pl.sb;
    } else {
        pl.a_slots = 4;
        pl.a_dbl = 0;
        pl.sb = 4;
    }
    if (pl.a_slots > MAX_AS || pl.sb > MAX_SB || pl.sb < 2) return false;
    pl.smem = (size_t)pl.a_slots * A_CHUNK_BYTES + (size_t)pl.sb * B_STAGE_BYTES + SMEM_EXTRA;
    // more than half of the SM's shared memory: one CTA per SM (it owns all 512 TMEM columns)
    if (pl.smem < 120 * 1024) pl.smem = 120 * 1024;
    pl.img_bytes = (size_t)q.embed_heads * pl.n_tiles * pl.kcn * B_STAGE_BYTES;
    pl.ne2_bytes = (size_t)q.embed_heads * pl.n_tiles * BN * 4;
    return pl.smem <= SMEM_MAX;
}

}  // namespace tc

bool vq_tc_supported(const VqView& q) {
    tc::Plan pl;
    const bool aligned = q.D % 4 == 0 && q.x_row_stride % 4 == 0 && q.x_head_stride % 4 == 0 &&
                         (reinterpret_cast<uintptr_t>(q.x) & 15) == 0;
    return aligned && q.D >= 16 && q.K >= 1 && tc::make_plan(q, pl);
}

int vq_assign_tc_workspace(const VqView& q, size_t* bytes) {
    tc::Plan pl;
    if (!tc::make_plan(q, pl)) return STEMGNN_EUNSUPPORTED;
    const size_t items = (size_t)(q.heads * q.rows);
    size_t b = 0;
    b += align_up(pl.img_bytes, 1024);
    b += align_up(pl.ne2_bytes, 256);
    b += align_up((size_t)q.embed_heads * 16, 256);          // head_stats
    b += align_up(items * tc::CAP * 4, 256);                 // candidates
    *bytes = b + 1024;
    return STEMGNN_OK;
}

int vq_assign_tc(const VqView& q, const float* e2, int32_t* best, int32_t** cand_out, void* ws_ptr,
                 size_t ws_bytes, cudaStream_t stream) {
    using namespace tc;
    Plan pl;
    if (!make_plan(q, pl)) {
        set_error("vq_assign_tc: shape does not fit the tcgen05 plan");
        return STEMGNN_EUNSUPPORTED;
    }
    size_t need = 0;
    vq_assign_tc_workspace(q, &need);
    if (ws_ptr == nullptr || ws_bytes < need) {
        set_error("vq_assign_tc: workspace too small (%zu < %zu)", ws_bytes, need);
        return STEMGNN_ENOMEM_WS;
    }
    // 1024-byte alignment for the bulk-copy source images
    char* base = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(ws_ptr) + 1023) & ~(uintptr_t)1023);
    Workspace ws(base, ws_bytes - (size_t)(base - reinterpret_cast<char*>(ws_ptr)));
    const size_t items = (size_t)(q.heads * q.rows);
    __half* img = reinterpret_cast<__half*>(ws.take<char>(align_up(pl.img_bytes, 1024)));
    float* ne2 = ws.take<float>(pl.ne2_bytes / 4);
    uint32_t* head_stats = ws.take<uint32_t>((size_t)q.embed_heads * 4);
    int32_t* cand = ws.take<int32_t>(items * CAP);
    *cand_out = cand;
    cudaMemsetAsync(head_stats, 0, (size_t)q.embed_heads * 16, stream);

    const int64_t ek = q.embed_heads * (int64_t)q.K;
    SG_LAUNCH(vq_tc_maxabs_kernel<<<(unsigned)((ek * 32 + 255) / 256), 256, 0, stream>>>(
        q.embed, q.embed_heads, q.K, q.D, head_stats));
    const int64_t img_threads = (int64_t)q.embed_heads * pl.n_tiles * pl.kcn * BN * 8;
    SG_LAUNCH(vq_tc_image_kernel<<<(unsigned)((img_threads + 255) / 256), 256, 0, stream>>>(
        q.embed, q.embed_heads, q.K, q.D, pl.n_tiles, pl.kcn, head_stats, img));
    const int64_t ne2_warps = (int64_t)q.embed_heads * pl.n_tiles * BN;
    SG_LAUNCH(vq_tc_ne2_kernel<<<(unsigned)((ne2_warps * 32 + 255) / 256), 256, 0, stream>>>(
        q.embed, e2, q.embed_heads, q.K, q.D, pl.n_tiles, q.metric, ne2, head_stats));

    TcParams p;
    p.x = q.x; p.x_row_stride = q.x_row_stride; p.x_head_stride = q.x_head_stride;
    p.heads = q.heads; p.embed_heads = q.embed_heads; p.rows = q.rows;
    p.K = q.K; p.D = q.D; p.n_tiles = pl.n_tiles; p.kcn = pl.kcn; p.sb = pl.sb; p.a_slots = pl.a_slots; p.a_dbl = pl.a_dbl;
    p.img = img; p.ne2 = ne2; p.head_stats = head_stats;
    // fp32 accumulation of D exact fp16 x fp16 products (tensor core) and of the fp32 re-score
    // chain: <= D 2^-22 |x||e| together.  The Euclidean score adds the fp32 roundings of
    // x2 + e2 - 2 dot and of acc*S - e2/2: a few ulps of (x2 + e2), covered by 2^-20 (x2 + e2max).
    p.c_acc = (float)q.D * 2.3841858e-7f;
    p.c2 = q.metric == 0 ? 9.5367432e-7f : 0.f;
    p.best = best; p.cand = cand;

    const int64_t n_items = ((q.rows + BM - 1) / BM) * q.heads;
    const int grid = (int)(n_items < sm_count() ? n_items : sm_count());
    cudaError_t e;
    if (pl.resident) {
        e = cudaFuncSetAttribute(vq_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem);
        if (e == cudaSuccess) SG_LAUNCH(vq_tc_kernel<true><<<grid, NUM_THREADS, pl.smem, stream>>>(p));
    } else {
        e = cudaFuncSetAttribute(vq_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem);
        if (e == cudaSuccess) SG_LAUNCH(vq_tc_kernel<false><<<grid, NUM_THREADS, pl.smem, stream>>>(p));
    }
    if (e != cudaSuccess) {
        set_error("vq_assign_tc: cudaFuncSetAttribute(%zu B smem): %s", pl.smem, cudaGetErrorString(e));
        return STEMGNN_ECUDA;
    }
    cudaError_t le = cudaGetLastError();
    if (le != cudaSuccess) {
        set_error("vq_assign_tc: CUDA error %s", cudaGetErrorString(le));
        return STEMGNN_ECUDA;
    }
    return STEMGNN_OK;
}

}  // namespace stemgnn
