// k12_negative.cu -- K12: negative edge sampling for the topology-reconstruction loss.
//
// Replaces torch_geometric.utils.negative_sampling(pos_edge_index, num_nodes) (2.3.0, "sparse"
// method) as called at reference STEM-GNN/model/pt_model.py:60.  PyG moves the
// edge list to the CPU, draws candidates with Python's `random.sample` and filters them with
// `numpy.isin`, up to three times.  Here, on the device:
//   * a pair (i, j), i != j, is the integer i*(N-1) + j - [i < j] in [0, N(N-1)) -- PyG's own
//     vectorisation, self loops cannot be drawn;
//   * the positive edges go into an open-addressing hash set of those integers;
//   * candidates are Philox draws keyed by (*seed, candidate number); each is looked up / inserted in
//     the same table: a positive is rejected, and among equal candidates the smallest candidate
//     number wins (atomicMin on the slot's owner), so duplicates are dropped deterministically;
//   * the survivors are compacted in candidate order (scan.cuh) into [2, capacity] and counted.
// The result is a set of distinct non-edges in a seed-determined order; which ones differs from
// PyG's by construction (different RNG), the distribution does not: uniform over the non-edges.
// Integer work; the table is 12 B per slot at load factor <= 1/2.
#include <limits.h>

#include "common.cuh"
#include "philox.cuh"
#include "scan.cuh"

namespace stemgnn {
namespace neg {

constexpr unsigned long long EMPTY = ~0ull;
constexpr int OWNER_POSITIVE = -1;
constexpr int SCAN_ITEMS = 8;

__device__ __forceinline__ uint64_t mix(uint64_t k) {       // splitmix64 finaliser
    k ^= k >> 30; k *= 0xbf58476d1ce4e5b9ull; k ^= k >> 27; k *= 0x94d049bb133111ebull; k ^= k >> 31;
    return k;
}

struct Table { unsigned long long* keys; int* owner; uint64_t mask; };

// slot holding `key` after this call (inserted if absent)
__device__ __forceinline__ uint64_t find_or_insert(const Table& t, unsigned long long key) {
    uint64_t s = mix(key) & t.mask;
    while (true) {
        unsigned long long cur = t.keys[s];
        if (cur == EMPTY) cur = atomicCAS(&t.keys[s], EMPTY, key);
        if (cur == EMPTY || cur == key) return s;
        s = (s + 1) & t.mask;
    }
}

__global__ void init_kernel(Table t) {
    const uint64_t n = t.mask + 1;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        t.keys[i] = EMPTY;
        t.owner[i] = INT_MAX;
    }
}

__global__ void insert_positive_kernel(Table t, const int64_t* __restrict__ ei, int64_t E, int64_t N) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= E) return;
    const int64_t r = ei[e], c = ei[E + e];
    if (r == c || r < 0 || c < 0 || r >= N || c >= N) return;       // self loops are not part of the population
    const unsigned long long key = (unsigned long long)(r * (N - 1) + c - (r < c ? 1 : 0));
    t.owner[find_or_insert(t, key)] = OWNER_POSITIVE;
}

__global__ void insert_candidate_kernel(Table t, int64_t M, uint64_t population, const uint64_t* __restrict__ seed,
                                        unsigned long long* __restrict__ cand_key, int32_t* __restrict__ cand_slot) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= M) return;
    const uint4 d = philox_draw4(*seed, (uint64_t)j >> 1, 12u);
    const uint64_t r64 = (j & 1) ? ((uint64_t)d.z << 32 | d.w) : ((uint64_t)d.x << 32 | d.y);
    const unsigned long long key = __umul64hi(r64, population);     // uniform in [0, population)
    const uint64_t s = find_or_insert(t, key);
    cand_key[j] = key;
    // a positive keeps owner -1 (atomicMin with j >= 0 cannot raise it); otherwise the smallest j wins
    atomicMin(&t.owner[s], (int)j);
    cand_slot[j] = (int32_t)s;
}

struct Keep {
    Table t; const int32_t* cand_slot;
    __device__ __forceinline__ int operator()(int64_t j) const { return t.owner[cand_slot[j]] == (int)j ? 1 : 0; }
};

struct Write {
    const unsigned long long* cand_key; int64_t N, cap; int64_t* out;
    __device__ __forceinline__ void operator()(int64_t j, int64_t pos, int) const {
        if (pos >= cap) return;
        const unsigned long long key = cand_key[j];
        const int64_t r = (int64_t)(key / (unsigned long long)(N - 1));
        int64_t c = (int64_t)(key - (unsigned long long)r * (unsigned long long)(N - 1));
        if (r <= c) c += 1;
        out[pos] = r;
        out[cap + pos] = c;
    }
};

static uint64_t table_slots(int64_t E, int64_t M) {
    uint64_t need = 2ull * (uint64_t)(E + M) + 2, n = 1024;
    while (n < need) n <<= 1;
    return n;
}

}  // namespace neg
}  // namespace stemgnn

using namespace stemgnn;

extern "C" int stemgnn_negative_sampling_workspace(int64_t num_edges, int64_t num_candidates, size_t* bytes) {
    SG_CHECK_ARG(bytes != nullptr, "bytes_host is NULL");
    SG_CHECK_ARG(num_edges >= 0 && num_candidates >= 0, "negative size");
    SG_CHECK_ARG(num_edges < (int64_t)2147483647 && num_candidates < (int64_t)2147483647 &&
                 num_edges + num_candidates < (int64_t)1 << 30, "E and the candidate count must stay below 2^30 together");
    const uint64_t slots = neg::table_slots(num_edges, num_candidates);
    size_t b = 0;
    b += align_up((size_t)slots * 8, 256);                                   // keys
    b += align_up((size_t)slots * 4, 256);                                   // owner
    b += align_up((size_t)num_candidates * 8, 256);                          // candidate keys
    b += align_up((size_t)num_candidates * 4, 256);                          // candidate slots
    b += align_up((size_t)scan::num_tiles(num_candidates, neg::SCAN_ITEMS) * 4, 256);
    *bytes = b;
    return STEMGNN_OK;
}

extern "C" int stemgnn_negative_sampling(const int64_t* edge_index, int64_t E, int64_t N, int64_t M,
                                         int64_t capacity, const uint64_t* seed, int64_t* neg_out,
                                         int64_t* count_out, void* workspace, size_t workspace_bytes,
                                         stemgnn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SG_CHECK_ARG(N >= 2, "negative sampling needs at least two nodes");
    SG_CHECK_ARG(N < (int64_t)3037000000ll, "num_nodes too large for 64-bit pair ids");
    SG_CHECK_ARG(E == 0 || edge_index, "edge_index is NULL");
    SG_CHECK_ARG(capacity >= 0 && M >= 0, "negative size");
    SG_CHECK_ARG(count_out && seed && (capacity == 0 || neg_out), "NULL pointer");
    size_t need = 0;
    if (int rc = stemgnn_negative_sampling_workspace(E, M, &need)) return rc;
    if (workspace == nullptr || workspace_bytes < need) {
        set_error("stemgnn_negative_sampling: workspace too small (%zu < %zu)", workspace_bytes, need);
        return STEMGNN_ENOMEM_WS;
    }
    Workspace ws(workspace, workspace_bytes);
    const uint64_t slots = neg::table_slots(E, M);
    neg::Table t{ws.take<unsigned long long>(slots), ws.take<int>(slots), slots - 1};
    unsigned long long* cand_key = ws.take<unsigned long long>(M);
    int32_t* cand_slot = ws.take<int32_t>(M);
    int32_t* tiles = ws.take<int32_t>(scan::num_tiles(M, neg::SCAN_ITEMS));
    const int T = 256;
    SG_LAUNCH(neg::init_kernel<<<(unsigned)(sm_count() * 8), T, 0, stream>>>(t));
    if (E > 0) SG_LAUNCH(neg::insert_positive_kernel<<<(unsigned)((E + T - 1) / T), T, 0, stream>>>(t, edge_index, E, N));
    const uint64_t population = (uint64_t)N * (uint64_t)(N - 1);
    if (M > 0) SG_LAUNCH(neg::insert_candidate_kernel<<<(unsigned)((M + T - 1) / T), T, 0, stream>>>(t, M, population, seed, cand_key, cand_slot));
    scan::ordered_emit<neg::SCAN_ITEMS>(neg::Keep{t, cand_slot}, neg::Write{cand_key, N, capacity, neg_out}, M, tiles,
                                        count_out, stream);
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}
