// k10_sampler.cu -- K10: fan-out neighbour sampling on the GPU.
//
// Replaces torch_geometric.loader.NeighborLoader's sampling step (torch_geometric 2.3.0 ->
// torch_sparse 0.6.15 `neighbor_sample`, replace=False, directed=True) as the reference drives it
// at STEM-GNN/pretrain.py:151-153 (num_neighbors=[10]*L, batch 1,024) and utils/loader.py:11-24
// ([10]*L training, [-1]*L inference):
//
//   nodes = seeds (in order);  frontier = all of them
//   for every hop: for every frontier node v in order: take min(deg_in(v), k) distinct in-edges
//   (all of them, in CSR order, when k < 0 or deg <= k); every sampled source not seen before is
//   appended to `nodes` in order of first appearance; the edge is stored as (local source, local v,
//   original edge id); the next frontier is the nodes appended in this hop.
//
// The CPU sampler walks this with a hash map; here each hop is a handful of ordered stages (scan.cuh)
// over launch BOUNDS known on the host (|frontier| * k), with the true sizes kept in device memory:
// no host synchronisation inside a batch, so the sampler can be captured into a CUDA graph with the
// step it feeds.  First appearance is decided by an atomicMax race on "claims" (-(2 + slot)) in a
// node -> local-id map that persists across batches (all INT32_MIN outside a call), which makes the
// relabelling deterministic and identical to the sequential algorithm.
// Random subsets use Floyd's algorithm on Philox draws keyed by (seed, frontier position, hop).
#include <limits.h>

#include "common.cuh"
#include "philox.cuh"
#include "scan.cuh"

namespace stemgnn {
namespace smp {

constexpr int SCAN_ITEMS = 1;    // one frontier node / edge slot per thread: spreads a 10k-node hop over the SMs
constexpr int MAX_FANOUT = 64;   // random subsets (deg > k) keep the chosen slots in local memory
enum { ST_BEGIN = 0, ST_END, ST_EDGE_BASE, ST_HOP_EDGES, ST_NEW_NODES, ST_ERROR, ST_WORDS = 8 };
enum { ERR_SEED_RANGE = 1, ERR_EDGE_CAP = 2, ERR_NODE_CAP = 4 };

__device__ __forceinline__ int claim_of(int64_t slot) { return -(int)(2 + slot); }

struct View {
    const int32_t* rowptr; const int32_t* col; const int32_t* perm;
    int64_t num_nodes, node_cap, edge_cap;
    int32_t* map; int64_t* n_id; int64_t* src; int64_t* dst; int64_t* e_id;
    int64_t* st;
};

__global__ void seed_claim_kernel(View g, const int64_t* __restrict__ seeds, int64_t B) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B) return;
    const int64_t v = seeds[i];
    g.n_id[i] = v;
    if (v < 0 || v >= g.num_nodes) { atomicOr((unsigned long long*)&g.st[ST_ERROR], (unsigned long long)ERR_SEED_RANGE); return; }
    atomicMax(&g.map[v], claim_of(i));
}

__global__ void seed_assign_kernel(View g, const int64_t* __restrict__ seeds, int64_t B) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B) return;
    const int64_t v = seeds[i];
    if (v < 0 || v >= g.num_nodes) return;
    if (g.map[v] == claim_of(i)) g.map[v] = (int)i;   // a duplicated seed keeps its first position
}

__global__ void state_init_kernel(int64_t* st, int64_t B) {
    st[ST_BEGIN] = 0; st[ST_END] = B; st[ST_EDGE_BASE] = 0; st[ST_HOP_EDGES] = 0; st[ST_NEW_NODES] = 0;
    st[ST_ERROR] = 0;
}

struct CountNbr {
    View g; int fanout;
    __device__ __forceinline__ int operator()(int64_t i) const {
        const int64_t b = g.st[ST_BEGIN];
        if (i >= g.st[ST_END] - b) return 0;
        const int64_t v = g.n_id[b + i];
        if (v < 0 || v >= g.num_nodes) return 0;
        const int deg = g.rowptr[v + 1] - g.rowptr[v];
        return (fanout < 0 || deg < fanout) ? deg : fanout;
    }
};

struct PickNbr {
    View g; int fanout; int hop; const uint64_t* seed;
    __device__ void operator()(int64_t i, int64_t pos, int c) const {
        const int64_t tgt = g.st[ST_BEGIN] + i;
        const int64_t base = g.st[ST_EDGE_BASE] + pos;
        if (base + c > g.edge_cap) { atomicOr((unsigned long long*)&g.st[ST_ERROR], (unsigned long long)ERR_EDGE_CAP); return; }
        const int64_t v = g.n_id[tgt];
        const int p0 = g.rowptr[v];
        const int deg = g.rowptr[v + 1] - p0;
        if (c == deg) {                       // every in-edge, in CSR (= original edge) order
            for (int q = 0; q < c; ++q) {
                g.src[base + q] = g.col[p0 + q];
                g.dst[base + q] = tgt;
                g.e_id[base + q] = g.perm[p0 + q];
            }
            return;
        }
        // Floyd: a uniform c-subset of {0..deg-1} in c draws
        int chosen[MAX_FANOUT];
        const uint64_t s = *seed;
        uint4 r = make_uint4(0u, 0u, 0u, 0u);
        for (int q = 0; q < c; ++q) {
            if ((q & 3) == 0) r = philox_draw4(s, (uint64_t)tgt, ((uint32_t)hop << 16) | (uint32_t)(16 + (q >> 2)));
            const uint32_t w = (q & 3) == 0 ? r.x : (q & 3) == 1 ? r.y : (q & 3) == 2 ? r.z : r.w;
            const int j = deg - c + q;
            int t = (int)(((uint64_t)w * (uint64_t)(j + 1)) >> 32);
            for (int u = 0; u < q; ++u)
                if (chosen[u] == t) { t = j; break; }
            chosen[q] = t;
            g.src[base + q] = g.col[p0 + t];
            g.dst[base + q] = tgt;
            g.e_id[base + q] = g.perm[p0 + t];
        }
    }
};

// global source id stored in slot s of the current hop, or -1 when the slot does not exist (beyond
// the hop's edge count, or beyond the output capacity after an ERR_EDGE_CAP)
__device__ __forceinline__ int64_t hop_source(const View& g, int64_t s) {
    if (s >= g.st[ST_HOP_EDGES]) return -1;
    const int64_t e = g.st[ST_EDGE_BASE] + s;
    if (e >= g.edge_cap) return -1;
    const int64_t v = g.src[e];
    return (v >= 0 && v < g.num_nodes) ? v : -1;
}

__global__ void claim_kernel(View g, int64_t bound) {
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= bound) return;
    const int64_t v = hop_source(g, s);
    if (v >= 0 && g.map[v] < 0) atomicMax(&g.map[v], claim_of(s));
}

struct IsFirst {
    View g;
    __device__ __forceinline__ int operator()(int64_t s) const {
        const int64_t v = hop_source(g, s);
        return (v >= 0 && g.map[v] == claim_of(s)) ? 1 : 0;
    }
};

struct AssignId {
    View g;
    __device__ __forceinline__ void operator()(int64_t s, int64_t pos, int) const {
        const int64_t id = g.st[ST_END] + pos;
        const int64_t v = g.src[g.st[ST_EDGE_BASE] + s];
        if (id >= g.node_cap) { atomicOr((unsigned long long*)&g.st[ST_ERROR], (unsigned long long)ERR_NODE_CAP); return; }
        g.n_id[id] = v;
        g.map[v] = (int)id;
    }
};

__global__ void relabel_kernel(View g, int64_t bound) {
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= bound) return;
    const int64_t v = hop_source(g, s);
    if (v >= 0) g.src[g.st[ST_EDGE_BASE] + s] = g.map[v];
}

__global__ void advance_kernel(int64_t* st, int64_t* counts, int hop) {
    st[ST_BEGIN] = st[ST_END];
    st[ST_END] += st[ST_NEW_NODES];
    st[ST_EDGE_BASE] += st[ST_HOP_EDGES];
    if (counts) { counts[2 + 2 * hop] = st[ST_END]; counts[3 + 2 * hop] = st[ST_EDGE_BASE]; }
}

__global__ void cleanup_kernel(View g, int64_t bound, int64_t* counts) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0 && counts) { counts[0] = g.st[ST_END]; counts[1] = g.st[ST_EDGE_BASE]; }
    if (i >= bound || i >= g.st[ST_END]) return;
    const int64_t v = g.n_id[i];
    if (v >= 0 && v < g.num_nodes) g.map[v] = INT_MIN;
}

__global__ void fill_i32_kernel(int32_t* p, int64_t n, int32_t v) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

struct Bounds { int64_t node_cap, edge_cap, max_stage; };

static int bounds(int64_t B, int64_t N, int64_t E, int hops, const int32_t* fanout, Bounds* out) {
    int64_t front = B, nodes = B, edges = 0, mx = B;
    for (int h = 0; h < hops; ++h) {
        const int k = fanout[h];
        if (k > MAX_FANOUT) return -1;
        int64_t eb = k < 0 ? E : front * (int64_t)k;
        if (h > 0 && eb > E) eb = E;      // a later frontier holds distinct nodes: at most every edge once
        edges += eb;
        front = eb < N ? eb : N;
        nodes += front;
        if (eb > mx) mx = eb;
        if (front > mx) mx = front;
    }
    out->node_cap = nodes; out->edge_cap = edges; out->max_stage = mx;
    return 0;
}

}  // namespace smp
}  // namespace stemgnn

using namespace stemgnn;

extern "C" int stemgnn_neighbor_sample_bounds(int64_t num_seeds, int64_t num_nodes, int64_t num_edges,
                                              int32_t num_hops, const int32_t* fanout_host,
                                              int64_t* node_cap_host, int64_t* edge_cap_host,
                                              size_t* workspace_bytes_host) {
    SG_CHECK_ARG(num_seeds >= 0 && num_nodes >= 0 && num_edges >= 0, "negative size");
    SG_CHECK_ARG(num_hops >= 0 && num_hops <= 16 && (num_hops == 0 || fanout_host), "num_hops must be in [0, 16]");
    SG_CHECK_ARG(num_nodes < (int64_t)2147483647 && num_edges < (int64_t)2147483647, "int32 CSR: E and N must be < 2^31");
    smp::Bounds b;
    if (smp::bounds(num_seeds, num_nodes, num_edges, num_hops, fanout_host, &b)) {
        set_error("stemgnn_neighbor_sample: fan-out above %d is not supported (use -1 for all neighbours)", smp::MAX_FANOUT);
        return STEMGNN_EUNSUPPORTED;
    }
    SG_CHECK_ARG(b.edge_cap < (int64_t)2147483000 && b.node_cap < (int64_t)2147483000, "sample bound exceeds 2^31");
    if (node_cap_host) *node_cap_host = b.node_cap;
    if (edge_cap_host) *edge_cap_host = b.edge_cap;
    if (workspace_bytes_host)
        *workspace_bytes_host = align_up(smp::ST_WORDS * 8, 256) + align_up((size_t)scan::num_tiles(b.max_stage, smp::SCAN_ITEMS) * 4, 256);
    return STEMGNN_OK;
}

extern "C" int stemgnn_node_map_init(int32_t* node_map, int64_t num_nodes, stemgnn_stream_t stream) {
    SG_CHECK_ARG(num_nodes >= 0 && (num_nodes == 0 || node_map), "bad arguments");
    if (num_nodes == 0) return STEMGNN_OK;
    SG_LAUNCH(smp::fill_i32_kernel<<<(unsigned)((num_nodes + 255) / 256), 256, 0, (cudaStream_t)stream>>>(node_map, num_nodes, INT_MIN));
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}

extern "C" int stemgnn_neighbor_sample(const stemgnn_neighbor_sample_args* a, void* workspace,
                                       size_t workspace_bytes, stemgnn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SG_CHECK_ARG(a != nullptr, "args is NULL");
    SG_CHECK_ARG(a->rowptr && (a->num_edges == 0 || (a->col && a->perm)) && (a->num_nodes == 0 || a->node_map), "NULL graph pointer");
    SG_CHECK_ARG(a->counts, "counts is NULL");
    SG_CHECK_ARG(a->num_seeds == 0 || a->seeds, "seeds is NULL");
    int64_t node_cap = 0, edge_cap = 0;
    size_t need = 0;
    int rc = stemgnn_neighbor_sample_bounds(a->num_seeds, a->num_nodes, a->num_edges, a->num_hops,
                                            a->fanout_host, &node_cap, &edge_cap, &need);
    if (rc) return rc;
    SG_CHECK_ARG(a->node_cap >= node_cap && a->edge_cap >= edge_cap, "output capacity below stemgnn_neighbor_sample_bounds");
    SG_CHECK_ARG(a->n_id && (edge_cap == 0 || (a->edge_index_out && a->e_id)), "NULL output pointer");
    bool random = false;
    for (int h = 0; h < a->num_hops; ++h) random = random || a->fanout_host[h] >= 0;
    SG_CHECK_ARG(!random || a->seed, "a Philox seed (device uint64) is needed when a fan-out is >= 0");
    if (workspace == nullptr || workspace_bytes < need) {
        set_error("stemgnn_neighbor_sample: workspace too small (%zu < %zu)", workspace_bytes, need);
        return STEMGNN_ENOMEM_WS;
    }
    Workspace ws(workspace, workspace_bytes);
    int64_t* st = ws.take<int64_t>(smp::ST_WORDS);
    int32_t* tiles = (int32_t*)(ws.base + ws.off);
    smp::View g{a->rowptr, a->col, a->perm, a->num_nodes, a->node_cap, a->edge_cap, a->node_map, a->n_id,
                a->edge_index_out, a->edge_index_out ? a->edge_index_out + a->edge_cap : nullptr, a->e_id, st};
    const int T = 256;
    const int64_t B = a->num_seeds;
    SG_LAUNCH(smp::state_init_kernel<<<1, 1, 0, stream>>>(st, B));
    if (B > 0) {
        SG_LAUNCH(smp::seed_claim_kernel<<<(unsigned)((B + T - 1) / T), T, 0, stream>>>(g, a->seeds, B));
        SG_LAUNCH(smp::seed_assign_kernel<<<(unsigned)((B + T - 1) / T), T, 0, stream>>>(g, a->seeds, B));
    }
    int64_t front = B;
    for (int h = 0; h < a->num_hops; ++h) {
        const int k = a->fanout_host[h];
        int64_t eb = k < 0 ? a->num_edges : front * (int64_t)k;
        if (h > 0 && eb > a->num_edges) eb = a->num_edges;
        scan::ordered_emit<smp::SCAN_ITEMS>(smp::CountNbr{g, k}, smp::PickNbr{g, k, h, a->seed}, front, tiles, st + smp::ST_HOP_EDGES, stream);
        if (eb > 0) SG_LAUNCH(smp::claim_kernel<<<(unsigned)((eb + T - 1) / T), T, 0, stream>>>(g, eb));
        scan::ordered_emit<smp::SCAN_ITEMS>(smp::IsFirst{g}, smp::AssignId{g}, eb, tiles, st + smp::ST_NEW_NODES, stream);
        if (eb > 0) SG_LAUNCH(smp::relabel_kernel<<<(unsigned)((eb + T - 1) / T), T, 0, stream>>>(g, eb));
        SG_LAUNCH(smp::advance_kernel<<<1, 1, 0, stream>>>(st, a->counts, h));
        front = eb < a->num_nodes ? eb : a->num_nodes;
    }
    const int64_t nb = a->node_cap > 0 ? a->node_cap : 1;
    SG_LAUNCH(smp::cleanup_kernel<<<(unsigned)((nb + T - 1) / T), T, 0, stream>>>(g, a->node_cap, a->counts));
    // counts[0] nodes, counts[1] edges, counts[2+2h], counts[3+2h] cumulative per hop; the error word
    // travels in counts[2 + 2*num_hops]
    SG_CHECK_CUDA(cudaMemcpyAsync(a->counts + 2 + 2 * a->num_hops, st + smp::ST_ERROR, 8, cudaMemcpyDeviceToDevice, stream));
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}
