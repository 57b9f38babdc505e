// k9_augment.cu -- K9: the graph augmentations of the pretrain step and the row gathers that
// feed it.
//
// Replaces, from reference STEM-GNN/pretrain.py:
//   :33-38  x = node_text_feat[data.x]; edge_attr = edge_text_feat[data.xe]      -> gather_rows
//   :41     aug_x, _ = mask_feature(x, p)                                          -> mask_fill
//   :42-44  dropout_adj(edge_index, edge_attr, p, force_undirected=True, ...)      -> dropout_adj
// (mask_feature / dropout_adj are torch_geometric.utils 2.3.0 functions; environment.yml:292.)
//
// dropout_adj is an ORDER-PRESERVING stream compaction: edge e survives iff rand[e] >= p and
// (when force_undirected) row[e] <= col[e]; survivors keep their relative order and the mirrored
// copies (col, row) follow in the second half, exactly as `edge_index[:, mask]` + `torch.cat`
// produce them.  Integer work, HBM-bound: E*(16+4) B read twice (count pass, write pass) and
// 8 B*(2 + 1) per output edge written (both index rows and the original edge id).
#include "common.cuh"
#include "philox.cuh"
#include "scan.cuh"

namespace stemgnn {
namespace aug {

constexpr int SCAN_ITEMS = 8;   // edge-list sized streams

struct KeepEdge {
    const int64_t* ei; int64_t E; const float* rnd; const uint64_t* seed; float p; int undirected;
    __device__ __forceinline__ int operator()(int64_t e) const {
        const float u = rnd ? __ldg(rnd + e) : philox_u01_at(*seed, (uint64_t)e, 9u);
        bool keep = u >= p;
        if (undirected) keep = keep && (__ldg(ei + e) <= __ldg(ei + E + e));
        return keep ? 1 : 0;
    }
};

struct WriteEdge {
    const int64_t* ei; int64_t E; int64_t kept; int undirected; int64_t* out; int64_t* eid;
    __device__ __forceinline__ void operator()(int64_t e, int64_t pos, int) const {
        const int64_t r = __ldg(ei + e), c = __ldg(ei + E + e);
        const int64_t ld = undirected ? 2 * kept : kept;
        out[pos] = r;
        out[ld + pos] = c;
        if (eid) eid[pos] = e;
        if (undirected) {
            out[kept + pos] = c;
            out[ld + kept + pos] = r;
            if (eid) eid[kept + pos] = e;
        }
    }
};

// out[r, c] = keep(r, c) ? x[r, c] : fill, 4 columns per thread
__global__ void __launch_bounds__(256)
mask_fill_kernel(const float* __restrict__ x, int64_t ldx, int64_t rows, int feat4,
                 const uint8_t* __restrict__ row_keep, const uint8_t* __restrict__ col_keep,
                 float p_elem, const uint64_t* __restrict__ seed_p, float fill,
                 float* __restrict__ out, int64_t ldo, uint8_t* __restrict__ mask_out) {
    const int64_t total = rows * feat4;
    const uint64_t seed = p_elem > 0.f ? *seed_p : 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / feat4;
        const int c = (int)(i - r * feat4) * 4;
        const float4 v = __ldg(reinterpret_cast<const float4*>(x + r * ldx + c));
        bool k0 = true, k1 = true, k2 = true, k3 = true;
        if (row_keep) k0 = k1 = k2 = k3 = row_keep[r] != 0;
        if (col_keep) {
            const uchar4 m = *reinterpret_cast<const uchar4*>(col_keep + c);
            k0 = k0 && m.x; k1 = k1 && m.y; k2 = k2 && m.z; k3 = k3 && m.w;
        }
        if (p_elem > 0.f) {
            const uint4 q = philox_draw4(seed, (uint64_t)i, 10u);
            k0 = k0 && u01(q.x) >= p_elem; k1 = k1 && u01(q.y) >= p_elem;
            k2 = k2 && u01(q.z) >= p_elem; k3 = k3 && u01(q.w) >= p_elem;
        }
        float4 o;
        o.x = k0 ? v.x : fill; o.y = k1 ? v.y : fill; o.z = k2 ? v.z : fill; o.w = k3 ? v.w : fill;
        *reinterpret_cast<float4*>(out + r * ldo + c) = o;
        if (mask_out) *reinterpret_cast<uchar4*>(mask_out + r * (int64_t)feat4 * 4 + c) =
            make_uchar4(k0, k1, k2, k3);
    }
}

// out[i, :] = src[ids[i], :]; LANES threads per row, 16 B per lane per trip; ids outside
// [0, n_src) give a row of NaN (torch raises a device assert)
template <int LANES>
__global__ void __launch_bounds__(256)
gather_rows_kernel(const float* __restrict__ src, int64_t lds, int64_t n_src,
                   const int64_t* __restrict__ ids, int64_t n_out, int feat4,
                   float* __restrict__ out, int64_t ldo) {
    const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / LANES;
    const int lane = threadIdx.x % LANES;
    if (row >= n_out) return;
    const int64_t id = __ldg(ids + row);
    const bool ok = id >= 0 && id < n_src;
    const float4* s = reinterpret_cast<const float4*>(src + (ok ? id : 0) * lds);
    float4* d = reinterpret_cast<float4*>(out + row * ldo);
    const float nan = __int_as_float(0x7fc00000);
#pragma unroll 4
    for (int c = lane; c < feat4; c += LANES)
        d[c] = ok ? __ldg(s + c) : make_float4(nan, nan, nan, nan);
}

}  // namespace aug
}  // namespace stemgnn

using namespace stemgnn;

extern "C" int stemgnn_dropout_adj_workspace(int64_t E, size_t* bytes) {
    SG_CHECK_ARG(bytes != nullptr, "bytes_host is NULL");
    SG_CHECK_ARG(E >= 0 && E < (int64_t)2147483647, "E must be in [0, 2^31)");
    *bytes = align_up((size_t)scan::num_tiles(E, aug::SCAN_ITEMS) * 4, 256);
    return STEMGNN_OK;
}

static int check_dropout_args(const int64_t* ei, int64_t E, const float* rnd, const uint64_t* seed,
                              float p, void* ws, size_t ws_bytes) {
    SG_CHECK_ARG(E >= 0 && E < (int64_t)2147483647, "E must be in [0, 2^31)");
    SG_CHECK_ARG(p >= 0.f && p <= 1.f, "dropout probability has to be between 0 and 1");
    SG_CHECK_ARG(E == 0 || ei, "edge_index is NULL");
    SG_CHECK_ARG(E == 0 || rnd || seed, "either rand or seed must be given");
    size_t need = 0;
    stemgnn_dropout_adj_workspace(E, &need);
    if (ws == nullptr || ws_bytes < need) {
        set_error("stemgnn_dropout_adj: workspace too small (%zu < %zu)", ws_bytes, need);
        return STEMGNN_ENOMEM_WS;
    }
    return STEMGNN_OK;
}

extern "C" int stemgnn_dropout_adj_count(const int64_t* edge_index, int64_t E, const float* rnd,
                                         const uint64_t* seed, float p, int force_undirected,
                                         int64_t* kept_out, void* workspace, size_t workspace_bytes,
                                         stemgnn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    int rc = check_dropout_args(edge_index, E, rnd, seed, p, workspace, workspace_bytes);
    if (rc) return rc;
    SG_CHECK_ARG(kept_out, "kept_out is NULL");
    int32_t* tiles = (int32_t*)workspace;
    const int nt = (int)scan::num_tiles(E, aug::SCAN_ITEMS);
    aug::KeepEdge keep{edge_index, E, rnd, seed, p, force_undirected};
    SG_LAUNCH(scan::tile_count_kernel<aug::SCAN_ITEMS><<<nt, scan::THREADS, 0, stream>>>(keep, E, tiles));
    SG_LAUNCH(scan::tile_scan_kernel<<<1, scan::THREADS, 0, stream>>>(tiles, nt, kept_out));
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}

extern "C" int stemgnn_dropout_adj_write(const int64_t* edge_index, int64_t E, const float* rnd,
                                         const uint64_t* seed, float p, int force_undirected,
                                         int64_t kept, int64_t* edge_index_out, int64_t* edge_id_out,
                                         const void* workspace, size_t workspace_bytes,
                                         stemgnn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    int rc = check_dropout_args(edge_index, E, rnd, seed, p, (void*)workspace, workspace_bytes);
    if (rc) return rc;
    SG_CHECK_ARG(kept >= 0 && kept <= E, "kept out of range");
    if (kept == 0 || E == 0) return STEMGNN_OK;
    SG_CHECK_ARG(edge_index_out, "edge_index_out is NULL");
    const int nt = (int)scan::num_tiles(E, aug::SCAN_ITEMS);
    aug::KeepEdge keep{edge_index, E, rnd, seed, p, force_undirected};
    aug::WriteEdge wr{edge_index, E, kept, force_undirected, edge_index_out, edge_id_out};
    SG_LAUNCH(scan::tile_emit_kernel<aug::SCAN_ITEMS><<<nt, scan::THREADS, 0, stream>>>(keep, wr, E, (const int32_t*)workspace));
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}

extern "C" int stemgnn_mask_fill(int64_t rows, int64_t feat, const float* x, int64_t ldx,
                                 const uint8_t* row_keep, const uint8_t* col_keep, float p_elem,
                                 const uint64_t* seed, float fill, float* out, int64_t ldo,
                                 uint8_t* mask_out, stemgnn_stream_t stream) {
    SG_CHECK_ARG(rows >= 0 && feat >= 0, "negative size");
    if (rows == 0 || feat == 0) return STEMGNN_OK;
    SG_CHECK_ARG(x && out, "NULL pointer");
    SG_CHECK_ARG(feat % 4 == 0 && ldx % 4 == 0 && ldo % 4 == 0, "feat, ldx, ldo must be multiples of 4");
    SG_CHECK_ARG(((uintptr_t)x | (uintptr_t)out) % 16 == 0, "x, out must be 16-byte aligned");
    SG_CHECK_ARG(p_elem >= 0.f && p_elem <= 1.f, "p must be between 0 and 1");
    SG_CHECK_ARG(p_elem == 0.f || seed, "element-wise masking needs a seed");
    SG_CHECK_ARG(!col_keep || (uintptr_t)col_keep % 4 == 0, "col_keep must be 4-byte aligned");
    SG_CHECK_ARG(!mask_out || (uintptr_t)mask_out % 4 == 0, "mask_out must be 4-byte aligned");
    const int64_t total = rows * (feat / 4);
    int64_t blocks = (total + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    SG_LAUNCH(aug::mask_fill_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        x, ldx, rows, (int)(feat / 4), row_keep, col_keep, p_elem, seed, fill, out, ldo, mask_out));
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}

extern "C" int stemgnn_gather_rows(const float* src, int64_t lds, int64_t n_src, const int64_t* ids,
                                   int64_t n_out, int64_t feat, float* out, int64_t ldo,
                                   stemgnn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SG_CHECK_ARG(n_out >= 0 && feat >= 0 && n_src >= 0, "negative size");
    if (n_out == 0 || feat == 0) return STEMGNN_OK;
    SG_CHECK_ARG(src && ids && out, "NULL pointer");
    SG_CHECK_ARG(feat % 4 == 0 && lds % 4 == 0 && ldo % 4 == 0, "feat, lds, ldo must be multiples of 4");
    SG_CHECK_ARG(((uintptr_t)src | (uintptr_t)out) % 16 == 0, "src, out must be 16-byte aligned");
    const int f4 = (int)(feat / 4);
    if (f4 <= 8) {
        const int64_t thr = n_out * 8;
        SG_LAUNCH(aug::gather_rows_kernel<8><<<(unsigned)((thr + 255) / 256), 256, 0, stream>>>(
            src, lds, n_src, ids, n_out, f4, out, ldo));
    } else {
        const int64_t thr = n_out * 32;
        SG_LAUNCH(aug::gather_rows_kernel<32><<<(unsigned)((thr + 255) / 256), 256, 0, stream>>>(
            src, lds, n_src, ids, n_out, f4, out, ldo));
    }
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}
