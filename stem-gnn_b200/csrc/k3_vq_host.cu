// k3_vq_host.cu -- C-ABI entry points of the VQ assignment (K3): argument checks, workspace
// carving and the kernel sequence  norms -> {exact fp32 scan | tcgen05 GEMM + rescoring} ->
// gather + commitment partials -> fixed-order reduction.  Replaces model/vq.py:472-482,
// 650-659, 79-80, 1007 of the reference.
#include "k3_vq_common.cuh"

using namespace stemgnn;

namespace {
struct VqPlan {
    int64_t items;   // heads * rows
    int64_t nblk;    // gather warps (entries of VQ_GATHER_ROWS rows) per head
    int path;
};

int make_view(const stemgnn_vq_assign_args* a, VqView& q, VqPlan& p) {
    SG_CHECK_ARG(a != nullptr, "args is NULL");
    SG_CHECK_ARG(a->heads >= 1 && a->rows >= 0 && a->codes >= 1 && a->dim >= 1, "bad sizes");
    SG_CHECK_ARG(a->embed_heads == 1 || a->embed_heads == a->heads, "embed_heads must be 1 or heads");
    SG_CHECK_ARG(a->metric == 0 || a->metric == 1, "metric must be 0 or 1");
    SG_CHECK_ARG(a->path >= 0 && a->path <= 2, "path must be 0, 1 or 2");
    SG_CHECK_ARG(a->codes < (1 << 24) && a->dim < (1 << 20), "codes/dim too large");
    SG_CHECK_ARG(a->heads * a->rows < 2147483647, "heads*rows must be < 2^31");
    q.x = a->x; q.x_row_stride = a->x_row_stride; q.x_head_stride = a->x_head_stride;
    q.embed = a->embed;
    q.heads = a->heads; q.embed_heads = a->embed_heads; q.rows = a->rows;
    q.K = (int)a->codes; q.D = (int)a->dim; q.metric = a->metric;
    q.ind = a->ind; q.ind_row_stride = a->ind_row_stride; q.ind_head_stride = a->ind_head_stride;
    q.quantize = a->quantize; q.q_row_stride = a->q_row_stride; q.q_head_stride = a->q_head_stride;
    p.items = a->heads * a->rows;
    p.nblk = (a->rows + VQ_GATHER_ROWS - 1) / VQ_GATHER_ROWS;
    p.path = a->path;
    if (p.path == 0) p.path = vq_tc_supported(q) ? 2 : 1;
    if (p.path == 2 && !vq_tc_supported(q)) {
        set_error("stemgnn_vq_assign: path 2 (tcgen05) needs dim %% 4 == 0, dim >= 16 and 16-byte aligned rows");
        return STEMGNN_EUNSUPPORTED;
    }
    return STEMGNN_OK;
}

size_t base_bytes(const VqView& q, const VqPlan& p) {
    size_t b = 0;
    b += align_up((size_t)p.items * 4, 256);                         // x2
    b += align_up((size_t)q.embed_heads * q.K * 4, 256);             // e2
    b += align_up((size_t)p.items * 4, 256);                         // best
    b += align_up((size_t)q.heads * (size_t)(p.nblk > 0 ? p.nblk : 1) * 8, 256);  // partials
    b += align_up((size_t)q.heads * (size_t)(p.nblk > 0 ? p.nblk : 1) * 4, 256);  // open_list
    b += align_up((size_t)q.heads * VQ_SQ_CHUNKS * 8, 256);                        // commitment chunk sums
    b += 256;                                                                      // open_count
    return b;
}
}  // namespace

extern "C" int stemgnn_vq_assign_workspace(const stemgnn_vq_assign_args* a, size_t* bytes) {
    SG_CHECK_ARG(bytes != nullptr, "bytes_host is NULL");
    VqView q; VqPlan p;
    int rc = make_view(a, q, p);
    if (rc) return rc;
    size_t b = base_bytes(q, p);
    if (p.path == 2) {
        size_t t = 0;
        rc = vq_assign_tc_workspace(q, &t);
        if (rc) return rc;
        b += t;
    }
    *bytes = b;
    return STEMGNN_OK;
}

extern "C" int stemgnn_vq_assign(const stemgnn_vq_assign_args* a, void* workspace,
                                 size_t workspace_bytes, stemgnn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    VqView q; VqPlan p;
    int rc = make_view(a, q, p);
    if (rc) return rc;
    if (a->stats) cudaMemsetAsync(a->stats, 0, 4 * sizeof(int64_t), stream);
    if (p.items == 0) {
        if (a->sq_err) cudaMemsetAsync(a->sq_err, 0, (size_t)a->heads * 8, stream);
        SG_CHECK_LAUNCH();
        return STEMGNN_OK;
    }
    SG_CHECK_ARG(a->x && a->embed, "x/embed is NULL");
    size_t need = 0;
    rc = stemgnn_vq_assign_workspace(a, &need);
    if (rc) return rc;
    if (workspace == nullptr || workspace_bytes < need) {
        set_error("stemgnn_vq_assign: workspace too small (%zu < %zu)", workspace_bytes, need);
        return STEMGNN_ENOMEM_WS;
    }
    Workspace ws(workspace, workspace_bytes);
    float* x2 = ws.take<float>(p.items);
    float* e2 = ws.take<float>((size_t)q.embed_heads * q.K);
    int32_t* best = ws.take<int32_t>(p.items);
    double* partials = ws.take<double>((size_t)q.heads * p.nblk);
    int32_t* open_list = ws.take<int32_t>((size_t)q.heads * p.nblk);
    double* chunk_sums = ws.take<double>((size_t)q.heads * VQ_SQ_CHUNKS);
    int32_t* open_count = ws.take<int32_t>(1);
    cudaMemsetAsync(open_count, 0, 4, stream);

    const int64_t ek = q.embed_heads * (int64_t)q.K;
    SG_LAUNCH(vq_sqnorm_kernel<<<(unsigned)((ek * 32 + 255) / 256), 256, 0, stream>>>(
        q.embed, ek, 1, q.D, q.D, 0, e2));
    int32_t* cand = nullptr;
    if (p.path == 1) {
        if (q.metric == 0)
            SG_LAUNCH(vq_sqnorm_kernel<<<(unsigned)((p.items * 32 + 255) / 256), 256, 0, stream>>>(
                q.x, q.rows, q.heads, q.D, q.x_row_stride, q.x_head_stride, x2));
        dim3 grid((unsigned)((q.rows + 63) / 64), (unsigned)q.heads);
        SG_LAUNCH(vq_scan_kernel<<<grid, 256, 0, stream>>>(q, x2, e2, nullptr, nullptr, best));
    } else {
        rc = vq_assign_tc(q, e2, best, &cand, ws.base + ws.off, workspace_bytes - ws.off, stream);
        if (rc) return rc;
    }
    dim3 ggrid((unsigned)((p.nblk + 7) / 8), (unsigned)q.heads);      // one warp per entry of VQ_GATHER_ROWS rows
    SG_LAUNCH(vq_gather_kernel<<<ggrid, 256, 0, stream>>>(
        q, best, cand, e2, a->sq_err ? partials : nullptr, open_list, open_count,
        reinterpret_cast<unsigned long long*>(a->stats), l2_pingpong() ? 1 : 0));
    if (p.path == 2)      // rows the tensor-core search left to the exact scan (normally a handful)
        SG_LAUNCH(vq_rescan_kernel<<<(unsigned)sm_count(), 1024, 0, stream>>>(
            q, best, e2, a->sq_err ? partials : nullptr, (int)p.nblk, open_list, open_count,
            reinterpret_cast<unsigned long long*>(a->stats)));
    if (a->sq_err) {
        SG_LAUNCH(vq_sqerr_chunk_kernel<<<dim3(VQ_SQ_CHUNKS, (unsigned)q.heads), 256, 0, stream>>>(partials, (int)p.nblk, chunk_sums));
        SG_LAUNCH(vq_sqerr_reduce_kernel<<<(unsigned)q.heads, 256, 0, stream>>>(chunk_sums, VQ_SQ_CHUNKS, a->sq_err));
    }
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}
