// k8_edge_dot.cu -- K8: the edge-wise inner product of InnerProductDecoder.
//
// Replaces `value = (z[edge_index[0]] * z[edge_index[1]]).sum(dim=1)` (+ sigmoid) of the
// reference (STEM-GNN/model/encoder.py:363-366, called from model/pt_model.py:62-63 on the
// positive and the sampled negative edges) without materialising the two E x F gathered
// matrices and their product: a sub-warp of F/4 (<= 32) lanes owns an edge, reads both rows
// with 128-bit loads and reduces with a fixed xor-shuffle tree (deterministic).  HBM/L2-bound
// gather: 2*F*4 B per edge, E*4 B out, E*16 B of indices.
// The backward is two weighted aggregations (K1, edge weights = upstream gradient) on the host
// side (stemgnn_b200/decoder.py); there is no separate backward kernel.
#include "common.cuh"

namespace stemgnn {
namespace edot {

template <int LANES>
__global__ void __launch_bounds__(256)
edge_dot_kernel(const int64_t* __restrict__ src, const int64_t* __restrict__ dst, int64_t E, int64_t N,
                int F, const float* __restrict__ z, int64_t ldz, int sigmoid, float* __restrict__ out) {
    constexpr int EPW = 32 / LANES;                  // edges per warp and trip
    const int lane = threadIdx.x & 31, sub = lane % LANES, slot = lane / LANES;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const bool vec4 = (F & 3) == 0 && (ldz & 3) == 0 && (reinterpret_cast<uintptr_t>(z) & 15) == 0;
    // two edges per sub-warp in flight: trip t covers edges [t*2*EPW, (t+1)*2*EPW) of this warp's stride
    for (int64_t base = warp * (2 * EPW); base < E; base += n_warps * (2 * EPW)) {
        float acc[2];
        int64_t e[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            e[u] = base + u * EPW + slot;
            acc[u] = 0.f;
        }
        const float* pa[2];
        const float* pb[2];
        bool ok[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            ok[u] = false;
            pa[u] = z; pb[u] = z;
            if (e[u] < E) {
                const int64_t s = __ldg(src + e[u]), d = __ldg(dst + e[u]);
                ok[u] = (uint64_t)s < (uint64_t)N && (uint64_t)d < (uint64_t)N;
                if (ok[u]) { pa[u] = z + s * ldz; pb[u] = z + d * ldz; }
            }
        }
        if (vec4) {
            for (int c = sub * 4; c < F; c += LANES * 4) {
                float4 a[2], b[2];
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    a[u] = make_float4(0.f, 0.f, 0.f, 0.f); b[u] = a[u];
                    if (ok[u]) {
                        a[u] = __ldg(reinterpret_cast<const float4*>(pa[u] + c));
                        b[u] = __ldg(reinterpret_cast<const float4*>(pb[u] + c));
                    }
                }
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    acc[u] = fmaf(a[u].x, b[u].x, acc[u]); acc[u] = fmaf(a[u].y, b[u].y, acc[u]);
                    acc[u] = fmaf(a[u].z, b[u].z, acc[u]); acc[u] = fmaf(a[u].w, b[u].w, acc[u]);
                }
            }
        } else {
            for (int c = sub; c < F; c += LANES) {
#pragma unroll
                for (int u = 0; u < 2; ++u)
                    if (ok[u]) acc[u] = fmaf(__ldg(pa[u] + c), __ldg(pb[u] + c), acc[u]);
            }
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
#pragma unroll
            for (int o = LANES / 2; o > 0; o >>= 1) acc[u] += __shfl_xor_sync(0xffffffffu, acc[u], o);
            if (sub == 0 && e[u] < E) {
                float v = ok[u] ? acc[u] : __int_as_float(0x7fc00000);   // index out of range: NaN
                if (sigmoid) v = 1.f / (1.f + expf(-v));
                out[e[u]] = v;
            }
        }
    }
}

}  // namespace edot
}  // namespace stemgnn

using namespace stemgnn;

extern "C" int stemgnn_edge_dot(const int64_t* edge_index, int64_t num_edges, int64_t num_nodes, int64_t feat,
                                const float* z, int64_t ldz, int apply_sigmoid, float* out,
                                stemgnn_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SG_CHECK_ARG(num_edges >= 0 && num_nodes >= 0 && feat >= 1 && feat < (1 << 20), "bad sizes");
    if (num_edges == 0) return STEMGNN_OK;
    SG_CHECK_ARG(edge_index && z && out, "NULL pointer");
    SG_CHECK_ARG(ldz >= feat, "ldz < feat");
    const int64_t* src = edge_index;
    const int64_t* dst = edge_index + num_edges;
    const int lanes = feat >= 128 ? 32 : (feat >= 64 ? 16 : (feat >= 32 ? 8 : 4));
    const int64_t epb = (int64_t)(256 / lanes) * 2;               // edges per block and trip
    int64_t blocks = (num_edges + epb - 1) / epb;
    const int64_t cap = (int64_t)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    const int F = (int)feat;
    switch (lanes) {
        case 32: SG_LAUNCH(edot::edge_dot_kernel<32><<<(unsigned)blocks, 256, 0, stream>>>(src, dst, num_edges, num_nodes, F, z, ldz, apply_sigmoid, out)); break;
        case 16: SG_LAUNCH(edot::edge_dot_kernel<16><<<(unsigned)blocks, 256, 0, stream>>>(src, dst, num_edges, num_nodes, F, z, ldz, apply_sigmoid, out)); break;
        case 8: SG_LAUNCH(edot::edge_dot_kernel<8><<<(unsigned)blocks, 256, 0, stream>>>(src, dst, num_edges, num_nodes, F, z, ldz, apply_sigmoid, out)); break;
        default: SG_LAUNCH(edot::edge_dot_kernel<4><<<(unsigned)blocks, 256, 0, stream>>>(src, dst, num_edges, num_nodes, F, z, ldz, apply_sigmoid, out)); break;
    }
    SG_CHECK_LAUNCH();
    return STEMGNN_OK;
}
