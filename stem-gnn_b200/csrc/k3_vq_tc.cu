// k3_vq_tc.cu -- K3 path 2 placeholder while the tcgen05 kernel is brought up: reports the
// tensor-core path as unsupported so path 0 (auto) resolves to the exact fp32 scan.
#include "k3_vq_common.cuh"

namespace stemgnn {
bool vq_tc_supported(const VqView&) { return false; }
int vq_assign_tc_workspace(const VqView&, size_t* bytes) { *bytes = 0; return STEMGNN_EUNSUPPORTED; }
int vq_assign_tc(const VqView&, const float*, const float*, int32_t*, int64_t*, void*, size_t,
                 cudaStream_t) {
    set_error("vq_assign_tc: tcgen05 path not built");
    return STEMGNN_EUNSUPPORTED;
}
}  // namespace stemgnn
