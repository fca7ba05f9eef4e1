// tcgen05 / TMEM / TMA GEMM backend (placeholder until the kernels land: every entry reports
// "shape not taken", so VNB_BACKEND_AUTO uses the SIMT backend and VNB_BACKEND_TC errors out).
#include "common.cuh"

namespace vnb {

int conv_fwd_tc(const void*, int, int, int, int, int, const void*, int, int, int, const float*, int, const float*,
                float*, int, cudaStream_t) {
    return 1;
}
int conv_wgrad_tc(const void*, const void*, int, int, int, int, int, int, int, int, int, float*, cudaStream_t) {
    return 1;
}
int vq_lookup_tc(const void*, const float*, const void*, const float*, int, int, int, int, int, int64_t*,
                 cudaStream_t) {
    return 1;
}

}  // namespace vnb
