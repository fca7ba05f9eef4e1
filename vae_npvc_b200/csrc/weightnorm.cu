// Weight-norm materialisation (g, v) -> GEMM operand planes, and its backward.
// Replaces the nn.utils.weight_norm pre-hook the reference installs on all 55 conv modules
// (vae_npvc/model/vqvae.py:203-208, 329-334).  SURVEY.md section 9 G; F8 for ConvTranspose.
#include "common.cuh"

namespace vnb {

// One block per index of dim 0.  v row = d1*k contiguous floats.
__global__ void __launch_bounds__(256) wn_prepare_kernel(const float* __restrict__ v, const float* __restrict__ g,
                                                         int d0, int d1, int k, int transposed, int P,
                                                         bf16* __restrict__ wf, bf16* __restrict__ wb,
                                                         float* __restrict__ w32) {
    __shared__ float red[32];
    const int i = blockIdx.x;
    const int n = d1 * k;
    const float* row = v + (int64_t)i * n;
    float scale = 1.f;
    if (g != nullptr) {
        float ss = 0.f;
        for (int e = threadIdx.x; e < n; e += blockDim.x) {
            float x = row[e];
            ss += x * x;
        }
        ss = block_sum(ss, red);
        scale = g[i] / sqrtf(ss);
    }
    const int co_n = transposed ? d1 : d0;
    const int ci_n = transposed ? d0 : d1;
    const int ci_p = (ci_n + 7) & ~7, co_p = (co_n + 7) & ~7;
    const int64_t wf_plane = (int64_t)k * co_n * ci_p;
    const int64_t wb_plane = (int64_t)k * ci_n * co_p;
    for (int e = threadIdx.x; e < n; e += blockDim.x) {
        const int j1 = e / k, j = e - j1 * k;
        const float w = row[e] * scale;
        if (w32) w32[(int64_t)i * n + e] = w;
        const int co = transposed ? j1 : i;
        const int ci = transposed ? i : j1;
        const int tap = transposed ? (k - 1 - j) : j;      // forward conv tap
        if (wf) store_planes(wf, ((int64_t)tap * co_n + co) * ci_p + ci, wf_plane, P, w);
        if (wb) store_planes(wb, ((int64_t)(k - 1 - tap) * ci_n + ci) * co_p + co, wb_plane, P, w);
    }
}

__global__ void __launch_bounds__(256) wn_backward_kernel(const float* __restrict__ dw, const float* __restrict__ v,
                                                          const float* __restrict__ g, int n,
                                                          float* __restrict__ dv, float* __restrict__ dg) {
    __shared__ float red[32];
    const int i = blockIdx.x;
    const float* vr = v + (int64_t)i * n;
    const float* dr = dw + (int64_t)i * n;
    float ss = 0.f, dot = 0.f;
    for (int e = threadIdx.x; e < n; e += blockDim.x) {
        float x = vr[e];
        ss += x * x;
        dot += x * dr[e];
    }
    ss = block_sum(ss, red);
    dot = block_sum(dot, red);
    const float norm = sqrtf(ss);
    const float gi = g[i];
    if (threadIdx.x == 0) dg[i] = dot / norm;
    const float a = gi / norm, b = gi * dot / (norm * ss);
    for (int e = threadIdx.x; e < n; e += blockDim.x) dv[(int64_t)i * n + e] = a * dr[e] - b * vr[e];
}

}  // namespace vnb

extern "C" int vnb_wn_prepare(const float* v, const float* g, int d0, int d1, int k, int transposed, int P,
                              void* wf, void* wb, float* w32, void* stream) {
    VNB_REQUIRE(v && d0 > 0 && d1 > 0 && k > 0 && (P == 1 || P == 3), "wn_prepare: bad arguments");
    vnb::wn_prepare_kernel<<<d0, 256, 0, (cudaStream_t)stream>>>(v, g, d0, d1, k, transposed, P, (vnb::bf16*)wf,
                                                                (vnb::bf16*)wb, w32);
    VNB_LAUNCH_CHECK("wn_prepare");
    return VNB_OK;
}

extern "C" int vnb_wn_backward(const float* dw, const float* v, const float* g, int d0, int d1, int k, float* dv,
                               float* dg, void* stream) {
    VNB_REQUIRE(dw && v && g && dv && dg && d0 > 0, "wn_backward: bad arguments");
    vnb::wn_backward_kernel<<<d0, 256, 0, (cudaStream_t)stream>>>(dw, v, g, d1 * k, dv, dg);
    VNB_LAUNCH_CHECK("wn_backward");
    return VNB_OK;
}
