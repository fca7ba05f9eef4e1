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

// ---- batched variants over the trainer's flat parameter / gradient buffers -------------------
// desc: int64[n_layers][WN_DESC] = {v_off, g_off (-1: plain weight), d0, d1, k, transposed, wf_off, wb_off,
// w32_off (-1: not wanted), row0};  row2layer: int32[total_rows].  One block per weight row of ANY layer.
constexpr int WN_DESC = 10;

__global__ void __launch_bounds__(256) wn_prepare_batch_kernel(const float* __restrict__ flat_p,
                                                               const int64_t* __restrict__ desc,
                                                               const int* __restrict__ row2layer, int P,
                                                               bf16* __restrict__ wbuf, float* __restrict__ w32buf) {
    __shared__ float red[32];
    const int64_t* L = desc + (int64_t)row2layer[blockIdx.x] * WN_DESC;
    const int i = blockIdx.x - (int)L[9];
    const int d0 = (int)L[2], d1 = (int)L[3], k = (int)L[4], transposed = (int)L[5];
    const int n = d1 * k;
    const float* row = flat_p + L[0] + (int64_t)i * n;
    float scale = 1.f;
    if (L[1] >= 0) {
        float ss = 0.f;
        for (int e = threadIdx.x; e < n; e += blockDim.x) {
            const float x = row[e];
            ss += x * x;
        }
        ss = block_sum(ss, red);
        scale = flat_p[L[1] + i] / sqrtf(ss);
    }
    const int co_n = transposed ? d1 : d0, ci_n = transposed ? d0 : d1;
    const int ci_p = (ci_n + 7) & ~7, co_p = (co_n + 7) & ~7;
    const int64_t wf_plane = (int64_t)k * co_n * ci_p, wb_plane = (int64_t)k * ci_n * co_p;
    bf16* wf = L[6] >= 0 ? wbuf + L[6] : nullptr;
    bf16* wb = L[7] >= 0 ? wbuf + L[7] : nullptr;
    float* w32 = L[8] >= 0 ? w32buf + L[8] : nullptr;
    for (int e = threadIdx.x; e < n; e += blockDim.x) {
        const int j1 = e / k, j = e - j1 * k;
        const float w = row[e] * scale;
        if (w32) w32[(int64_t)i * n + e] = w;
        const int co = transposed ? j1 : i, ci = transposed ? i : j1;
        const int tap = transposed ? (k - 1 - j) : j;
        if (wf) store_planes(wf, ((int64_t)tap * co_n + co) * ci_p + ci, wf_plane, P, w);
        if (wb) store_planes(wb, ((int64_t)(k - 1 - tap) * ci_n + ci) * co_p + co, wb_plane, P, w);
    }
}

// in place on flat_g: the v slot holds dw on entry, dv on exit; the g slot receives dg
__global__ void __launch_bounds__(256) wn_backward_batch_kernel(const float* __restrict__ flat_p,
                                                                float* __restrict__ flat_g,
                                                                const int64_t* __restrict__ desc,
                                                                const int* __restrict__ row2layer) {
    __shared__ float red[32];
    const int64_t* L = desc + (int64_t)row2layer[blockIdx.x] * WN_DESC;
    if (L[1] < 0) return;                       // plain weight: dw already is its gradient
    const int i = blockIdx.x - (int)L[9];
    const int n = (int)L[3] * (int)L[4];
    const float* vr = flat_p + L[0] + (int64_t)i * n;
    float* dr = flat_g + L[0] + (int64_t)i * n;
    float ss = 0.f, dot = 0.f;
    for (int e = threadIdx.x; e < n; e += blockDim.x) {
        const float x = vr[e];
        ss += x * x;
        dot += x * dr[e];
    }
    ss = block_sum(ss, red);
    dot = block_sum(dot, red);
    const float norm = sqrtf(ss);
    const float gi = flat_p[L[1] + i];
    if (threadIdx.x == 0) flat_g[L[1] + i] = dot / norm;
    const float a = gi / norm, b = gi * dot / (norm * ss);
    for (int e = threadIdx.x; e < n; e += blockDim.x) dr[e] = a * dr[e] - b * vr[e];
}

}  // namespace vnb

extern "C" int vnb_wn_prepare_batch(const float* flat_p, const int64_t* desc, const int* row2layer, int total_rows,
                                    int P, void* wbuf, float* w32buf, void* stream) {
    VNB_REQUIRE(flat_p && desc && row2layer && wbuf && total_rows > 0 && (P == 1 || P == 3), "wn_prepare_batch: bad arguments");
    vnb::wn_prepare_batch_kernel<<<total_rows, 256, 0, (cudaStream_t)stream>>>(flat_p, desc, row2layer, P,
                                                                               (vnb::bf16*)wbuf, w32buf);
    VNB_LAUNCH_CHECK("wn_prepare_batch");
    return VNB_OK;
}

extern "C" int vnb_wn_backward_batch(const float* flat_p, float* flat_g, const int64_t* desc, const int* row2layer,
                                     int total_rows, void* stream) {
    VNB_REQUIRE(flat_p && flat_g && desc && row2layer && total_rows > 0, "wn_backward_batch: bad arguments");
    vnb::wn_backward_batch_kernel<<<total_rows, 256, 0, (cudaStream_t)stream>>>(flat_p, flat_g, desc, row2layer);
    VNB_LAUNCH_CHECK("wn_backward_batch");
    return VNB_OK;
}

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
