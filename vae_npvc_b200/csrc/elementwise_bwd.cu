// Backward-side memory-bound kernels: GroupNorm / gate gradient row statistics, their
// finalisation, gradient application into operand planes, activation masks, bias sums.
// Math: SURVEY.md section 9 A/B (restated from vae_npvc/model/layers.py:151-176, 229-247).
#include "common.cuh"

namespace vnb {

constexpr int EW_THREADS = 256;
static inline int ew_grid(int64_t n) {
    int64_t g = (n + EW_THREADS - 1) / EW_THREADS;
    const int64_t cap = 148 * 16;
    return (int)(g < cap ? g : cap);
}
#define GRID_STRIDE(i, n) \
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < (n); i += (int64_t)gridDim.x * blockDim.x)

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }

// gradient wrt the normalised+affine gate input row c2 in [0,2C) at (b,t), recomputed from dg
__device__ __forceinline__ float gate_dn(const float* __restrict__ dg, const float* __restrict__ a,
                                         const float* __restrict__ mean, const float* __restrict__ rstd,
                                         const float* __restrict__ gamma, const float* __restrict__ beta, int b, int c2,
                                         int t, int C, int T, float& xhat) {
    const int c = c2 < C ? c2 : c2 - C;
    const int64_t i1 = ((int64_t)b * 2 * C + c) * T + t, i2 = i1 + (int64_t)C * T;
    const float x1 = (a[i1] - mean[2 * b]) * rstd[2 * b];
    const float x2 = (a[i2] - mean[2 * b + 1]) * rstd[2 * b + 1];
    const float th = tanhf(x1 * gamma[c] + beta[c]);
    const float sg = sigmoidf_(x2 * gamma[C + c] + beta[C + c]);
    const float d = dg[((int64_t)b * C + c) * T + t];
    if (c2 < C) {
        xhat = x1;
        return d * sg * (1.f - th * th);
    }
    xhat = x2;
    return d * th * sg * (1.f - sg);
}

// one warp per (b,c) row: R[row] = { sum_t dn, sum_t dn*xhat, sum_t xhat }
__global__ void __launch_bounds__(256) gn_bwd_rows_kernel(const float* __restrict__ dn, const float* __restrict__ h,
                                                          const float* __restrict__ mean,
                                                          const float* __restrict__ rstd, int B, int C, int T,
                                                          int groups, float* __restrict__ R) {
    const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= (int64_t)B * C) return;
    const int lane = threadIdx.x & 31;
    const int b = (int)(row / C), c = (int)(row - (int64_t)b * C);
    const int gi = b * groups + c / (C / groups);
    const float m = mean[gi], rs = rstd[gi];
    float s0 = 0.f, s1 = 0.f, s2 = 0.f;
    for (int t = lane; t < T; t += 32) {
        const float d = dn[row * T + t];
        const float xh = (h[row * T + t] - m) * rs;
        s0 += d;
        s1 += d * xh;
        s2 += xh;
    }
    s0 = warp_sum(s0);
    s1 = warp_sum(s1);
    s2 = warp_sum(s2);
    if (lane == 0) {
        R[row * 3 + 0] = s0;
        R[row * 3 + 1] = s1;
        R[row * 3 + 2] = s2;
    }
}

__global__ void __launch_bounds__(256) gate_bwd_rows_kernel(const float* __restrict__ dg, const float* __restrict__ a,
                                                            const float* __restrict__ mean,
                                                            const float* __restrict__ rstd,
                                                            const float* __restrict__ gamma,
                                                            const float* __restrict__ beta, int B, int C, int T,
                                                            float* __restrict__ R) {
    const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= (int64_t)B * 2 * C) return;
    const int lane = threadIdx.x & 31;
    const int b = (int)(row / (2 * C)), c2 = (int)(row - (int64_t)b * 2 * C);
    float s0 = 0.f, s1 = 0.f, s2 = 0.f;
    for (int t = lane; t < T; t += 32) {
        float xh;
        const float d = gate_dn(dg, a, mean, rstd, gamma, beta, b, c2, t, C, T, xh);
        s0 += d;
        s1 += d * xh;
        s2 += xh;
    }
    s0 = warp_sum(s0);
    s1 = warp_sum(s1);
    s2 = warp_sum(s2);
    if (lane == 0) {
        R[row * 3 + 0] = s0;
        R[row * 3 + 1] = s1;
        R[row * 3 + 2] = s2;
    }
}

// S[b,g,0..1] = sum_{c in g} gamma_c * R{0,1}[b,c]     (one block per (b,g))
__global__ void __launch_bounds__(256) gn_bwd_groupsum_kernel(const float* __restrict__ R,
                                                              const float* __restrict__ gamma, int C, int groups,
                                                              float* __restrict__ S) {
    __shared__ float red[32];
    const int b = blockIdx.x / groups, g = blockIdx.x % groups;
    const int cg = C / groups;
    float s0 = 0.f, s1 = 0.f;
    for (int j = threadIdx.x; j < cg; j += blockDim.x) {
        const int c = g * cg + j;
        const float gm = gamma[c];
        const float* r = R + ((int64_t)b * C + c) * 3;
        s0 += gm * r[0];
        s1 += gm * r[1];
    }
    s0 = block_sum(s0, red);
    s1 = block_sum(s1, red);
    if (threadIdx.x == 0) {
        S[blockIdx.x * 2 + 0] = s0;
        S[blockIdx.x * 2 + 1] = s1;
    }
}

// per channel: dgamma, dbeta, and the time/batch sums of the pre-norm gradient
__global__ void gn_bwd_channel_kernel(const float* __restrict__ R, const float* __restrict__ gamma,
                                      const float* __restrict__ rstd, const float* __restrict__ S, int B, int C, int T,
                                      int groups, float* __restrict__ dgamma, float* __restrict__ dbeta,
                                      float* __restrict__ dpre_bsum, float* __restrict__ dpre_csum) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const int cg = C / groups, g = c / cg;
    const float inv_n = 1.f / ((float)cg * (float)T);
    const float gm = gamma[c];
    float dgm = 0.f, dbt = 0.f, dcs = 0.f;
    for (int b = 0; b < B; b++) {
        const float* r = R + ((int64_t)b * C + c) * 3;
        dbt += r[0];
        dgm += r[1];
        if (dpre_bsum || dpre_csum) {
            const float* s = S + (b * groups + g) * 2;
            const float v = rstd[b * groups + g] * (gm * r[0] - (float)T * s[0] * inv_n - r[2] * s[1] * inv_n);
            if (dpre_bsum) dpre_bsum[(int64_t)b * C + c] = v;
            dcs += v;
        }
    }
    if (dgamma) dgamma[c] = dgm;
    if (dbeta) dbeta[c] = dbt;
    if (dpre_csum) dpre_csum[c] = dcs;
}

__global__ void gn_bwd_apply_kernel(const float* __restrict__ dn, const float* __restrict__ h,
                                    const float* __restrict__ mean, const float* __restrict__ rstd,
                                    const float* __restrict__ gamma, const float* __restrict__ S, int B, int C, int T,
                                    int groups, int Tp, int P, bf16* __restrict__ dpre_planes,
                                    bf16* __restrict__ dn_planes) {
    const int64_t n = (int64_t)B * C * T, plane = (int64_t)B * C * Tp;
    const int cg = C / groups;
    const float inv_n = 1.f / ((float)cg * (float)T);
    GRID_STRIDE(i, n) {
        const int64_t r = i / T;
        const int t = (int)(i - r * T);
        const int b = (int)(r / C), c = (int)(r - (int64_t)b * C);
        const int gi = b * groups + c / cg;
        const float rs = rstd[gi];
        const float xh = (h[i] - mean[gi]) * rs;
        const float d = dn[i];
        const float v = rs * (gamma[c] * d - S[gi * 2] * inv_n - xh * S[gi * 2 + 1] * inv_n);
        const int64_t o = r * Tp + t;
        store_planes(dpre_planes, o, plane, P, v);
        if (dn_planes) store_planes(dn_planes, o, plane, P, d);
    }
}

__global__ void gate_bwd_apply_kernel(const float* __restrict__ dg, const float* __restrict__ a,
                                      const float* __restrict__ mean, const float* __restrict__ rstd,
                                      const float* __restrict__ gamma, const float* __restrict__ beta,
                                      const float* __restrict__ S, int B, int C, int T, int Tp, int P,
                                      bf16* __restrict__ da_planes) {
    const int C2 = 2 * C;
    const int64_t n = (int64_t)B * C2 * T, plane = (int64_t)B * C2 * Tp;
    const float inv_n = 1.f / ((float)C * (float)T);
    GRID_STRIDE(i, n) {
        const int64_t r = i / T;
        const int t = (int)(i - r * T);
        const int b = (int)(r / C2), c2 = (int)(r - (int64_t)b * C2);
        const int gi = b * 2 + (c2 >= C ? 1 : 0);
        float xh;
        const float d = gate_dn(dg, a, mean, rstd, gamma, beta, b, c2, t, C, T, xh);
        const float v = rstd[gi] * (gamma[c2] * d - S[gi * 2] * inv_n - xh * S[gi * 2 + 1] * inv_n);
        store_planes(da_planes, r * Tp + t, plane, P, v);
    }
}

__global__ void act_bwd_add_kernel(const float* __restrict__ du, const float* __restrict__ c,
                                   const float* __restrict__ add, int64_t n, int act, float scale,
                                   float* __restrict__ out) {
    GRID_STRIDE(i, n) {
        float v = scale * du[i] * act_grad(c[i] * scale, act);
        if (add) v += add[i];
        out[i] = v;
    }
}

// one block per channel
__global__ void __launch_bounds__(256) channel_sum_kernel(const float* __restrict__ x, int B, int C, int T,
                                                          float* __restrict__ out) {
    __shared__ float red[32];
    const int c = blockIdx.x;
    float s = 0.f;
    for (int b = 0; b < B; b++) {
        const float* p = x + ((int64_t)b * C + c) * T;
        for (int t = threadIdx.x; t < T; t += blockDim.x) s += p[t];
    }
    s = block_sum(s, red);
    if (threadIdx.x == 0) out[c] = s;
}

}  // namespace vnb
using namespace vnb;

extern "C" int vnb_gn_bwd_rows(const float* dn, const float* h, const float* mean, const float* rstd, int B, int C,
                               int T, int groups, float* R, void* stream) {
    VNB_REQUIRE(dn && h && mean && rstd && R && B > 0 && C > 0 && T > 0 && groups > 0 && C % groups == 0, "gn_bwd_rows: bad arguments");
    gn_bwd_rows_kernel<<<ceil_div((int64_t)B * C, 8), 256, 0, (cudaStream_t)stream>>>(dn, h, mean, rstd, B, C, T, groups, R);
    VNB_LAUNCH_CHECK("gn_bwd_rows");
    return VNB_OK;
}

extern "C" int vnb_gate_bwd_rows(const float* dg, const float* a, const float* mean, const float* rstd,
                                 const float* gamma, const float* beta, int B, int C, int T, float* R, void* stream) {
    VNB_REQUIRE(dg && a && mean && rstd && gamma && beta && R && B > 0 && C > 0 && T > 0, "gate_bwd_rows: bad arguments");
    gate_bwd_rows_kernel<<<ceil_div((int64_t)B * 2 * C, 8), 256, 0, (cudaStream_t)stream>>>(dg, a, mean, rstd, gamma, beta, B, C, T, R);
    VNB_LAUNCH_CHECK("gate_bwd_rows");
    return VNB_OK;
}

extern "C" int vnb_gn_bwd_finalize(const float* R, const float* gamma, const float* rstd, int B, int C, int T,
                                   int groups, float* dgamma, float* dbeta, float* S, float* dpre_bsum,
                                   float* dpre_csum, void* stream) {
    VNB_REQUIRE(R && gamma && rstd && S && B > 0 && C > 0 && groups > 0 && C % groups == 0, "gn_bwd_finalize: bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    gn_bwd_groupsum_kernel<<<B * groups, 256, 0, st>>>(R, gamma, C, groups, S);
    VNB_LAUNCH_CHECK("gn_bwd_groupsum");
    gn_bwd_channel_kernel<<<ceil_div(C, 128), 128, 0, st>>>(R, gamma, rstd, S, B, C, T, groups, dgamma, dbeta, dpre_bsum, dpre_csum);
    VNB_LAUNCH_CHECK("gn_bwd_channel");
    return VNB_OK;
}

extern "C" int vnb_gn_bwd_apply(const float* dn, const float* h, const float* mean, const float* rstd,
                                const float* gamma, const float* S, int B, int C, int T, int groups, int Tp, int P,
                                void* dpre_planes, void* dn_planes, void* stream) {
    VNB_REQUIRE(dn && h && mean && rstd && gamma && S && dpre_planes, "gn_bwd_apply: null pointer");
    VNB_REQUIRE(B > 0 && C > 0 && T > 0 && Tp >= T && groups > 0 && C % groups == 0 && (P == 1 || P == 3), "gn_bwd_apply: bad shape");
    gn_bwd_apply_kernel<<<ew_grid((int64_t)B * C * T), EW_THREADS, 0, (cudaStream_t)stream>>>(
        dn, h, mean, rstd, gamma, S, B, C, T, groups, Tp, P, (bf16*)dpre_planes, (bf16*)dn_planes);
    VNB_LAUNCH_CHECK("gn_bwd_apply");
    return VNB_OK;
}

extern "C" int vnb_gate_bwd_apply(const float* dg, const float* a, const float* mean, const float* rstd,
                                  const float* gamma, const float* beta, const float* S, int B, int C, int T, int Tp,
                                  int P, void* da_planes, void* stream) {
    VNB_REQUIRE(dg && a && mean && rstd && gamma && beta && S && da_planes, "gate_bwd_apply: null pointer");
    VNB_REQUIRE(B > 0 && C > 0 && T > 0 && Tp >= T && (P == 1 || P == 3), "gate_bwd_apply: bad shape");
    gate_bwd_apply_kernel<<<ew_grid((int64_t)B * 2 * C * T), EW_THREADS, 0, (cudaStream_t)stream>>>(
        dg, a, mean, rstd, gamma, beta, S, B, C, T, Tp, P, (bf16*)da_planes);
    VNB_LAUNCH_CHECK("gate_bwd_apply");
    return VNB_OK;
}

extern "C" int vnb_act_bwd_add(const float* du, const float* c, const float* add, int64_t n, int act, float scale,
                               float* out, void* stream) {
    VNB_REQUIRE(du && c && out && n > 0, "act_bwd_add: bad arguments");
    act_bwd_add_kernel<<<ew_grid(n), EW_THREADS, 0, (cudaStream_t)stream>>>(du, c, add, n, act, scale, out);
    VNB_LAUNCH_CHECK("act_bwd_add");
    return VNB_OK;
}

extern "C" int vnb_channel_sum(const float* x, int B, int C, int T, float* out, void* stream) {
    VNB_REQUIRE(x && out && B > 0 && C > 0 && T > 0, "channel_sum: bad arguments");
    channel_sum_kernel<<<C, 256, 0, (cudaStream_t)stream>>>(x, B, C, T, out);
    VNB_LAUNCH_CHECK("channel_sum");
    return VNB_OK;
}
