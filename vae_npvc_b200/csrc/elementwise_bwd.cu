// Backward-side memory-bound kernels: GroupNorm / gate gradient row statistics, their
// finalisation, gradient application into operand planes, activation masks, bias sums.
// Math: SURVEY.md section 9 A/B (restated from vae_npvc/model/layers.py:151-176, 229-247).
#include "common.cuh"

namespace vnb {

constexpr int EW_THREADS = 256;
static inline int ew_grid(int64_t n) {
    int64_t g = (n + EW_THREADS - 1) / EW_THREADS;
    const int64_t cap = 148 * 8;
    return (int)(g < cap ? g : cap);
}
#define GRID_STRIDE(i, n) \
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < (n); i += (int64_t)gridDim.x * blockDim.x)

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }

// gradient wrt the normalised+affine gate input channel c2 in [0,2C) of frame row r (channels-last:
// a (rows, 2C), dg (rows, C)), recomputed from dg
__device__ __forceinline__ float gate_dn(const float* __restrict__ dg, const float* __restrict__ a,
                                         const float* __restrict__ mean, const float* __restrict__ rstd,
                                         const float* __restrict__ gamma, const float* __restrict__ beta, int b,
                                         int64_t r, int c2, int C, float& xhat) {
    const int c = c2 < C ? c2 : c2 - C;
    const float x1 = (a[r * 2 * C + c] - mean[2 * b]) * rstd[2 * b];
    const float x2 = (a[r * 2 * C + C + c] - mean[2 * b + 1]) * rstd[2 * b + 1];
    const float th = tanhf(x1 * gamma[c] + beta[c]);
    const float sg = sigmoidf_(x2 * gamma[C + c] + beta[C + c]);
    const float d = dg[r * C + c];
    if (c2 < C) {
        xhat = x1;
        return d * sg * (1.f - th * th);
    }
    xhat = x2;
    return d * th * sg * (1.f - sg);
}

// R[b,c] = { sum_t dn, sum_t dn*xhat, sum_t xhat }.  grid (ceil(C/32), B), block 32 x 8: a warp reads
// 32 consecutive channels of one frame, the 8 warps stride over frames, shared-memory reduce.
template <bool GATE>
__global__ void __launch_bounds__(256) bwd_rows_kernel(const float* __restrict__ dn, const float* __restrict__ h,
                                                       const float* __restrict__ mean, const float* __restrict__ rstd,
                                                       const float* __restrict__ gamma, const float* __restrict__ beta,
                                                       int Cn, int T, int groups, float* __restrict__ R) {
    __shared__ float red[3][8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int b = blockIdx.y, c = blockIdx.x * 32 + tx;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f;
    if (c < Cn) {
        const int gi = b * groups + c / (Cn / groups);
        const float m = mean[gi], rs = rstd[gi];
        for (int t = ty; t < T; t += 8) {
            const int64_t r = (int64_t)b * T + t;
            float d, xh;
            if (GATE) d = gate_dn(dn, h, mean, rstd, gamma, beta, b, r, c, Cn / 2, xh);
            else {
                d = dn[r * Cn + c];
                xh = (h[r * Cn + c] - m) * rs;
            }
            s0 += d;
            s1 += d * xh;
            s2 += xh;
        }
    }
    red[0][ty][tx] = s0;
    red[1][ty][tx] = s1;
    red[2][ty][tx] = s2;
    __syncthreads();
    if (ty < 3 && c < Cn) {
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < 8; j++) s += red[ty][j][tx];
        R[((int64_t)b * Cn + c) * 3 + ty] = s;
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
                                      float* __restrict__ dpre_bsum, float* __restrict__ dpre_csum,
                                      float* __restrict__ dbeta2, float* __restrict__ dpre_csum2) {
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
        if (dpre_bsum || dpre_csum || dpre_csum2) {
            const float* s = S + (b * groups + g) * 2;
            const float v = rstd[b * groups + g] * (gm * r[0] - (float)T * s[0] * inv_n - r[2] * s[1] * inv_n);
            if (dpre_bsum) dpre_bsum[(int64_t)b * C + c] = v;
            dcs += v;
        }
    }
    if (dgamma) dgamma[c] = dgm;
    if (dbeta) dbeta[c] = dbt;
    if (dpre_csum) dpre_csum[c] = dcs;
    if (dbeta2) dbeta2[c] = dbt;
    if (dpre_csum2) dpre_csum2[c] = dcs;
}

template <int VEC>
__global__ void gn_bwd_apply_kernel(const float* __restrict__ dn, const float* __restrict__ h,
                                    const float* __restrict__ mean, const float* __restrict__ rstd,
                                    const float* __restrict__ gamma, const float* __restrict__ S, int64_t rows, int C,
                                    int T, int groups, int Cp, int P, bf16* __restrict__ dpre_planes,
                                    bf16* __restrict__ dn_planes) {
    const int Cv = C / VEC;
    const int64_t n = rows * Cv, plane = rows * Cp;
    const int cg = C / groups;
    const float inv_n = 1.f / ((float)cg * (float)T);
    GRID_STRIDE(i, n) {
        const int64_t r = i / Cv;
        const int c = (int)(i - r * Cv) * VEC;
        const int b = (int)(r / T);
        float d[VEC], hv[VEC], v[VEC];
        ld_vec<VEC>(dn + r * C + c, d);
        ld_vec<VEC>(h + r * C + c, hv);
#pragma unroll
        for (int j = 0; j < VEC; j++) {
            const int gi = b * groups + (c + j) / cg;
            const float rs = rstd[gi];
            const float xh = (hv[j] - mean[gi]) * rs;
            v[j] = rs * (gamma[c + j] * d[j] - S[gi * 2] * inv_n - xh * S[gi * 2 + 1] * inv_n);
        }
        const int64_t o = r * Cp + c;
        st_planes_vec<VEC>(dpre_planes, o, plane, P, v);
        if (dn_planes) st_planes_vec<VEC>(dn_planes, o, plane, P, d);
    }
}

// da planes (rows, (2C)p) from dg (rows, C), a (rows, 2C)
template <int VEC>
__global__ void gate_bwd_apply_kernel(const float* __restrict__ dg, const float* __restrict__ a,
                                      const float* __restrict__ mean, const float* __restrict__ rstd,
                                      const float* __restrict__ gamma, const float* __restrict__ beta,
                                      const float* __restrict__ S, int64_t rows, int C, int T, int Cp, int P,
                                      bf16* __restrict__ da_planes) {
    const int C2 = 2 * C, Cv = C2 / VEC;
    const int64_t n = rows * Cv, plane = rows * Cp;
    const float inv_n = 1.f / ((float)C * (float)T);
    GRID_STRIDE(i, n) {
        const int64_t r = i / Cv;
        const int c2 = (int)(i - r * Cv) * VEC;
        const int b = (int)(r / T);
        float v[VEC];
#pragma unroll
        for (int j = 0; j < VEC; j++) {
            const int gi = b * 2 + (c2 + j >= C ? 1 : 0);
            float xh;
            const float d = gate_dn(dg, a, mean, rstd, gamma, beta, b, r, c2 + j, C, xh);
            v[j] = rstd[gi] * (gamma[c2 + j] * d - S[gi * 2] * inv_n - xh * S[gi * 2 + 1] * inv_n);
        }
        st_planes_vec<VEC>(da_planes, r * Cp + c2, plane, P, v);
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

// out[c] = sum over the B*T frame rows; one block (32 x 32) per 32 channels
__global__ void __launch_bounds__(1024) channel_sum_kernel(const float* __restrict__ x, int64_t rows, int C,
                                                           float* __restrict__ out) {
    __shared__ float red[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + tx;
    float s = 0.f;
    if (c < C)
        for (int64_t r = ty; r < rows; r += 32) s += x[r * C + c];
    red[ty][tx] = s;
    __syncthreads();
    if (ty == 0 && c < C) {
        float t = 0.f;
#pragma unroll
        for (int j = 0; j < 32; j++) t += red[j][tx];
        out[c] = t;
    }
}

}  // namespace vnb
using namespace vnb;

extern "C" int vnb_gn_bwd_rows(const float* dn, const float* h, const float* mean, const float* rstd, int B, int C,
                               int T, int groups, float* R, void* stream) {
    VNB_REQUIRE(dn && h && mean && rstd && R && B > 0 && B <= 65535 && C > 0 && T > 0 && groups > 0 && C % groups == 0,
                "gn_bwd_rows: bad arguments");
    bwd_rows_kernel<false><<<dim3(ceil_div(C, 32), B), 256, 0, (cudaStream_t)stream>>>(dn, h, mean, rstd, nullptr, nullptr,
                                                                                       C, T, groups, R);
    VNB_LAUNCH_CHECK("gn_bwd_rows");
    return VNB_OK;
}

extern "C" int vnb_gate_bwd_rows(const float* dg, const float* a, const float* mean, const float* rstd,
                                 const float* gamma, const float* beta, int B, int C, int T, float* R, void* stream) {
    VNB_REQUIRE(dg && a && mean && rstd && gamma && beta && R && B > 0 && B <= 65535 && C > 0 && T > 0,
                "gate_bwd_rows: bad arguments");
    bwd_rows_kernel<true><<<dim3(ceil_div(2 * C, 32), B), 256, 0, (cudaStream_t)stream>>>(dg, a, mean, rstd, gamma, beta,
                                                                                          2 * C, T, 2, R);
    VNB_LAUNCH_CHECK("gate_bwd_rows");
    return VNB_OK;
}

extern "C" int vnb_gn_bwd_finalize(const float* R, const float* gamma, const float* rstd, int B, int C, int T,
                                   int groups, float* dgamma, float* dbeta, float* S, float* dpre_bsum,
                                   float* dpre_csum, float* dbeta2, float* dpre_csum2, void* stream) {
    VNB_REQUIRE(R && gamma && rstd && S && B > 0 && C > 0 && groups > 0 && C % groups == 0, "gn_bwd_finalize: bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    gn_bwd_groupsum_kernel<<<B * groups, 256, 0, st>>>(R, gamma, C, groups, S);
    VNB_LAUNCH_CHECK("gn_bwd_groupsum");
    gn_bwd_channel_kernel<<<ceil_div(C, 128), 128, 0, st>>>(R, gamma, rstd, S, B, C, T, groups, dgamma, dbeta, dpre_bsum, dpre_csum, dbeta2,
                                                            dpre_csum2);
    VNB_LAUNCH_CHECK("gn_bwd_channel");
    return VNB_OK;
}

#define VEC_DISPATCH(cond4, KERNEL, n_scalar, st, ...)                                        \
    do {                                                                                      \
        if (cond4)                                                                            \
            KERNEL<4><<<ew_grid((n_scalar) / 4), EW_THREADS, 0, st>>>(__VA_ARGS__);           \
        else                                                                                  \
            KERNEL<1><<<ew_grid(n_scalar), EW_THREADS, 0, st>>>(__VA_ARGS__);                 \
    } while (0)

extern "C" int vnb_gn_bwd_apply(const float* dn, const float* h, const float* mean, const float* rstd,
                                const float* gamma, const float* S, int B, int C, int T, int groups, int P,
                                void* dpre_planes, void* dn_planes, void* stream) {
    VNB_REQUIRE(dn && h && mean && rstd && gamma && S && dpre_planes, "gn_bwd_apply: null pointer");
    VNB_REQUIRE(B > 0 && C > 0 && T > 0 && groups > 0 && C % groups == 0 && (P == 1 || P == 3), "gn_bwd_apply: bad shape");
    const int64_t rows = (int64_t)B * T;
    VEC_DISPATCH((C & 3) == 0, gn_bwd_apply_kernel, rows * C, (cudaStream_t)stream, dn, h, mean, rstd, gamma, S, rows, C, T,
                 groups, round_up8(C), P, (bf16*)dpre_planes, (bf16*)dn_planes);
    VNB_LAUNCH_CHECK("gn_bwd_apply");
    return VNB_OK;
}

extern "C" int vnb_gate_bwd_apply(const float* dg, const float* a, const float* mean, const float* rstd,
                                  const float* gamma, const float* beta, const float* S, int B, int C, int T, int P,
                                  void* da_planes, void* stream) {
    VNB_REQUIRE(dg && a && mean && rstd && gamma && beta && S && da_planes, "gate_bwd_apply: null pointer");
    VNB_REQUIRE(B > 0 && C > 0 && T > 0 && (P == 1 || P == 3), "gate_bwd_apply: bad shape");
    const int64_t rows = (int64_t)B * T;
    VEC_DISPATCH((C & 3) == 0, gate_bwd_apply_kernel, rows * 2 * C, (cudaStream_t)stream, dg, a, mean, rstd, gamma, beta, S,
                 rows, C, T, round_up8(2 * C), P, (bf16*)da_planes);
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
    channel_sum_kernel<<<ceil_div(C, 32), 1024, 0, (cudaStream_t)stream>>>(x, (int64_t)B * T, C, out);
    VNB_LAUNCH_CHECK("channel_sum");
    return VNB_OK;
}
