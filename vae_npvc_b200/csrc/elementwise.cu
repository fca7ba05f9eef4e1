// Memory-bound fused kernels around the GEMMs: operand-plane writers, GroupNorm statistics /
// apply / backward, the decoder gate, residual/skip bookkeeping, bias reductions.
// Reference op sequences: vae_npvc/model/layers.py:151-155,168-176 (encoder residual block),
// :229-247 (gated decoder block); math in SURVEY.md section 9 A/B.
#include "common.cuh"

namespace vnb {

constexpr int EW_THREADS = 256;
static inline int ew_grid(int64_t n) {
    int64_t g = (n + EW_THREADS - 1) / EW_THREADS;
    const int64_t cap = 148 * 16;          // grid-stride beyond 16 CTAs per SM
    return (int)(g < cap ? g : cap);
}
#define GRID_STRIDE(i, n) \
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < (n); i += (int64_t)gridDim.x * blockDim.x)

// ------------------------------------------------------------------------------------------
__global__ void make_planes_kernel(const float* __restrict__ x, int64_t rows, int T, int Tp, int P, float scale,
                                   int act, bf16* __restrict__ out_id, bf16* __restrict__ out_act) {
    const int64_t n = rows * T, plane = rows * Tp;
    GRID_STRIDE(i, n) {
        const int64_t r = i / T;
        const int t = (int)(i - r * T);
        const float v = x[i] * scale;
        const int64_t o = r * Tp + t;
        if (out_id) store_planes(out_id, o, plane, P, v);
        if (out_act) store_planes(out_act, o, plane, P, apply_act(v, act));
    }
}

__global__ void concat_planes_kernel(const float* __restrict__ xa, int Ca, const float* __restrict__ xb, int Cb, int B,
                                     int T, int Tp, int P, bf16* __restrict__ out) {
    const int Cc = Ca + Cb;
    const int64_t n = (int64_t)B * Cc * T, plane = (int64_t)B * Cc * Tp;
    GRID_STRIDE(i, n) {
        const int64_t r = i / T;
        const int t = (int)(i - r * T);
        const int b = (int)(r / Cc), c = (int)(r - (int64_t)b * Cc);
        const float v = c < Ca ? xa[((int64_t)b * Ca + c) * T + t] : xb[((int64_t)b * Cb + (c - Ca)) * T + t];
        store_planes(out, r * Tp + t, plane, P, v);
    }
}

// one block per (sample, group); the group's (C/groups)*T values are contiguous in NCT
__global__ void __launch_bounds__(512) gn_stats_kernel(const float* __restrict__ h, int64_t n, float eps,
                                                       float* __restrict__ mean, float* __restrict__ rstd) {
    __shared__ float red[32];
    const float* p = h + (int64_t)blockIdx.x * n;
    float s = 0.f;
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) s += p[i];
    const float m = block_sum(s, red) / (float)n;
    float q = 0.f;
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
        const float d = p[i] - m;
        q += d * d;
    }
    const float var = block_sum(q, red) / (float)n;
    if (threadIdx.x == 0) {
        mean[blockIdx.x] = m;
        rstd[blockIdx.x] = rsqrtf(var + eps);
    }
}

__global__ void enc_out_kernel(const float* __restrict__ h, const float* __restrict__ s, const float* __restrict__ mean,
                               const float* __restrict__ rstd, const float* __restrict__ gamma,
                               const float* __restrict__ beta, int B, int C, int T, int Tp, int P,
                               float* __restrict__ y, bf16* __restrict__ y_planes, bf16* __restrict__ u_planes) {
    const int64_t n = (int64_t)B * C * T, plane = (int64_t)B * C * Tp;
    GRID_STRIDE(i, n) {
        const int64_t r = i / T;
        const int t = (int)(i - r * T);
        const int b = (int)(r / C), c = (int)(r - (int64_t)b * C);
        float v = (h[i] - mean[b]) * rstd[b] * gamma[c] + beta[c];
        if (s) v += s[i];
        if (y) y[i] = v;
        const int64_t o = r * Tp + t;
        if (y_planes) store_planes(y_planes, o, plane, P, v);
        if (u_planes) store_planes(u_planes, o, plane, P, lrelu02(v));
    }
}

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }

__global__ void dec_gate_kernel(const float* __restrict__ a, const float* __restrict__ mean,
                                const float* __restrict__ rstd, const float* __restrict__ gamma,
                                const float* __restrict__ beta, int B, int C, int T, int Tp, int P,
                                bf16* __restrict__ g_planes) {
    const int64_t n = (int64_t)B * C * T, plane = (int64_t)B * C * Tp;
    GRID_STRIDE(i, n) {
        const int64_t r = i / T;
        const int t = (int)(i - r * T);
        const int b = (int)(r / C), c = (int)(r - (int64_t)b * C);
        const int64_t i1 = ((int64_t)b * 2 * C + c) * T + t, i2 = i1 + (int64_t)C * T;
        const float h1 = (a[i1] - mean[2 * b]) * rstd[2 * b] * gamma[c] + beta[c];
        const float h2 = (a[i2] - mean[2 * b + 1]) * rstd[2 * b + 1] * gamma[C + c] + beta[C + c];
        store_planes(g_planes, r * Tp + t, plane, P, tanhf(h1) * sigmoidf_(h2));
    }
}

__global__ void dec_out_kernel(const float* __restrict__ r, const float* __restrict__ x, int B, int C, int S, int T,
                               int Tp, int P, float* __restrict__ x_out, bf16* __restrict__ x_planes,
                               float* __restrict__ skip_sum, int first) {
    const int Cc = C + S;
    const int64_t n = (int64_t)B * Cc * T, plane = (int64_t)B * C * Tp;
    GRID_STRIDE(i, n) {
        const int64_t row = i / T;
        const int t = (int)(i - row * T);
        const int b = (int)(row / Cc), c = (int)(row - (int64_t)b * Cc);
        const float v = r[i];
        if (c < C) {
            const int64_t o = ((int64_t)b * C + c) * T + t;
            const float xo = v + x[o];
            if (x_out) x_out[o] = xo;
            if (x_planes) store_planes(x_planes, ((int64_t)b * C + c) * Tp + t, plane, P, xo);
        } else if (skip_sum) {
            const int64_t o = ((int64_t)b * S + (c - C)) * T + t;
            skip_sum[o] = first ? v : skip_sum[o] + v;
        }
    }
}

}  // namespace vnb
using namespace vnb;

extern "C" int vnb_make_planes(const float* x, int B, int C, int T, int Tp, int P, float scale, int act, void* out_id,
                               void* out_act, void* stream) {
    VNB_REQUIRE(x && (out_id || out_act) && B > 0 && C > 0 && T > 0 && Tp >= T && (P == 1 || P == 3), "make_planes: bad arguments");
    const int64_t rows = (int64_t)B * C;
    make_planes_kernel<<<ew_grid(rows * T), EW_THREADS, 0, (cudaStream_t)stream>>>(x, rows, T, Tp, P, scale, act,
                                                                                  (bf16*)out_id, (bf16*)out_act);
    VNB_LAUNCH_CHECK("make_planes");
    return VNB_OK;
}

extern "C" int vnb_concat_planes(const float* xa, int Ca, const float* xb, int Cb, int B, int T, int Tp, int P,
                                 void* out, void* stream) {
    VNB_REQUIRE(xa && xb && out && Ca > 0 && Cb > 0 && B > 0 && T > 0 && Tp >= T && (P == 1 || P == 3), "concat_planes: bad arguments");
    concat_planes_kernel<<<ew_grid((int64_t)B * (Ca + Cb) * T), EW_THREADS, 0, (cudaStream_t)stream>>>(
        xa, Ca, xb, Cb, B, T, Tp, P, (bf16*)out);
    VNB_LAUNCH_CHECK("concat_planes");
    return VNB_OK;
}

extern "C" int vnb_gn_stats(const float* h, int B, int C, int T, int groups, float eps, float* mean, float* rstd,
                            void* stream) {
    VNB_REQUIRE(h && mean && rstd && B > 0 && groups > 0 && C % groups == 0 && T > 0, "gn_stats: bad arguments");
    gn_stats_kernel<<<B * groups, 512, 0, (cudaStream_t)stream>>>(h, (int64_t)(C / groups) * T, eps, mean, rstd);
    VNB_LAUNCH_CHECK("gn_stats");
    return VNB_OK;
}

extern "C" int vnb_enc_out(const float* h, const float* s, const float* mean, const float* rstd, const float* gamma,
                           const float* beta, int B, int C, int T, int Tp, int P, float* y, void* y_planes,
                           void* u_planes, void* stream) {
    VNB_REQUIRE(h && mean && rstd && gamma && beta && (y || y_planes || u_planes), "enc_out: null pointer");
    VNB_REQUIRE(B > 0 && C > 0 && T > 0 && Tp >= T && (P == 1 || P == 3), "enc_out: bad shape");
    enc_out_kernel<<<ew_grid((int64_t)B * C * T), EW_THREADS, 0, (cudaStream_t)stream>>>(
        h, s, mean, rstd, gamma, beta, B, C, T, Tp, P, y, (bf16*)y_planes, (bf16*)u_planes);
    VNB_LAUNCH_CHECK("enc_out");
    return VNB_OK;
}

extern "C" int vnb_dec_gate(const float* a, const float* mean, const float* rstd, const float* gamma,
                            const float* beta, int B, int C, int T, int Tp, int P, void* g_planes, void* stream) {
    VNB_REQUIRE(a && mean && rstd && gamma && beta && g_planes, "dec_gate: null pointer");
    VNB_REQUIRE(B > 0 && C > 0 && T > 0 && Tp >= T && (P == 1 || P == 3), "dec_gate: bad shape");
    dec_gate_kernel<<<ew_grid((int64_t)B * C * T), EW_THREADS, 0, (cudaStream_t)stream>>>(a, mean, rstd, gamma, beta, B,
                                                                                         C, T, Tp, P, (bf16*)g_planes);
    VNB_LAUNCH_CHECK("dec_gate");
    return VNB_OK;
}

extern "C" int vnb_dec_out(const float* r, const float* x, int B, int C, int S, int T, int Tp, int P, float* x_out,
                           void* x_planes, float* skip_sum, int first, void* stream) {
    VNB_REQUIRE(r && x && B > 0 && C > 0 && S >= 0 && T > 0 && Tp >= T && (P == 1 || P == 3), "dec_out: bad arguments");
    dec_out_kernel<<<ew_grid((int64_t)B * (C + S) * T), EW_THREADS, 0, (cudaStream_t)stream>>>(
        r, x, B, C, S, T, Tp, P, x_out, (bf16*)x_planes, skip_sum, first);
    VNB_LAUNCH_CHECK("dec_out");
    return VNB_OK;
}
