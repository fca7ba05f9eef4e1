// Second-generation streaming kernels for the memory-bound plane writers (make_planes, enc_out, dec_gate,
// gn_bwd_apply, gate_bwd_apply).  What the first versions left on the table at large frame counts (profiles/r01_kernels_final.txt:
// 0.45-0.62 of the measured HBM peak at 2^17 frames):
//   * ONE 16-byte load in flight per thread per tensor and loop iteration (2048 threads x 16 B = 32 KB per SM, about
//     the bandwidth-delay product of one SM's share of HBM) - the kernels with three input tensors reached the roof,
//     the ones with one or two did not.  Here each thread issues the loads of U rows before it touches any of them.
//   * a 64-bit division per element to split the flat index into (row, channel), and the per-channel affine
//     parameters re-loaded for every element.  Here the grid is sized so that the total thread count is a multiple
//     of the channel-vector count: a thread's channel is loop-invariant (affine parameters live in registers), rows
//     advance by a constant, and there is no division in the loop except the row -> sample one.
//   * runtime P (1 or 3 planes) branches around every store: P is a template parameter.
//   * (bf16 operand mode only) libm tanhf / expf / IEEE division in the gate: ~50 instructions per element, which
//     made dec_gate / gate_bwd_apply issue-bound.  With P == 1 the result is rounded to bf16 (8 significant bits)
//     anyway, so tanh.approx.f32 (2^-11) and __expf / __fdividef are used; P == 3 keeps the accurate functions.
// Selected by option VNB_OPT_EW_V2 (+ VNB_OPT_FAST_GATE for the approximate gate math); shapes the scheme does not
// cover fall back to the first versions.
#pragma once
#include "common.cuh"

namespace vnb {

// runtime switches (vnb_set_option / environment, csrc/runtime.cu)
static inline bool ew_v2_enabled() { return option(VNB_OPT_EW_V2) != 0; }
static inline bool fast_gate_enabled() { return option(VNB_OPT_FAST_GATE) != 0; }

constexpr int V2_THREADS = 256;
static inline int v2_gcd(int a, int b) {
    while (b) {
        const int t = a % b;
        a = b;
        b = t;
    }
    return a;
}
// Number of CTAs such that grid * 256 is a multiple of Cv (channel vectors per row); 0 = scheme not applicable.
static inline int v2_grid(int64_t rows, int Cv) {
    if (Cv <= 0 || rows <= 0 || rows >= (1 << 30)) return 0;
    const int g0 = Cv / v2_gcd(V2_THREADS, Cv);
    const int cap = 148 * 8;
    if (g0 > cap) return 0;
    int64_t g = (rows * Cv + V2_THREADS - 1) / V2_THREADS;
    g = (g + g0 - 1) / g0 * g0;
    if (g > cap) g = cap / g0 * g0;
    return (int)g;
}

template <int P>
__device__ __forceinline__ void st_planes4(bf16* __restrict__ base, int64_t off, int64_t plane_stride,
                                           const float (&x)[4]) {
    bf16 h[3][4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        if constexpr (P == 1) h[0][j] = __float2bfloat16_rn(x[j]);
        else split3(x[j], h[0][j], h[1][j], h[2][j]);
    }
#pragma unroll
    for (int p = 0; p < P; p++) {
        uint2 u;
        u.x = (uint32_t)__bfloat16_as_ushort(h[p][0]) | ((uint32_t)__bfloat16_as_ushort(h[p][1]) << 16);
        u.y = (uint32_t)__bfloat16_as_ushort(h[p][2]) | ((uint32_t)__bfloat16_as_ushort(h[p][3]) << 16);
        *reinterpret_cast<uint2*>(base + off + p * plane_stride) = u;
    }
}

template <bool FAST>
__device__ __forceinline__ float gate_tanh(float x) {
    if constexpr (FAST) {
        float y;
        asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
        return y;
    } else {
        return tanhf(x);
    }
}
template <bool FAST>
__device__ __forceinline__ float gate_sigmoid(float x) {
    if constexpr (FAST) return __fdividef(1.f, 1.f + __expf(-x));
    else return 1.f / (1.f + expf(-x));
}

// per-thread loop frame shared by the kernels below: channel c (4 wide) fixed, rows r0, r0 + rstep, ...
#define V2_FRAME(Cv_)                                                                   \
    const uint32_t v2_tid = blockIdx.x * (uint32_t)V2_THREADS + threadIdx.x;            \
    const uint32_t v2_nthr = gridDim.x * (uint32_t)V2_THREADS;                          \
    const int c = (int)(v2_tid % (uint32_t)(Cv_)) * 4;                                  \
    const int rstep = (int)(v2_nthr / (uint32_t)(Cv_));                                 \
    const int r_first = (int)(v2_tid / (uint32_t)(Cv_))

// Round 2: the consumers of a conv's fused GroupNorm statistics finalise them THEMSELVES (one less launch per
// normalisation on the forward critical path): every CTA turns the fp64 (sum, sum of squares) pairs into (mean, rstd)
// in shared memory - the arithmetic of gn_finalize_kernel, so the values are bit-identical - and CTA 0 also stores them
// for the backward pass.
__device__ __forceinline__ void v2_finalize_stats(const double* __restrict__ sums, int BG, double n, float eps,
                                                  float* __restrict__ sm, float* __restrict__ sr,
                                                  float* __restrict__ mean_out, float* __restrict__ rstd_out) {
    for (int i = threadIdx.x; i < BG; i += blockDim.x) {
        const double m = sums[2 * i] / n;
        double var = sums[2 * i + 1] / n - m * m;
        if (var < 0.0) var = 0.0;
        const float mf = (float)m, rf = (float)(1.0 / sqrt(var + (double)eps));
        sm[i] = mf;
        sr[i] = rf;
        if (blockIdx.x == 0) {
            mean_out[i] = mf;
            rstd_out[i] = rf;
        }
    }
    __syncthreads();
}

// ---- fp32 activation -> operand planes (identity and/or activated) -------------------------------------------
template <int P, int U>
__global__ void __launch_bounds__(V2_THREADS) make_planes_v2_kernel(const float* __restrict__ x, int rows, int C, int Cp,
                                                                    float scale, int act, bf16* __restrict__ out_id,
                                                                    bf16* __restrict__ out_act) {
    pdl_wait();
    V2_FRAME(C >> 2);
    const int64_t plane = (int64_t)rows * Cp;
    for (int r0 = r_first; r0 < rows; r0 += U * rstep) {
        float v[U][4];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int r = r0 + u * rstep;
            if (r < rows) ld_vec<4>(x + (int64_t)r * C + c, v[u]);
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int r = r0 + u * rstep;
            if (r >= rows) continue;
            float w[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                v[u][j] *= scale;
                w[j] = apply_act(v[u][j], act);
            }
            const int64_t o = (int64_t)r * Cp + c;
            if (out_id) st_planes4<P>(out_id, o, plane, v[u]);
            if (out_act) st_planes4<P>(out_act, o, plane, w);
        }
    }
}

// ---- encoder block epilogue: y = GN1(h)*gamma+beta (+ s) -> y fp32, planes(y), planes(lrelu(y)) ---------------
template <int P, int U>
__global__ void __launch_bounds__(V2_THREADS) enc_out_v2_kernel(const float* __restrict__ h, const float* __restrict__ s,
                                                                const float* __restrict__ mean_in,
                                                                const float* __restrict__ rstd_in,
                                                                const float* __restrict__ gamma,
                                                                const float* __restrict__ beta, int rows, int C, int T,
                                                                int Cp, float* __restrict__ y, bf16* __restrict__ y_planes,
                                                                bf16* __restrict__ u_planes, const double* __restrict__ sums,
                                                                double n_stat, float eps, float* __restrict__ mean_out,
                                                                float* __restrict__ rstd_out) {
    pdl_wait();
    extern __shared__ float v2_stats[];
    const float* mean = mean_in;
    const float* rstd = rstd_in;
    if (sums) {
        const int BG = rows / T;
        v2_finalize_stats(sums, BG, n_stat, eps, v2_stats, v2_stats + BG, mean_out, rstd_out);
        mean = v2_stats;
        rstd = v2_stats + BG;
    }
    V2_FRAME(C >> 2);
    const int64_t plane = (int64_t)rows * Cp;
    float gm[4], bt[4];
    ld_vec<4>(gamma + c, gm);
    ld_vec<4>(beta + c, bt);
    for (int r0 = r_first; r0 < rows; r0 += U * rstep) {
        float v[U][4], sv[U][4];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int r = r0 + u * rstep;
            if (r < rows) {
                ld_vec<4>(h + (int64_t)r * C + c, v[u]);
                if (s) ld_vec<4>(s + (int64_t)r * C + c, sv[u]);
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int r = r0 + u * rstep;
            if (r >= rows) continue;
            const int b = r / T;
            const float m = mean[b], rs = rstd[b];
            float w[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                v[u][j] = (v[u][j] - m) * rs * gm[j] + bt[j];
                if (s) v[u][j] += sv[u][j];
                w[j] = lrelu02(v[u][j]);
            }
            if (y) st_vec<4>(y + (int64_t)r * C + c, v[u]);
            const int64_t o = (int64_t)r * Cp + c;
            if (y_planes) st_planes4<P>(y_planes, o, plane, v[u]);
            if (u_planes) st_planes4<P>(u_planes, o, plane, w);
        }
    }
}

// ---- decoder gate: a (rows, 2C) -> planes of tanh(GN2(a)[:C]) * sigmoid(GN2(a)[C:]) ---------------------------
template <int P, int U, bool FAST>
__global__ void __launch_bounds__(V2_THREADS) dec_gate_v2_kernel(const float* __restrict__ a, const float* __restrict__ mean_in,
                                                                 const float* __restrict__ rstd_in,
                                                                 const float* __restrict__ gamma,
                                                                 const float* __restrict__ beta, int rows, int C, int T,
                                                                 int Cp, bf16* __restrict__ g_planes,
                                                                 const double* __restrict__ sums, double n_stat, float eps,
                                                                 float* __restrict__ mean_out, float* __restrict__ rstd_out) {
    pdl_wait();
    extern __shared__ float v2_stats[];
    const float* mean = mean_in;
    const float* rstd = rstd_in;
    if (sums) {
        const int BG = 2 * (rows / T);
        v2_finalize_stats(sums, BG, n_stat, eps, v2_stats, v2_stats + BG, mean_out, rstd_out);
        mean = v2_stats;
        rstd = v2_stats + BG;
    }
    V2_FRAME(C >> 2);
    const int64_t plane = (int64_t)rows * Cp;
    float g1[4], g2[4], b1[4], b2[4];
    ld_vec<4>(gamma + c, g1);
    ld_vec<4>(gamma + C + c, g2);
    ld_vec<4>(beta + c, b1);
    ld_vec<4>(beta + C + c, b2);
    for (int r0 = r_first; r0 < rows; r0 += U * rstep) {
        float a1[U][4], a2[U][4];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int r = r0 + u * rstep;
            if (r < rows) {
                ld_vec<4>(a + (int64_t)r * 2 * C + c, a1[u]);
                ld_vec<4>(a + (int64_t)r * 2 * C + C + c, a2[u]);
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int r = r0 + u * rstep;
            if (r >= rows) continue;
            const int b = r / T;
            const float m1 = mean[2 * b], r1 = rstd[2 * b], m2 = mean[2 * b + 1], r2 = rstd[2 * b + 1];
            float o[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const float h1 = (a1[u][j] - m1) * r1 * g1[j] + b1[j];
                const float h2 = (a2[u][j] - m2) * r2 * g2[j] + b2[j];
                o[j] = gate_tanh<FAST>(h1) * gate_sigmoid<FAST>(h2);
            }
            st_planes4<P>(g_planes, (int64_t)r * Cp + c, plane, o);
        }
    }
}

// ---- GroupNorm backward apply: dpre = rstd * (gamma*dn - S0/n - xhat*S1/n) -> planes (+ planes of dn) ----------
// requires (C / groups) % 4 == 0 so that a thread's four channels share a group
template <int P, int U>
__global__ void __launch_bounds__(V2_THREADS) gn_bwd_apply_v2_kernel(const float* __restrict__ dn, const float* __restrict__ h,
                                                                     const float* __restrict__ mean,
                                                                     const float* __restrict__ rstd,
                                                                     const float* __restrict__ gamma,
                                                                     const float* __restrict__ S, int rows, int C, int T,
                                                                     int groups, int Cp, bf16* __restrict__ dpre_planes,
                                                                     bf16* __restrict__ dn_planes) {
    pdl_wait();
    V2_FRAME(C >> 2);
    const int64_t plane = (int64_t)rows * Cp;
    const int cg = C / groups;
    const int g = c / cg;
    const float inv_n = 1.f / ((float)cg * (float)T);
    float gm[4];
    ld_vec<4>(gamma + c, gm);
    for (int r0 = r_first; r0 < rows; r0 += U * rstep) {
        float d[U][4], hv[U][4];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int r = r0 + u * rstep;
            if (r < rows) {
                ld_vec<4>(dn + (int64_t)r * C + c, d[u]);
                ld_vec<4>(h + (int64_t)r * C + c, hv[u]);
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int r = r0 + u * rstep;
            if (r >= rows) continue;
            const int gi = (r / T) * groups + g;
            const float rs = rstd[gi], mu = mean[gi];
            const float s0 = S[gi * 2] * inv_n, s1 = S[gi * 2 + 1] * inv_n;
            float v[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const float xh = (hv[u][j] - mu) * rs;
                v[j] = rs * (gm[j] * d[u][j] - s0 - xh * s1);
            }
            const int64_t o = (int64_t)r * Cp + c;
            st_planes4<P>(dpre_planes, o, plane, v);
            if (dn_planes) st_planes4<P>(dn_planes, o, plane, d[u]);
        }
    }
}

// ---- gate backward apply: dg (rows, C), a (rows, 2C) -> da planes (rows, (2C)p), both halves per thread -------
template <int P, int U, bool FAST>
__global__ void __launch_bounds__(V2_THREADS) gate_bwd_apply_v2_kernel(const float* __restrict__ dg, const float* __restrict__ a,
                                                                       const float* __restrict__ mean,
                                                                       const float* __restrict__ rstd,
                                                                       const float* __restrict__ gamma,
                                                                       const float* __restrict__ beta,
                                                                       const float* __restrict__ S, int rows, int C, int T,
                                                                       int Cp, bf16* __restrict__ da_planes) {
    pdl_wait();
    V2_FRAME(C >> 2);
    const int64_t plane = (int64_t)rows * Cp;
    const float inv_n = 1.f / ((float)C * (float)T);
    float g1[4], g2[4], b1[4], b2[4];
    ld_vec<4>(gamma + c, g1);
    ld_vec<4>(gamma + C + c, g2);
    ld_vec<4>(beta + c, b1);
    ld_vec<4>(beta + C + c, b2);
    for (int r0 = r_first; r0 < rows; r0 += U * rstep) {
        float a1[U][4], a2[U][4], d[U][4];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int r = r0 + u * rstep;
            if (r < rows) {
                ld_vec<4>(a + (int64_t)r * 2 * C + c, a1[u]);
                ld_vec<4>(a + (int64_t)r * 2 * C + C + c, a2[u]);
                ld_vec<4>(dg + (int64_t)r * C + c, d[u]);
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int r = r0 + u * rstep;
            if (r >= rows) continue;
            const int b = r / T;
            const float m1 = mean[2 * b], m2 = mean[2 * b + 1], rs1 = rstd[2 * b], rs2 = rstd[2 * b + 1];
            const float S10 = S[4 * b] * inv_n, S11 = S[4 * b + 1] * inv_n, S20 = S[4 * b + 2] * inv_n,
                        S21 = S[4 * b + 3] * inv_n;
            float v1[4], v2[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const float x1 = (a1[u][j] - m1) * rs1, x2 = (a2[u][j] - m2) * rs2;
                const float th = gate_tanh<FAST>(x1 * g1[j] + b1[j]);
                const float sg = gate_sigmoid<FAST>(x2 * g2[j] + b2[j]);
                const float dn1 = d[u][j] * sg * (1.f - th * th), dn2 = d[u][j] * th * sg * (1.f - sg);
                v1[j] = rs1 * (g1[j] * dn1 - S10 - x1 * S11);
                v2[j] = rs2 * (g2[j] * dn2 - S20 - x2 * S21);
            }
            st_planes4<P>(da_planes, (int64_t)r * Cp + c, plane, v1);
            st_planes4<P>(da_planes, (int64_t)r * Cp + C + c, plane, v2);
        }
    }
}

}  // namespace vnb
