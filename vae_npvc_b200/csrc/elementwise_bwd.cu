// Backward-side memory-bound kernels: GroupNorm / gate gradient row statistics, their
// finalisation, gradient application into operand planes, activation masks, bias sums.
// Math: SURVEY.md section 9 A/B (restated from vae_npvc/model/layers.py:151-176, 229-247).
#include "common.cuh"
#include "ew_v2.cuh"

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

// Round 2: the row-sum kernels can also produce the per-(sample, group) sums S[b,g,k] = sum_{c in g} gamma_c R[b,c,k]
// (k = 0, 1) that the apply kernels need, so that gn_bwd_finalize (7 us for a B x groups problem) leaves the critical
// path: a block's 32 channels lie in one group (host checks 32 | C / groups), each of its two sum warps adds
// gamma-weighted partials to a ZEROED fp64 accumulator, and the LAST block to finish (ticket counter) publishes the
// sums as fp32 for the apply kernel.
__device__ __forceinline__ void publish_group_sums(float t, float gm, bool valid, int k, int slot, double* S64, float* S32,
                                                   unsigned* counter, int n_slots2) {
    __shared__ int last_flag;
    if (k < 2) {
        float v = valid ? t * gm : 0.f;
        v = warp_sum(v);
        if ((threadIdx.x & 31) == 0) atomicAdd(S64 + (size_t)slot * 2 + k, (double)v);
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) last_flag = atomicAdd(counter, 1u) == gridDim.x * gridDim.y - 1 ? 1 : 0;
    __syncthreads();
    if (last_flag) {
        __threadfence();
        for (int i = threadIdx.x; i < n_slots2; i += blockDim.x) S32[i] = (float)__ldcg(S64 + i);
    }
}

// R[b,c] = { sum_t dn, sum_t dn*xhat, sum_t xhat }.  grid (ceil(C/32), B), block 32 x NW: a warp reads
// 32 consecutive channels of one frame, the NW warps stride over frames, shared-memory reduce.  NW = 32 (opt-in, see
// gn_rows_nw32()): at 16 x 128 frames the kernel is latency-bound (four dependent rounds of loads with 8 warps, one with 32).
template <bool GATE, int NW>
__global__ void __launch_bounds__(32 * NW) bwd_rows_kernel(const float* __restrict__ dn, const float* __restrict__ h,
                                                       const float* __restrict__ mean, const float* __restrict__ rstd,
                                                       const float* __restrict__ gamma, const float* __restrict__ beta,
                                                       int Cn, int T, int groups, float* __restrict__ R,
                                                       double* __restrict__ S64, float* __restrict__ S32,
                                                       unsigned* __restrict__ counter) {
    pdl_wait();
    __shared__ float red[3][NW][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int b = blockIdx.y, c = blockIdx.x * 32 + tx;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f;
    if (c < Cn) {
        const int gi = b * groups + c / (Cn / groups);
        const float m = mean[gi], rs = rstd[gi];
        if (GATE) {
            for (int t = ty; t < T; t += NW) {
                const int64_t r = (int64_t)b * T + t;
                float xh;
                const float d = gate_dn(dn, h, mean, rstd, gamma, beta, b, r, c, Cn / 2, xh);
                s0 += d;
                s1 += d * xh;
                s2 += xh;
            }
        } else {
            for (int t0 = ty; t0 < T; t0 += 4 * NW) {      // four frames in flight
                float d[4], hv[4];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int t = t0 + NW * u;
                    const int64_t r = (int64_t)b * T + (t < T ? t : 0);
                    d[u] = t < T ? dn[r * Cn + c] : 0.f;
                    hv[u] = h[r * Cn + c];
                }
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    if (t0 + NW * u >= T) continue;
                    const float xh = (hv[u] - m) * rs;
                    s0 += d[u];
                    s1 += d[u] * xh;
                    s2 += xh;
                }
            }
        }
    }
    red[0][ty][tx] = s0;
    red[1][ty][tx] = s1;
    red[2][ty][tx] = s2;
    __syncthreads();
    float s = 0.f;
    if (ty < 3 && c < Cn) {
#pragma unroll
        for (int j = 0; j < NW; j++) s += red[ty][j][tx];
        R[((int64_t)b * Cn + c) * 3 + ty] = s;
    }
    if (S64) {
        const int c0 = blockIdx.x * 32;
        publish_group_sums(s, (c < Cn && ty < 2) ? gamma[c] : 0.f, c < Cn, ty, b * groups + c0 / (Cn / groups), S64, S32, counter,
                           gridDim.y * groups * 2);
    }
}

// GroupNorm backward finalisation in ONE launch (was: a (b,g) group-sum kernel, then a per-channel kernel on 4-8
// blocks that walked the samples serially: 4 + 10 us for 200 KB of input).
// grid ceil(C/32), block 16 warps.  Phase 1 (every block, redundantly - R is L2 resident): warp w computes
//   S[b,g,0..1] = sum_{c in g} gamma_c * R{0,1}[b,c]   for b = w, w+16, ... into shared memory (block 0 also
//   publishes S for gn_bwd_apply / gate_bwd_apply).
// Phase 2: lane = channel, warp = sample: dgamma, dbeta and the time / batch sums of the pre-norm gradient
//   dpre = rstd * (gamma*dn - S0/n - xhat*S1/n), reduced over the 16 warps in shared memory.
__global__ void __launch_bounds__(512) gn_bwd_finalize_kernel(const float* __restrict__ R, const float* __restrict__ gamma,
                                                              const float* __restrict__ rstd, int B, int C, int T, int groups,
                                                              float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                              float* __restrict__ S, float* __restrict__ dpre_bsum,
                                                              float* __restrict__ dpre_csum, float* __restrict__ dbeta2,
                                                              float* __restrict__ dpre_csum2) {
    pdl_wait();
    extern __shared__ float fin_smem[];
    float* Ss = fin_smem;                              // [B][groups][2]
    float (*red)[16][33] = reinterpret_cast<float (*)[16][33]>(fin_smem + ((B * groups * 2 + 3) & ~3));
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int cg = C / groups;
    for (int b = w; b < B; b += 16) {
        for (int g = 0; g < groups; g++) {
            float s0 = 0.f, s1 = 0.f;
            for (int j = lane; j < cg; j += 32) {
                const int c = g * cg + j;
                const float gm = gamma[c];
                const float* r = R + ((int64_t)b * C + c) * 3;
                s0 = fmaf(gm, r[0], s0);
                s1 = fmaf(gm, r[1], s1);
            }
            s0 = warp_sum(s0);
            s1 = warp_sum(s1);
            if (lane == 0) {
                Ss[(b * groups + g) * 2] = s0;
                Ss[(b * groups + g) * 2 + 1] = s1;
                if (blockIdx.x == 0) {
                    S[(b * groups + g) * 2] = s0;
                    S[(b * groups + g) * 2 + 1] = s1;
                }
            }
        }
    }
    __syncthreads();
    const int c = blockIdx.x * 32 + lane;
    float dgm = 0.f, dbt = 0.f, dcs = 0.f;
    if (c < C) {
        const int g = c / cg;
        const float inv_n = 1.f / ((float)cg * (float)T);
        const float gm = gamma[c];
        const bool want = dpre_bsum || dpre_csum || dpre_csum2;
        for (int b = w; b < B; b += 16) {
            const float* r = R + ((int64_t)b * C + c) * 3;
            const float r0 = r[0], r1 = r[1];
            dbt += r0;
            dgm += r1;
            if (want) {
                const float* sp = Ss + (b * groups + g) * 2;
                const float v = rstd[b * groups + g] * (gm * r0 - (float)T * sp[0] * inv_n - r[2] * sp[1] * inv_n);
                if (dpre_bsum) dpre_bsum[(int64_t)b * C + c] = v;
                dcs += v;
            }
        }
    }
    red[0][w][lane] = dgm;
    red[1][w][lane] = dbt;
    red[2][w][lane] = dcs;
    __syncthreads();
    if (w < 3 && c < C) {
        float t = 0.f;
#pragma unroll
        for (int j = 0; j < 16; j++) t += red[w][j][lane];
        if (w == 0) {
            if (dgamma) dgamma[c] = t;
        } else if (w == 1) {
            if (dbeta) dbeta[c] = t;
            if (dbeta2) dbeta2[c] = t;
        } else {
            if (dpre_csum) dpre_csum[c] = t;
            if (dpre_csum2) dpre_csum2[c] = t;
        }
    }
}

template <int VEC>
__global__ void gn_bwd_apply_kernel(const float* __restrict__ dn, const float* __restrict__ h,
                                    const float* __restrict__ mean, const float* __restrict__ rstd,
                                    const float* __restrict__ gamma, const float* __restrict__ S, int64_t rows, int C,
                                    int T, int groups, int Cp, int P, bf16* __restrict__ dpre_planes,
                                    bf16* __restrict__ dn_planes) {
    pdl_wait();
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

// da planes (rows, (2C)p) from dg (rows, C), a (rows, 2C).  A thread takes VEC channels c < C and produces BOTH
// halves (tanh half at c, sigmoid half at C + c) from one tanh / exp pair and vector loads (the first version
// recomputed the pair for each half with scalar loads: 14.7 us for 12 MB in, 4 MB out).
template <int VEC>
__global__ void gate_bwd_apply_kernel(const float* __restrict__ dg, const float* __restrict__ a,
                                      const float* __restrict__ mean, const float* __restrict__ rstd,
                                      const float* __restrict__ gamma, const float* __restrict__ beta,
                                      const float* __restrict__ S, int64_t rows, int C, int T, int Cp, int P,
                                      bf16* __restrict__ da_planes) {
    pdl_wait();
    const int Cv = C / VEC;
    const int64_t n = rows * Cv, plane = rows * Cp;
    const float inv_n = 1.f / ((float)C * (float)T);
    GRID_STRIDE(i, n) {
        const int64_t r = i / Cv;
        const int c = (int)(i - r * Cv) * VEC;
        const int b = (int)(r / T);
        const float m1 = mean[2 * b], m2 = mean[2 * b + 1], rs1 = rstd[2 * b], rs2 = rstd[2 * b + 1];
        const float S10 = S[4 * b] * inv_n, S11 = S[4 * b + 1] * inv_n, S20 = S[4 * b + 2] * inv_n, S21 = S[4 * b + 3] * inv_n;
        float a1[VEC], a2[VEC], d[VEC], g1[VEC], g2[VEC], b1[VEC], b2[VEC], v1[VEC], v2[VEC];
        ld_vec<VEC>(a + r * 2 * C + c, a1);
        ld_vec<VEC>(a + r * 2 * C + C + c, a2);
        ld_vec<VEC>(dg + r * C + c, d);
        ld_vec<VEC>(gamma + c, g1);
        ld_vec<VEC>(gamma + C + c, g2);
        ld_vec<VEC>(beta + c, b1);
        ld_vec<VEC>(beta + C + c, b2);
#pragma unroll
        for (int j = 0; j < VEC; j++) {
            const float x1 = (a1[j] - m1) * rs1, x2 = (a2[j] - m2) * rs2;
            const float th = tanhf(x1 * g1[j] + b1[j]);
            const float sg = sigmoidf_(x2 * g2[j] + b2[j]);
            const float dn1 = d[j] * sg * (1.f - th * th), dn2 = d[j] * th * sg * (1.f - sg);
            v1[j] = rs1 * (g1[j] * dn1 - S10 - x1 * S11);
            v2[j] = rs2 * (g2[j] * dn2 - S20 - x2 * S21);
        }
        st_planes_vec<VEC>(da_planes, r * Cp + c, plane, P, v1);
        st_planes_vec<VEC>(da_planes, r * Cp + C + c, plane, P, v2);
    }
}

__global__ void act_bwd_add_kernel(const float* __restrict__ du, const float* __restrict__ c,
                                   const float* __restrict__ add, int64_t n, int act, float scale,
                                   float* __restrict__ out) {
    pdl_wait();
    GRID_STRIDE(i, n) {
        float v = scale * du[i] * act_grad(c[i] * scale, act);
        if (add) v += add[i];
        out[i] = v;
    }
}

// out[c] += sum over a slice of the B*T frame rows; grid (ceil(C/32), row slices), block 32 x 8; `out` zeroed by
// the host wrapper.  (One 1024-thread block per 32 channels used 16 SMs and took 11 us for 4 MB.)
__global__ void __launch_bounds__(256) channel_sum_kernel(const float* __restrict__ x, int64_t rows, int C,
                                                          float* __restrict__ out) {
    pdl_wait();
    __shared__ float red[8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + tx;
    const int64_t per = (rows + gridDim.y - 1) / gridDim.y;
    const int64_t r0 = per * blockIdx.y, r1 = r0 + per < rows ? r0 + per : rows;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    if (c < C) {
        int64_t r = r0 + ty;
        for (; r + 24 < r1; r += 32) {          // four independent loads in flight
            s0 += x[r * C + c];
            s1 += x[(r + 8) * C + c];
            s2 += x[(r + 16) * C + c];
            s3 += x[(r + 24) * C + c];
        }
        for (; r < r1; r += 8) s0 += x[r * C + c];
    }
    red[ty][tx] = (s0 + s1) + (s2 + s3);
    __syncthreads();
    if (ty == 0 && c < C) {
        float t = 0.f;
#pragma unroll
        for (int j = 0; j < 8; j++) t += red[j][tx];
        atomicAdd(out + c, t);
    }
}

// Wide blocks for the row-sum kernels (32 / 16 warps: VNB_GN_ROWS_NW32=1, VNB_GATE_ROWS_NW16=1) are 1.7 / 1.2 us faster per
// launch on one GPU (9.6 -> 7.9, 10.8 -> 9.6 us) but cost the 2-GPU step 0.1 ms (3.19 -> 3.29 ms): blocks that fill an SM's
// thread slots keep the overlapped all-reduce's CTAs off the SMs.  Default: 8 warps.
static inline bool gn_rows_nw32() {
    static const int v = getenv("VNB_GN_ROWS_NW32") ? atoi(getenv("VNB_GN_ROWS_NW32")) : 0;
    return v != 0;
}
}  // namespace vnb
using namespace vnb;

extern "C" int vnb_gn_bwd_rows(const float* dn, const float* h, const float* mean, const float* rstd, int B, int C,
                               int T, int groups, float* R, void* stream) {
    VNB_REQUIRE(dn && h && mean && rstd && R && B > 0 && B <= 65535 && C > 0 && T > 0 && groups > 0 && C % groups == 0,
                "gn_bwd_rows: bad arguments");
    if (T >= 128 && gn_rows_nw32())
        vnb::launch_k(bwd_rows_kernel<false, 32>, dim3(ceil_div(C, 32), B), 1024, 0, (cudaStream_t)stream, dn, h, mean, rstd, nullptr, nullptr,
                                                                                       C, T, groups, R, nullptr, nullptr, nullptr);
    else
        vnb::launch_k(bwd_rows_kernel<false, 8>, dim3(ceil_div(C, 32), B), 256, 0, (cudaStream_t)stream, dn, h, mean, rstd, nullptr, nullptr,
                                                                                       C, T, groups, R, nullptr, nullptr, nullptr);
    VNB_LAUNCH_CHECK("gn_bwd_rows");
    return VNB_OK;
}

extern "C" int vnb_gn_bwd_rows_s(const float* dn, const float* h, const float* mean, const float* rstd, const float* gamma,
                                 int B, int C, int T, int groups, float* R, double* S64, float* S, void* stream) {
    VNB_REQUIRE(dn && h && mean && rstd && gamma && R && S64 && S && B > 0 && B <= 65535 && C > 0 && T > 0 && groups > 0 &&
                    C % groups == 0 && ((C / groups) & 31) == 0, "gn_bwd_rows_s: bad arguments (32 must divide C / groups)");
    // S64: B * groups * 2 ZEROED doubles followed by one zeroed 8-byte ticket slot
    if (T >= 128 && gn_rows_nw32())
        vnb::launch_k(bwd_rows_kernel<false, 32>, dim3(ceil_div(C, 32), B), 1024, 0, (cudaStream_t)stream, dn, h, mean, rstd, gamma, nullptr,
                  C, T, groups, R, S64, S, reinterpret_cast<unsigned*>(S64 + (size_t)B * groups * 2));
    else
        vnb::launch_k(bwd_rows_kernel<false, 8>, dim3(ceil_div(C, 32), B), 256, 0, (cudaStream_t)stream, dn, h, mean, rstd, gamma, nullptr,
                  C, T, groups, R, S64, S, reinterpret_cast<unsigned*>(S64 + (size_t)B * groups * 2));
    VNB_LAUNCH_CHECK("gn_bwd_rows_s");
    return VNB_OK;
}

namespace vnb {
// Gate variant of bwd_rows: a thread owns channel c < C and accumulates the three row sums of BOTH halves
// (tanh half c, sigmoid half C + c) from one tanh / exp pair per element, four frames in flight.
// grid (ceil(C/32), B), block 32 x NW (8 by default; 16 with VNB_GATE_ROWS_NW16=1: 10.8 -> 9.6 us at 16 x 128 frames, but see
// gn_rows_nw32() below for why it is off; 32 warps - 64 registers per thread around libm tanh / exp - measured slower).
template <bool FAST, int NW>
__global__ void __launch_bounds__(32 * NW) gate_bwd_rows_kernel(const float* __restrict__ dg, const float* __restrict__ a,
                                                            const float* __restrict__ mean, const float* __restrict__ rstd,
                                                            const float* __restrict__ gamma, const float* __restrict__ beta,
                                                            int C, int T, float* __restrict__ R, double* __restrict__ S64,
                                                            float* __restrict__ S32, unsigned* __restrict__ counter) {
    pdl_wait();
    __shared__ float red[6][NW][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int b = blockIdx.y, c = blockIdx.x * 32 + tx;
    float s[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (c < C) {
        const float m1 = mean[2 * b], m2 = mean[2 * b + 1], rs1 = rstd[2 * b], rs2 = rstd[2 * b + 1];
        const float g1 = gamma[c], g2 = gamma[C + c], b1 = beta[c], b2 = beta[C + c];
        for (int t0 = ty; t0 < T; t0 += 4 * NW) {
            float a1[4], a2[4], d[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int t = t0 + NW * u;
                const int64_t r = (int64_t)b * T + (t < T ? t : 0);
                a1[u] = a[r * 2 * C + c];
                a2[u] = a[r * 2 * C + C + c];
                d[u] = t < T ? dg[r * C + c] : 0.f;
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
                if (t0 + NW * u >= T) continue;
                const float x1 = (a1[u] - m1) * rs1, x2 = (a2[u] - m2) * rs2;
                const float th = gate_tanh<FAST>(x1 * g1 + b1);
                const float sg = gate_sigmoid<FAST>(x2 * g2 + b2);
                const float dn1 = d[u] * sg * (1.f - th * th), dn2 = d[u] * th * sg * (1.f - sg);
                s[0] += dn1; s[1] += dn1 * x1; s[2] += x1;
                s[3] += dn2; s[4] += dn2 * x2; s[5] += x2;
            }
        }
    }
#pragma unroll
    for (int k = 0; k < 6; k++) red[k][ty][tx] = s[k];
    __syncthreads();
    float t = 0.f;
    const int ch = ty < 3 ? c : C + c;
    if (ty < 6 && c < C) {
#pragma unroll
        for (int j = 0; j < NW; j++) t += red[ty][j][tx];
        R[((int64_t)b * 2 * C + ch) * 3 + (ty % 3)] = t;
    }
    if (S64) {
        // warps 0..2 hold the tanh half (group 0), 3..5 the sigmoid half (group 1); warps 6 and up only take part in the barrier
        const int k = ty < 6 ? ty % 3 : 2;
        publish_group_sums(t, (c < C && ty < 6 && k < 2) ? gamma[ch] : 0.f, c < C && ty < 6, k, b * 2 + (ty >= 3 && ty < 6 ? 1 : 0), S64, S32,
                           counter, gridDim.y * 4);
    }
}
}  // namespace vnb

extern "C" int vnb_gate_bwd_rows_p(const float* dg, const float* a, const float* mean, const float* rstd,
                                   const float* gamma, const float* beta, int B, int C, int T, int P, float* R,
                                   void* stream) {
    VNB_REQUIRE(dg && a && mean && rstd && gamma && beta && R && B > 0 && B <= 65535 && C > 0 && T > 0 &&
                    (P >= 1 && P <= 3), "gate_bwd_rows: bad arguments");
    // the gate's approximate functions only where the operand planes are bf16 anyway (see ew_v2.cuh)
    static const int nw16_env = getenv("VNB_GATE_ROWS_NW16") ? atoi(getenv("VNB_GATE_ROWS_NW16")) : 0;
    const bool nw16 = nw16_env && T >= 64;
    if (P == 1 && ew_v2_enabled() && fast_gate_enabled())
        (nw16 ? vnb::launch_k(gate_bwd_rows_kernel<true, 16>, dim3(ceil_div(C, 32), B), 512, 0, (cudaStream_t)stream, dg, a, mean, rstd, gamma, beta, C, T, R,
                      nullptr, nullptr, nullptr)
              : vnb::launch_k(gate_bwd_rows_kernel<true, 8>, dim3(ceil_div(C, 32), B), 256, 0, (cudaStream_t)stream, dg, a, mean, rstd, gamma, beta, C, T, R,
                      nullptr, nullptr, nullptr));
    else
        (nw16 ? vnb::launch_k(gate_bwd_rows_kernel<false, 16>, dim3(ceil_div(C, 32), B), 512, 0, (cudaStream_t)stream, dg, a, mean, rstd, gamma, beta, C, T, R,
                      nullptr, nullptr, nullptr)
              : vnb::launch_k(gate_bwd_rows_kernel<false, 8>, dim3(ceil_div(C, 32), B), 256, 0, (cudaStream_t)stream, dg, a, mean, rstd, gamma, beta, C, T, R,
                      nullptr, nullptr, nullptr));
    VNB_LAUNCH_CHECK("gate_bwd_rows");
    return VNB_OK;
}

extern "C" int vnb_gate_bwd_rows_s(const float* dg, const float* a, const float* mean, const float* rstd,
                                   const float* gamma, const float* beta, int B, int C, int T, int P, float* R, double* S64,
                                   float* S, void* stream) {
    VNB_REQUIRE(dg && a && mean && rstd && gamma && beta && R && S64 && S && B > 0 && B <= 65535 && C > 0 && (C & 31) == 0 &&
                    T > 0 && (P >= 1 && P <= 3), "gate_bwd_rows_s: bad arguments (32 must divide C)");
    unsigned* counter = reinterpret_cast<unsigned*>(S64 + (size_t)B * 4);
    static const int nw16_env = getenv("VNB_GATE_ROWS_NW16") ? atoi(getenv("VNB_GATE_ROWS_NW16")) : 0;
    const bool nw16 = nw16_env && T >= 64;
    if (P == 1 && ew_v2_enabled() && fast_gate_enabled())
        (nw16 ? vnb::launch_k(gate_bwd_rows_kernel<true, 16>, dim3(ceil_div(C, 32), B), 512, 0, (cudaStream_t)stream, dg, a, mean, rstd, gamma, beta, C, T, R,
                      S64, S, counter)
              : vnb::launch_k(gate_bwd_rows_kernel<true, 8>, dim3(ceil_div(C, 32), B), 256, 0, (cudaStream_t)stream, dg, a, mean, rstd, gamma, beta, C, T, R,
                      S64, S, counter));
    else
        (nw16 ? vnb::launch_k(gate_bwd_rows_kernel<false, 16>, dim3(ceil_div(C, 32), B), 512, 0, (cudaStream_t)stream, dg, a, mean, rstd, gamma, beta, C, T, R,
                      S64, S, counter)
              : vnb::launch_k(gate_bwd_rows_kernel<false, 8>, dim3(ceil_div(C, 32), B), 256, 0, (cudaStream_t)stream, dg, a, mean, rstd, gamma, beta, C, T, R,
                      S64, S, counter));
    VNB_LAUNCH_CHECK("gate_bwd_rows_s");
    return VNB_OK;
}

extern "C" int vnb_gate_bwd_rows(const float* dg, const float* a, const float* mean, const float* rstd,
                                 const float* gamma, const float* beta, int B, int C, int T, float* R, void* stream) {
    return vnb_gate_bwd_rows_p(dg, a, mean, rstd, gamma, beta, B, C, T, 3, R, stream);
}

extern "C" int vnb_gn_bwd_finalize(const float* R, const float* gamma, const float* rstd, int B, int C, int T,
                                   int groups, float* dgamma, float* dbeta, float* S, float* dpre_bsum,
                                   float* dpre_csum, float* dbeta2, float* dpre_csum2, void* stream) {
    VNB_REQUIRE(R && gamma && rstd && S && B > 0 && C > 0 && groups > 0 && C % groups == 0, "gn_bwd_finalize: bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    const size_t smem = ((size_t)((B * groups * 2 + 3) & ~3) + 3 * 16 * 33) * sizeof(float);
    VNB_REQUIRE(smem <= 48 * 1024, "gn_bwd_finalize: B*groups = %d too large", B * groups);
    vnb::launch_k(gn_bwd_finalize_kernel, ceil_div(C, 32), 512, smem, st, R, gamma, rstd, B, C, T, groups, dgamma, dbeta, S, dpre_bsum,
                                                             dpre_csum, dbeta2, dpre_csum2);
    VNB_LAUNCH_CHECK("gn_bwd_finalize");
    return VNB_OK;
}

#define VEC_DISPATCH(cond4, KERNEL, n_scalar, st, ...)                                        \
    do {                                                                                      \
        if (cond4)                                                                            \
            vnb::launch_k(KERNEL<4>, ew_grid((n_scalar) / 4), EW_THREADS, 0, st, __VA_ARGS__);           \
        else                                                                                  \
            vnb::launch_k(KERNEL<1>, ew_grid(n_scalar), EW_THREADS, 0, st, __VA_ARGS__);                 \
    } while (0)

extern "C" int vnb_gn_bwd_apply(const float* dn, const float* h, const float* mean, const float* rstd,
                                const float* gamma, const float* S, int B, int C, int T, int groups, int P,
                                void* dpre_planes, void* dn_planes, void* stream) {
    VNB_REQUIRE(dn && h && mean && rstd && gamma && S && dpre_planes, "gn_bwd_apply: null pointer");
    VNB_REQUIRE(B > 0 && C > 0 && T > 0 && groups > 0 && C % groups == 0 && (P >= 1 && P <= 3), "gn_bwd_apply: bad shape");
    const int64_t rows = (int64_t)B * T;
    if ((C & 3) == 0 && ((C / groups) & 3) == 0 && ew_v2_enabled()) {
        const int g2 = v2_grid(rows, C / 4);
        if (g2) {
            cudaStream_t st = (cudaStream_t)stream;
            if (P == 1) vnb::launch_k(gn_bwd_apply_v2_kernel<1, 2>, g2, V2_THREADS, 0, st, dn, h, mean, rstd, gamma, S, (int)rows, C, T, groups, round_up8(C), (bf16*)dpre_planes, (bf16*)dn_planes);
            else if (P == 2) vnb::launch_k(gn_bwd_apply_v2_kernel<2, 2>, g2, V2_THREADS, 0, st, dn, h, mean, rstd, gamma, S, (int)rows, C, T, groups, round_up8(C), (bf16*)dpre_planes, (bf16*)dn_planes);
            else vnb::launch_k(gn_bwd_apply_v2_kernel<3, 2>, g2, V2_THREADS, 0, st, dn, h, mean, rstd, gamma, S, (int)rows, C, T, groups, round_up8(C), (bf16*)dpre_planes, (bf16*)dn_planes);
            VNB_LAUNCH_CHECK("gn_bwd_apply_v2");
            return VNB_OK;
        }
    }
    VEC_DISPATCH((C & 3) == 0, gn_bwd_apply_kernel, rows * C, (cudaStream_t)stream, dn, h, mean, rstd, gamma, S, rows, C, T,
                 groups, round_up8(C), P, (bf16*)dpre_planes, (bf16*)dn_planes);
    VNB_LAUNCH_CHECK("gn_bwd_apply");
    return VNB_OK;
}

extern "C" int vnb_gate_bwd_apply(const float* dg, const float* a, const float* mean, const float* rstd,
                                  const float* gamma, const float* beta, const float* S, int B, int C, int T, int P,
                                  void* da_planes, void* stream) {
    VNB_REQUIRE(dg && a && mean && rstd && gamma && beta && S && da_planes, "gate_bwd_apply: null pointer");
    VNB_REQUIRE(B > 0 && C > 0 && T > 0 && (P >= 1 && P <= 3), "gate_bwd_apply: bad shape");
    const int64_t rows = (int64_t)B * T;
    if ((C & 3) == 0 && ew_v2_enabled()) {
        const int g2 = v2_grid(rows, C / 4);
        if (g2) {
            cudaStream_t st = (cudaStream_t)stream;
            const int Cp = round_up8(2 * C);
            if (P == 1 && fast_gate_enabled()) vnb::launch_k(gate_bwd_apply_v2_kernel<1, 2, true>, g2, V2_THREADS, 0, st, dg, a, mean, rstd, gamma, beta, S, (int)rows, C, T, Cp, (bf16*)da_planes);
            else if (P == 1) vnb::launch_k(gate_bwd_apply_v2_kernel<1, 2, false>, g2, V2_THREADS, 0, st, dg, a, mean, rstd, gamma, beta, S, (int)rows, C, T, Cp, (bf16*)da_planes);
            else if (P == 2) vnb::launch_k(gate_bwd_apply_v2_kernel<2, 2, false>, g2, V2_THREADS, 0, st, dg, a, mean, rstd, gamma, beta, S, (int)rows, C, T, Cp, (bf16*)da_planes);
            else vnb::launch_k(gate_bwd_apply_v2_kernel<3, 2, false>, g2, V2_THREADS, 0, st, dg, a, mean, rstd, gamma, beta, S, (int)rows, C, T, Cp, (bf16*)da_planes);
            VNB_LAUNCH_CHECK("gate_bwd_apply_v2");
            return VNB_OK;
        }
    }
    VEC_DISPATCH((C & 3) == 0, gate_bwd_apply_kernel, rows * C, (cudaStream_t)stream, dg, a, mean, rstd, gamma, beta, S,
                 rows, C, T, round_up8(2 * C), P, (bf16*)da_planes);
    VNB_LAUNCH_CHECK("gate_bwd_apply");
    return VNB_OK;
}

extern "C" int vnb_act_bwd_add(const float* du, const float* c, const float* add, int64_t n, int act, float scale,
                               float* out, void* stream) {
    VNB_REQUIRE(du && c && out && n > 0, "act_bwd_add: bad arguments");
    vnb::launch_k(act_bwd_add_kernel, ew_grid(n), EW_THREADS, 0, (cudaStream_t)stream, du, c, add, n, act, scale, out);
    VNB_LAUNCH_CHECK("act_bwd_add");
    return VNB_OK;
}

extern "C" int vnb_channel_sum(const float* x, int B, int C, int T, float* out, void* stream) {
    VNB_REQUIRE(x && out && B > 0 && C > 0 && T > 0, "channel_sum: bad arguments");
    const int64_t rows = (int64_t)B * T;
    int slices = (int)(rows / 128);
    slices = slices < 1 ? 1 : slices > 32 ? 32 : slices;
    if (cudaMemsetAsync(out, 0, (size_t)C * sizeof(float), (cudaStream_t)stream) != cudaSuccess) {
        set_error("channel_sum: memset failed");
        return VNB_E_CUDA;
    }
    vnb::launch_k(channel_sum_kernel, dim3(ceil_div(C, 32), slices), 256, 0, (cudaStream_t)stream, x, rows, C, out);
    VNB_LAUNCH_CHECK("channel_sum");
    return VNB_OK;
}
