// Memory-bound fused kernels around the GEMMs (forward side): layout conversion, operand-plane
// writers, GroupNorm statistics, encoder-block epilogue, decoder gate, residual/skip bookkeeping.
// Everything is CHANNELS-LAST: activation element (b, c, t) lives at (b*T + t)*C + c, operand
// planes at (b*T + t)*Cp + c, so consecutive threads touch consecutive channels (float4 / 4 x bf16
// per thread when C % 4 == 0) and no kernel transposes.
// Reference op sequences: vae_npvc/model/layers.py:151-155,168-176 (encoder residual block),
// :229-247 (gated decoder block); math in SURVEY.md section 9 A/B.
#include "common.cuh"
#include "ew_v2.cuh"

namespace vnb {

constexpr int EW_THREADS = 256;
static inline int ew_grid(int64_t n) {
    int64_t g = (n + EW_THREADS - 1) / EW_THREADS;
    const int64_t cap = 148 * 8;           // grid-stride beyond 8 CTAs per SM
    return (int)(g < cap ? g : cap);
}
#define GRID_STRIDE(i, n) \
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < (n); i += (int64_t)gridDim.x * blockDim.x)

// launch a VEC-templated kernel with VEC = 4 when every channel count involved is a multiple of 4
#define VEC_DISPATCH(cond4, KERNEL, n_scalar, st, ...)                                        \
    do {                                                                                      \
        if (cond4)                                                                            \
            vnb::launch_k(KERNEL<4>, ew_grid((n_scalar) / 4), EW_THREADS, 0, st, __VA_ARGS__);           \
        else                                                                                  \
            vnb::launch_k(KERNEL<1>, ew_grid(n_scalar), EW_THREADS, 0, st, __VA_ARGS__);                 \
    } while (0)

// ------------------------------------------------------------------------------------------
// (B, R, S) -> (B, S, R), 32x32 shared-memory tiles (both sides coalesced)
__global__ void __launch_bounds__(256) transpose_kernel(const float* __restrict__ in, int R, int S,
                                                        float* __restrict__ out) {
    pdl_wait();
    __shared__ float tile[32][33];
    const int64_t base = (int64_t)blockIdx.z * R * S;
    const int s0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int j = ty; j < 32; j += 8) {
        const int r = r0 + j, s = s0 + tx;
        if (r < R && s < S) tile[j][tx] = in[base + (int64_t)r * S + s];
    }
    __syncthreads();
    for (int j = ty; j < 32; j += 8) {
        const int s = s0 + j, r = r0 + tx;
        if (r < R && s < S) out[base + (int64_t)s * R + r] = tile[tx][j];
    }
}

// ------------------------------------------------------------------------------------------
template <int VEC>
__global__ void make_planes_kernel(const float* __restrict__ x, int64_t rows, int C, int Cp, int P, float scale,
                                   int act, bf16* __restrict__ out_id, bf16* __restrict__ out_act,
                                   float* __restrict__ csum) {
    pdl_wait();
    extern __shared__ float csum_s[];                 // [C] when csum != nullptr: per-channel sums of x*scale
    if (csum) {
        for (int c = threadIdx.x; c < C; c += blockDim.x) csum_s[c] = 0.f;
        __syncthreads();
    }
    const int Cv = C / VEC;
    const int64_t n = rows * Cv, plane = rows * Cp;
    GRID_STRIDE(i, n) {
        const int64_t r = i / Cv;
        const int c = (int)(i - r * Cv) * VEC;
        float v[VEC], u[VEC];
        ld_vec<VEC>(x + r * C + c, v);
#pragma unroll
        for (int j = 0; j < VEC; j++) {
            v[j] *= scale;
            u[j] = apply_act(v[j], act);
        }
        const int64_t o = r * Cp + c;
        if (out_id) st_planes_vec<VEC>(out_id, o, plane, P, v);
        if (out_act) st_planes_vec<VEC>(out_act, o, plane, P, u);
        if (csum) {
#pragma unroll
            for (int j = 0; j < VEC; j++) atomicAdd(csum_s + c + j, v[j]);
        }
    }
    if (csum) {
        __syncthreads();
        for (int c = threadIdx.x; c < C; c += blockDim.x) atomicAdd(csum + c, csum_s[c]);
    }
}

template <int VEC>
__global__ void concat_planes_kernel(const float* __restrict__ xa, int Ca, const float* __restrict__ xb, int Cb,
                                     int64_t rows, int Cp, int P, bf16* __restrict__ out, float* __restrict__ csum) {
    pdl_wait();
    extern __shared__ float csum_s[];                 // [Ca+Cb] when csum != nullptr
    const int Cc = Ca + Cb, Cv = Cc / VEC;
    if (csum) {
        for (int c = threadIdx.x; c < Cc; c += blockDim.x) csum_s[c] = 0.f;
        __syncthreads();
    }
    const int64_t n = rows * Cv, plane = rows * Cp;
    GRID_STRIDE(i, n) {
        const int64_t r = i / Cv;
        const int c = (int)(i - r * Cv) * VEC;
        float v[VEC];
        if (c < Ca) {
            if (xa) ld_vec<VEC>(xa + r * Ca + c, v);
            else {
#pragma unroll
                for (int j = 0; j < VEC; j++) v[j] = 0.f;       // null first operand = zeros
            }
        } else ld_vec<VEC>(xb + r * Cb + (c - Ca), v);
        st_planes_vec<VEC>(out, r * Cp + c, plane, P, v);
        if (csum) {
#pragma unroll
            for (int j = 0; j < VEC; j++) atomicAdd(csum_s + c + j, v[j]);
        }
    }
    if (csum) {
        __syncthreads();
        for (int c = threadIdx.x; c < Cc; c += blockDim.x) atomicAdd(csum + c, csum_s[c]);
    }
}

// Second generation of concat_planes for the case WITH channel sums (the res/skip bias gradient, 10 launches per step):
// the thread count is a multiple of the channel-vector count, so a thread's four channels never change - it walks
// its rows with four loads in flight, keeps the sums in registers and touches the shared / global accumulators ONCE
// (the first generation did a shared atomic per element and 296 x C global atomics: 8 us at 16 x 128 frames).
template <int P>
__global__ void __launch_bounds__(256) concat_planes_v2_kernel(const float* __restrict__ xa, int Ca, const float* __restrict__ xb,
                                                               int Cb, int rows, int Cp, bf16* __restrict__ out,
                                                               float* __restrict__ csum) {
    pdl_wait();
    extern __shared__ float csum_s[];                 // [Ca + Cb]
    const int Cc = Ca + Cb, Cv = Cc >> 2;
    for (int c = threadIdx.x; c < Cc; c += blockDim.x) csum_s[c] = 0.f;
    __syncthreads();
    const int tid = blockIdx.x * blockDim.x + threadIdx.x, nthr = gridDim.x * blockDim.x;
    const int c = (tid % Cv) * 4, r0 = tid / Cv, rstep = nthr / Cv;
    const float* __restrict__ src = c < Ca ? (xa ? xa + c : nullptr) : xb + (c - Ca);      // xa null: zeros
    const int stride = c < Ca ? Ca : Cb;
    const int64_t plane = (int64_t)rows * Cp;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int r = r0; r < rows; r += 4 * rstep) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int ru = r + u * rstep;
            v[u] = (ru < rows && src) ? __ldg(reinterpret_cast<const float4*>(src + (int64_t)ru * stride)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int ru = r + u * rstep;
            if (ru >= rows) break;
            const float x[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
            st_planes4<P>(out, (int64_t)ru * Cp + c, plane, x);
            acc[0] += x[0]; acc[1] += x[1]; acc[2] += x[2]; acc[3] += x[3];
        }
    }
#pragma unroll
    for (int j = 0; j < 4; j++) atomicAdd(csum_s + c + j, acc[j]);
    __syncthreads();
    for (int cc = threadIdx.x; cc < Cc; cc += blockDim.x) atomicAdd(csum + cc, csum_s[cc]);
}

// ------------------------------------------------------------------------------------------
// GroupNorm statistics, two launches.  Pass 1: grid (chunks, B); each CTA owns `rows_per` frames
// of one sample and writes, per group, (local mean, local M2) of its slice - a CTA-local two-pass,
// so there is no E[x^2]-E[x]^2 cancellation.  Pass 2 merges the chunks with Chan's formula.
constexpr int GN_MAX_CHUNKS = 64;
constexpr int GN_MAX_GROUPS = 8;

__global__ void __launch_bounds__(256) gn_stats_partial_kernel(const float* __restrict__ h, int C, int T, int groups,
                                                               int rows_per, float* __restrict__ ws) {
    pdl_wait();
    __shared__ float acc[GN_MAX_GROUPS];
    __shared__ float lmean[GN_MAX_GROUPS];
    const int b = blockIdx.y, chunk = blockIdx.x, chunks = gridDim.x;
    const int t0 = chunk * rows_per;
    const int t1 = min(T, t0 + rows_per);
    const int nrow = max(0, t1 - t0);
    const int cg = C / groups;
    const float* p = h + ((int64_t)b * T + t0) * C;
    const int n = nrow * C;
    for (int pass = 0; pass < 2; pass++) {
        if (threadIdx.x < GN_MAX_GROUPS) acc[threadIdx.x] = 0.f;
        __syncthreads();
        // a thread's channel advances by blockDim mod C per step; keep a per-group register pair
        float s[2] = {0.f, 0.f};
        int cur = -1;
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            const int g = (i % C) / cg;
            if (g != cur) {
                if (cur >= 0) atomicAdd(&acc[cur], s[0]);
                s[0] = 0.f;
                cur = g;
            }
            const float x = p[i];
            if (pass == 0) s[0] += x;
            else {
                const float d = x - lmean[g];
                s[0] += d * d;
            }
        }
        if (cur >= 0) atomicAdd(&acc[cur], s[0]);
        __syncthreads();
        if (threadIdx.x < groups) {
            const float cnt = (float)nrow * (float)cg;
            float* o = ws + (((int64_t)b * groups + threadIdx.x) * chunks + chunk) * 2;
            if (pass == 0) lmean[threadIdx.x] = cnt > 0.f ? acc[threadIdx.x] / cnt : 0.f;
            else {
                o[0] = lmean[threadIdx.x];
                o[1] = acc[threadIdx.x];
            }
        }
        __syncthreads();
    }
}

// one warp per (b, group)
__global__ void gn_stats_merge_kernel(const float* __restrict__ ws, int BG, int chunks, int T, int rows_per, int cg,
                                      float eps, float* __restrict__ mean, float* __restrict__ rstd) {
    pdl_wait();
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (w >= BG) return;
    const float* p = ws + (int64_t)w * chunks * 2;
    float sm = 0.f;
    for (int c = lane; c < chunks; c += 32) {
        const int nrow = max(0, min(T, (c + 1) * rows_per) - c * rows_per);
        sm += p[2 * c] * (float)nrow * (float)cg;
    }
    const float n = (float)T * (float)cg;
    const float m = warp_sum(sm) / n;
    float m2 = 0.f;
    for (int c = lane; c < chunks; c += 32) {
        const int nrow = max(0, min(T, (c + 1) * rows_per) - c * rows_per);
        const float d = p[2 * c] - m;
        m2 += p[2 * c + 1] + (float)nrow * (float)cg * d * d;
    }
    m2 = warp_sum(m2);
    if (lane == 0) {
        mean[w] = m;
        rstd[w] = rsqrtf(m2 / n + eps);
    }
}

// mean / rstd from the fp64 (sum, sum of squares) pairs accumulated by the conv epilogue
__global__ void gn_finalize_kernel(const double* __restrict__ sums, int BG, double n, float eps,
                                   float* __restrict__ mean, float* __restrict__ rstd) {
    pdl_wait();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= BG) return;
    const double m = sums[2 * i] / n;
    double var = sums[2 * i + 1] / n - m * m;
    if (var < 0.0) var = 0.0;
    mean[i] = (float)m;
    rstd[i] = (float)(1.0 / sqrt(var + (double)eps));
}

// ------------------------------------------------------------------------------------------
// encoder block epilogue: y = GN1(h)*gamma+beta (+ s) -> y fp32, planes(y), planes(lrelu(y))
template <int VEC>
__global__ void enc_out_kernel(const float* __restrict__ h, const float* __restrict__ s, const float* __restrict__ mean,
                               const float* __restrict__ rstd, const float* __restrict__ gamma,
                               const float* __restrict__ beta, int64_t rows, int C, int T, int Cp, int P,
                               float* __restrict__ y, bf16* __restrict__ y_planes, bf16* __restrict__ u_planes) {
    pdl_wait();
    const int Cv = C / VEC;
    const int64_t n = rows * Cv, plane = rows * Cp;
    GRID_STRIDE(i, n) {
        const int64_t r = i / Cv;
        const int c = (int)(i - r * Cv) * VEC;
        const int b = (int)(r / T);
        const float m = mean[b], rs = rstd[b];
        float v[VEC], gm[VEC], bt[VEC], u[VEC];
        ld_vec<VEC>(h + r * C + c, v);
        ld_vec<VEC>(gamma + c, gm);
        ld_vec<VEC>(beta + c, bt);
#pragma unroll
        for (int j = 0; j < VEC; j++) v[j] = (v[j] - m) * rs * gm[j] + bt[j];
        if (s) {
            float sv[VEC];
            ld_vec<VEC>(s + r * C + c, sv);
#pragma unroll
            for (int j = 0; j < VEC; j++) v[j] += sv[j];
        }
#pragma unroll
        for (int j = 0; j < VEC; j++) u[j] = lrelu02(v[j]);
        if (y) st_vec<VEC>(y + r * C + c, v);
        const int64_t o = r * Cp + c;
        if (y_planes) st_planes_vec<VEC>(y_planes, o, plane, P, v);
        if (u_planes) st_planes_vec<VEC>(u_planes, o, plane, P, u);
    }
}

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }

// decoder gate: a (rows, 2C); hn = GN2(a)*gamma+beta; g = tanh(hn[:C]) * sigmoid(hn[C:]) -> planes (rows, Cp)
template <int VEC>
__global__ void dec_gate_kernel(const float* __restrict__ a, const float* __restrict__ mean,
                                const float* __restrict__ rstd, const float* __restrict__ gamma,
                                const float* __restrict__ beta, int64_t rows, int C, int T, int Cp, int P,
                                bf16* __restrict__ g_planes) {
    pdl_wait();
    const int Cv = C / VEC;
    const int64_t n = rows * Cv, plane = rows * Cp;
    GRID_STRIDE(i, n) {
        const int64_t r = i / Cv;
        const int c = (int)(i - r * Cv) * VEC;
        const int b = (int)(r / T);
        const float m1 = mean[2 * b], r1 = rstd[2 * b], m2 = mean[2 * b + 1], r2 = rstd[2 * b + 1];
        float a1[VEC], a2[VEC], g1[VEC], g2[VEC], b1[VEC], b2[VEC], o[VEC];
        ld_vec<VEC>(a + r * 2 * C + c, a1);
        ld_vec<VEC>(a + r * 2 * C + C + c, a2);
        ld_vec<VEC>(gamma + c, g1);
        ld_vec<VEC>(gamma + C + c, g2);
        ld_vec<VEC>(beta + c, b1);
        ld_vec<VEC>(beta + C + c, b2);
#pragma unroll
        for (int j = 0; j < VEC; j++) {
            const float h1 = (a1[j] - m1) * r1 * g1[j] + b1[j];
            const float h2 = (a2[j] - m2) * r2 * g2[j] + b2[j];
            o[j] = tanhf(h1) * sigmoidf_(h2);
        }
        st_planes_vec<VEC>(g_planes, r * Cp + c, plane, P, o);
    }
}

// decoder residual/skip split: r (rows, C+S): x_out = r[:C] + x ; skip_sum (+)= r[C:]
template <int VEC>
__global__ void dec_out_kernel(const float* __restrict__ r, const float* __restrict__ x, int64_t rows, int C, int S,
                               int Cp, int P, float* __restrict__ x_out, bf16* __restrict__ x_planes,
                               float* __restrict__ skip_sum, int first) {
    pdl_wait();
    const int Cc = C + S, Cv = Cc / VEC;
    const int64_t n = rows * Cv, plane = rows * Cp;
    GRID_STRIDE(i, n) {
        const int64_t row = i / Cv;
        const int c = (int)(i - row * Cv) * VEC;
        float v[VEC];
        ld_vec<VEC>(r + row * Cc + c, v);
        if (c < C) {
            float xv[VEC];
            ld_vec<VEC>(x + row * C + c, xv);
#pragma unroll
            for (int j = 0; j < VEC; j++) v[j] += xv[j];
            if (x_out) st_vec<VEC>(x_out + row * C + c, v);
            if (x_planes) st_planes_vec<VEC>(x_planes, row * Cp + c, plane, P, v);
        } else if (skip_sum) {
            float* sp = skip_sum + row * S + (c - C);
            if (!first) {
                float sv[VEC];
                ld_vec<VEC>(sp, sv);
#pragma unroll
                for (int j = 0; j < VEC; j++) v[j] += sv[j];
            }
            st_vec<VEC>(sp, v);
        }
    }
}

}  // namespace vnb
using namespace vnb;

#define ST ((cudaStream_t)stream)

extern "C" int vnb_transpose(const float* in, int B, int R, int S, float* out, void* stream) {
    VNB_REQUIRE(in && out && B > 0 && R > 0 && S > 0 && B <= 65535, "transpose: bad arguments");
    dim3 grid(ceil_div(S, 32), ceil_div(R, 32), B);
    vnb::launch_k(transpose_kernel, grid, 256, 0, ST, in, R, S, out);
    VNB_LAUNCH_CHECK("transpose");
    return VNB_OK;
}

// csum (optional, C floats, ZEROED by the caller): also accumulates the per-channel sums over all frames of the
// values written to the identity plane (x*scale) - the bias gradient, without a second pass over the tensor.
extern "C" int vnb_make_planes(const float* x, int B, int C, int T, int P, float scale, int act, void* out_id,
                               void* out_act, float* csum, void* stream) {
    VNB_REQUIRE(x && (out_id || out_act) && B > 0 && C > 0 && T > 0 && (P >= 1 && P <= 3), "make_planes: bad arguments");
    VNB_REQUIRE(!csum || C <= 12000, "make_planes: too many channels for the fused channel sums");
    const int64_t rows = (int64_t)B * T;
    const bool v4 = (C & 3) == 0;
    if (v4 && !csum && ew_v2_enabled()) {
        const int g2 = v2_grid(rows, C / 4);
        if (g2) {
            if (P == 1) vnb::launch_k(make_planes_v2_kernel<1, 4>, g2, V2_THREADS, 0, ST, x, (int)rows, C, round_up8(C), scale, act, (bf16*)out_id, (bf16*)out_act);
            else if (P == 2) vnb::launch_k(make_planes_v2_kernel<2, 2>, g2, V2_THREADS, 0, ST, x, (int)rows, C, round_up8(C), scale, act, (bf16*)out_id, (bf16*)out_act);
            else vnb::launch_k(make_planes_v2_kernel<3, 2>, g2, V2_THREADS, 0, ST, x, (int)rows, C, round_up8(C), scale, act, (bf16*)out_id, (bf16*)out_act);
            VNB_LAUNCH_CHECK("make_planes_v2");
            return VNB_OK;
        }
    }
    const int64_t n = v4 ? rows * C / 4 : rows * C;
    int grid = ew_grid(n);
    if (csum && grid > 296) grid = 296;                 // fewer flushes of the shared-memory sums
    const size_t smem = csum ? (size_t)C * sizeof(float) : 0;
    if (v4) vnb::launch_k(make_planes_kernel<4>, grid, EW_THREADS, smem, ST, x, rows, C, round_up8(C), P, scale, act, (bf16*)out_id, (bf16*)out_act, csum);
    else vnb::launch_k(make_planes_kernel<1>, grid, EW_THREADS, smem, ST, x, rows, C, round_up8(C), P, scale, act, (bf16*)out_id, (bf16*)out_act, csum);
    VNB_LAUNCH_CHECK("make_planes");
    return VNB_OK;
}

extern "C" int vnb_concat_planes(const float* xa, int Ca, const float* xb, int Cb, int B, int T, int P, void* out,
                                 float* csum, void* stream) {
    VNB_REQUIRE(xb && out && Ca > 0 && Cb > 0 && B > 0 && T > 0 && (P >= 1 && P <= 3), "concat_planes: bad arguments");
    VNB_REQUIRE(!csum || Ca + Cb <= 12000, "concat_planes: too many channels for the fused channel sums");
    const int64_t rows = (int64_t)B * T;
    const bool v4 = ((Ca | Cb) & 3) == 0;
    const int64_t n = v4 ? rows * (Ca + Cb) / 4 : rows * (Ca + Cb);
    if (csum && v4 && ew_v2_enabled() && rows < (1 << 30)) {
        // CTAs: a multiple of Cv / gcd(256, Cv) (thread count divisible by the channel-vector count), about one per SM
        const int Cv = (Ca + Cb) / 4, g0 = Cv / v2_gcd(V2_THREADS, Cv);
        static const int per_sm = getenv("VNB_CONCAT_CTAS_PER_SM") ? atoi(getenv("VNB_CONCAT_CTAS_PER_SM")) : 1;
        if (g0 <= 148) {
            int grid2 = (148 * per_sm / g0) * g0;
            const int64_t need = (n + V2_THREADS - 1) / V2_THREADS;
            if (need < grid2) grid2 = (int)((need + g0 - 1) / g0 * g0);
            const size_t smem2 = (size_t)(Ca + Cb) * sizeof(float);
            const int Cp = round_up8(Ca + Cb);
            if (P == 1) vnb::launch_k(concat_planes_v2_kernel<1>, grid2, V2_THREADS, smem2, ST, xa, Ca, xb, Cb, (int)rows, Cp, (bf16*)out, csum);
            else if (P == 2) vnb::launch_k(concat_planes_v2_kernel<2>, grid2, V2_THREADS, smem2, ST, xa, Ca, xb, Cb, (int)rows, Cp, (bf16*)out, csum);
            else vnb::launch_k(concat_planes_v2_kernel<3>, grid2, V2_THREADS, smem2, ST, xa, Ca, xb, Cb, (int)rows, Cp, (bf16*)out, csum);
            VNB_LAUNCH_CHECK("concat_planes_v2");
            return VNB_OK;
        }
    }
    int grid = ew_grid(n);
    if (csum && grid > 296) grid = 296;
    const size_t smem = csum ? (size_t)(Ca + Cb) * sizeof(float) : 0;
    if (v4) vnb::launch_k(concat_planes_kernel<4>, grid, EW_THREADS, smem, ST, xa, Ca, xb, Cb, rows, round_up8(Ca + Cb), P, (bf16*)out, csum);
    else vnb::launch_k(concat_planes_kernel<1>, grid, EW_THREADS, smem, ST, xa, Ca, xb, Cb, rows, round_up8(Ca + Cb), P, (bf16*)out, csum);
    VNB_LAUNCH_CHECK("concat_planes");
    return VNB_OK;
}

extern "C" int vnb_gn_stats_ws_floats(int B, int groups) { return B * groups * GN_MAX_CHUNKS * 2; }

extern "C" int vnb_gn_stats(const float* h, int B, int C, int T, int groups, float eps, float* ws, float* mean,
                            float* rstd, void* stream) {
    VNB_REQUIRE(h && ws && mean && rstd && B > 0 && B <= 65535 && groups > 0 && groups <= GN_MAX_GROUPS &&
                    C % groups == 0 && T > 0, "gn_stats: bad arguments");
    int rows_per = 8192 / C;
    if (rows_per < 1) rows_per = 1;
    int chunks = ceil_div(T, rows_per);
    if (chunks > GN_MAX_CHUNKS) {
        chunks = GN_MAX_CHUNKS;
        rows_per = ceil_div(T, chunks);
        chunks = ceil_div(T, rows_per);
    }
    vnb::launch_k(gn_stats_partial_kernel, dim3(chunks, B), 256, 0, ST, h, C, T, groups, rows_per, ws);
    VNB_LAUNCH_CHECK("gn_stats_partial");
    const int BG = B * groups;
    vnb::launch_k(gn_stats_merge_kernel, ceil_div((int64_t)BG * 32, 128), 128, 0, ST, ws, BG, chunks, T, rows_per, C / groups, eps,
                                                                           mean, rstd);
    VNB_LAUNCH_CHECK("gn_stats_merge");
    return VNB_OK;
}

// ---- vnb_gn_stats without a caller-provided workspace (no counterpart in the reference; added
// for callers that cannot allocate: CUDA-graph capture, serving loops).  The chunk-wise mean / M2
// workspace lives in this file-scope device array: 256 KB held by every GPU that loads the library, until exit.
// Every call shares it: calls that overlap on the GPU (different streams, no ordering between them) RACE.  Nothing
// in it has to start at zero - gn_stats_partial_kernel writes every (sample, group, chunk) entry the merge reads.
constexpr int GN_STATIC_WS_FLOATS = 65536;
static __device__ __align__(256) float g_gn_static_ws[GN_STATIC_WS_FLOATS];

// host side: the symbol's name is NOT a device address - resolve it once per device, with that device current
static float* gn_static_ws_ptr() {
    constexpr int MAX_DEV = 64;
    static float* ptrs[MAX_DEV] = {};
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEV) {
        set_error("gn_stats_static: cudaGetDevice failed or device index >= %d", MAX_DEV);
        return nullptr;
    }
    float* p = __atomic_load_n(&ptrs[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_gn_static_ws);
        if (e != cudaSuccess || !p) {
            set_error("gn_stats_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            return nullptr;
        }
        __atomic_store_n(&ptrs[dev], p, __ATOMIC_RELEASE);     // every thread resolves the same address: benign if two race here
    }
    return p;
}

extern "C" int vnb_gn_stats_static_max_floats(void) { return GN_STATIC_WS_FLOATS; }

extern "C" int vnb_gn_stats_static(const float* h, int B, int C, int T, int groups, float eps, float* mean, float* rstd,
                                   void* stream) {
    VNB_REQUIRE(B > 0 && groups > 0 && groups <= GN_MAX_GROUPS, "gn_stats_static: bad arguments");
    const int64_t need = (int64_t)B * groups * GN_MAX_CHUNKS * 2;
    if (need > GN_STATIC_WS_FLOATS) {
        set_error("gn_stats_static: B * groups = %lld needs %lld workspace floats, the static workspace holds %d "
                  "(B * groups <= %d); use vnb_gn_stats with a caller-provided workspace",
                  (long long)B * groups, (long long)need, GN_STATIC_WS_FLOATS, GN_STATIC_WS_FLOATS / (GN_MAX_CHUNKS * 2));
        return VNB_E_ARG;
    }
    float* ws = gn_static_ws_ptr();
    if (!ws) return VNB_E_CUDA;
    return vnb_gn_stats(h, B, C, T, groups, eps, ws, mean, rstd, stream);
}

extern "C" int vnb_gn_finalize(const double* gn_sums, int BG, double n, float eps, float* mean, float* rstd,
                               void* stream) {
    VNB_REQUIRE(gn_sums && mean && rstd && BG > 0 && n > 0, "gn_finalize: bad arguments");
    vnb::launch_k(gn_finalize_kernel, ceil_div(BG, 128), 128, 0, ST, gn_sums, BG, n, eps, mean, rstd);
    VNB_LAUNCH_CHECK("gn_finalize");
    return VNB_OK;
}

static int enc_out_impl(const float* h, const float* s, const float* mean, const float* rstd, const float* gamma,
                        const float* beta, int B, int C, int T, int P, float* y, void* y_planes, void* u_planes,
                        const double* sums, float eps, float* mean_out, float* rstd_out, void* stream) {
    VNB_REQUIRE(h && (sums ? (mean_out && rstd_out) : (mean && rstd)) && gamma && beta && (y || y_planes || u_planes), "enc_out: null pointer");
    VNB_REQUIRE(B > 0 && C > 0 && T > 0 && (P >= 1 && P <= 3), "enc_out: bad shape");
    const int64_t rows = (int64_t)B * T;
    if ((C & 3) == 0 && ew_v2_enabled()) {
        const int g2 = v2_grid(rows, C / 4);
        if (g2 && (!sums || B <= 4096)) {
            const int Cp = round_up8(C);
            const size_t smem = sums ? (size_t)B * 2 * sizeof(float) : 0;
            const double n = (double)C * (double)T;
            if (P == 1) vnb::launch_k(enc_out_v2_kernel<1, 2>, g2, V2_THREADS, smem, ST, h, s, mean, rstd, gamma, beta, (int)rows, C, T, Cp, y, (bf16*)y_planes, (bf16*)u_planes, sums, n, eps, mean_out, rstd_out);
            else if (P == 2) vnb::launch_k(enc_out_v2_kernel<2, 2>, g2, V2_THREADS, smem, ST, h, s, mean, rstd, gamma, beta, (int)rows, C, T, Cp, y, (bf16*)y_planes, (bf16*)u_planes, sums, n, eps, mean_out, rstd_out);
            else vnb::launch_k(enc_out_v2_kernel<3, 2>, g2, V2_THREADS, smem, ST, h, s, mean, rstd, gamma, beta, (int)rows, C, T, Cp, y, (bf16*)y_planes, (bf16*)u_planes, sums, n, eps, mean_out, rstd_out);
            VNB_LAUNCH_CHECK("enc_out_v2");
            return VNB_OK;
        }
    }
    if (sums) {                                   // shapes the fused path does not take: finalise separately
        vnb::launch_k(gn_finalize_kernel, ceil_div(B, 128), 128, 0, ST, sums, B, (double)C * (double)T, eps, mean_out, rstd_out);
        VNB_LAUNCH_CHECK("gn_finalize");
        mean = mean_out;
        rstd = rstd_out;
    }
    VEC_DISPATCH((C & 3) == 0, enc_out_kernel, rows * C, ST, h, s, mean, rstd, gamma, beta, rows, C, T, round_up8(C), P,
                 y, (bf16*)y_planes, (bf16*)u_planes);
    VNB_LAUNCH_CHECK("enc_out");
    return VNB_OK;
}

extern "C" int vnb_enc_out(const float* h, const float* s, const float* mean, const float* rstd, const float* gamma,
                           const float* beta, int B, int C, int T, int P, float* y, void* y_planes, void* u_planes,
                           void* stream) {
    return enc_out_impl(h, s, mean, rstd, gamma, beta, B, C, T, P, y, y_planes, u_planes, nullptr, 0.f, nullptr, nullptr, stream);
}

extern "C" int vnb_enc_out_s(const float* h, const float* s, const double* gn_sums, float eps, const float* gamma,
                             const float* beta, int B, int C, int T, int P, float* y, void* y_planes, void* u_planes,
                             float* mean_out, float* rstd_out, void* stream) {
    VNB_REQUIRE(gn_sums, "enc_out_s: null statistics");
    return enc_out_impl(h, s, nullptr, nullptr, gamma, beta, B, C, T, P, y, y_planes, u_planes, gn_sums, eps, mean_out, rstd_out, stream);
}

static int dec_gate_impl(const float* a, const float* mean, const float* rstd, const float* gamma, const float* beta, int B,
                         int C, int T, int P, void* g_planes, const double* sums, float eps, float* mean_out, float* rstd_out,
                         void* stream) {
    VNB_REQUIRE(a && (sums ? (mean_out && rstd_out) : (mean && rstd)) && gamma && beta && g_planes, "dec_gate: null pointer");
    VNB_REQUIRE(B > 0 && C > 0 && T > 0 && (P >= 1 && P <= 3), "dec_gate: bad shape");
    const int64_t rows = (int64_t)B * T;
    if ((C & 3) == 0 && ew_v2_enabled()) {
        const int g2 = v2_grid(rows, C / 4);
        if (g2 && (!sums || B <= 2048)) {
            const int Cp = round_up8(C);
            const size_t smem = sums ? (size_t)B * 4 * sizeof(float) : 0;
            const double n = (double)C * (double)T;
            if (P == 1 && fast_gate_enabled()) vnb::launch_k(dec_gate_v2_kernel<1, 2, true>, g2, V2_THREADS, smem, ST, a, mean, rstd, gamma, beta, (int)rows, C, T, Cp, (bf16*)g_planes, sums, n, eps, mean_out, rstd_out);
            else if (P == 1) vnb::launch_k(dec_gate_v2_kernel<1, 2, false>, g2, V2_THREADS, smem, ST, a, mean, rstd, gamma, beta, (int)rows, C, T, Cp, (bf16*)g_planes, sums, n, eps, mean_out, rstd_out);
            else if (P == 2) vnb::launch_k(dec_gate_v2_kernel<2, 2, false>, g2, V2_THREADS, smem, ST, a, mean, rstd, gamma, beta, (int)rows, C, T, Cp, (bf16*)g_planes, sums, n, eps, mean_out, rstd_out);
            else vnb::launch_k(dec_gate_v2_kernel<3, 2, false>, g2, V2_THREADS, smem, ST, a, mean, rstd, gamma, beta, (int)rows, C, T, Cp, (bf16*)g_planes, sums, n, eps, mean_out, rstd_out);
            VNB_LAUNCH_CHECK("dec_gate_v2");
            return VNB_OK;
        }
    }
    if (sums) {
        vnb::launch_k(gn_finalize_kernel, ceil_div(2 * B, 128), 128, 0, ST, sums, 2 * B, (double)C * (double)T, eps, mean_out, rstd_out);
        VNB_LAUNCH_CHECK("gn_finalize");
        mean = mean_out;
        rstd = rstd_out;
    }
    VEC_DISPATCH((C & 3) == 0, dec_gate_kernel, rows * C, ST, a, mean, rstd, gamma, beta, rows, C, T, round_up8(C), P,
                 (bf16*)g_planes);
    VNB_LAUNCH_CHECK("dec_gate");
    return VNB_OK;
}

extern "C" int vnb_dec_gate(const float* a, const float* mean, const float* rstd, const float* gamma,
                            const float* beta, int B, int C, int T, int P, void* g_planes, void* stream) {
    return dec_gate_impl(a, mean, rstd, gamma, beta, B, C, T, P, g_planes, nullptr, 0.f, nullptr, nullptr, stream);
}

extern "C" int vnb_dec_gate_s(const float* a, const double* gn_sums, float eps, const float* gamma, const float* beta, int B,
                              int C, int T, int P, void* g_planes, float* mean_out, float* rstd_out, void* stream) {
    VNB_REQUIRE(gn_sums, "dec_gate_s: null statistics");
    return dec_gate_impl(a, nullptr, nullptr, gamma, beta, B, C, T, P, g_planes, gn_sums, eps, mean_out, rstd_out, stream);
}

extern "C" int vnb_dec_out(const float* r, const float* x, int B, int C, int S, int T, int P, float* x_out,
                           void* x_planes, float* skip_sum, int first, void* stream) {
    VNB_REQUIRE(r && x && B > 0 && C > 0 && S >= 0 && T > 0 && (P >= 1 && P <= 3), "dec_out: bad arguments");
    const int64_t rows = (int64_t)B * T;
    VEC_DISPATCH(((C | S) & 3) == 0, dec_out_kernel, rows * (C + S), ST, r, x, rows, C, S, round_up8(C), P, x_out,
                 (bf16*)x_planes, skip_sum, first);
    VNB_LAUNCH_CHECK("dec_out");
    return VNB_OK;
}
