// CUDA-core (fp32 FFMA) GEMM backend: conv forward / data-gradient, weight-gradient and the VQ
// distance+argmin.  It consumes exactly the same bf16 operand planes as the tcgen05 backend
// (summing the P planes back to fp32 on load), so it is a drop-in cross-check for it and the
// backend for shapes the tensor-core kernel does not take.  Not a CPU fallback: CUDA only.
#include "common.cuh"

namespace vnb {

constexpr int TM = 64, TN = 64, KC = 16;

// out[b,t,co] = (accumulate ? out : 0) + bias + residual + sum_{tap,ci} W[tap][co][ci] * in[b,t+off(tap),ci]
// channels-last: rows = frames (one sample per blockIdx.z so that the zero padding is per sample)
template <int P>
__global__ void __launch_bounds__(256) conv_fwd_simt_kernel(
    const bf16* __restrict__ in, int B, int Cin, int T, const bf16* __restrict__ w, int taps, int dil, int Cout,
    int Cin_p, const float* __restrict__ bias, int bias_bstride, const float* __restrict__ residual,
    float* __restrict__ out, int accumulate, double* __restrict__ gn_sums, int gn_groups) {
    pdl_wait();
    __shared__ float Xs[KC][TM + 4];
    __shared__ float Ws[KC][TN + 4];
    const int t0 = blockIdx.x * TM, co0 = blockIdx.y * TN, b = blockIdx.z;
    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    const int64_t in_plane = (int64_t)B * T * Cin_p;
    const int64_t w_plane = (int64_t)taps * Cout * Cin_p;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = 0.f;

    for (int tap = 0; tap < taps; tap++) {
        const int off = (tap - (taps - 1) / 2) * dil;
        for (int c0 = 0; c0 < Cin; c0 += KC) {
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const int idx = tid * 4 + e;
                const int m = idx >> 4, kk = idx & 15;       // 64 rows x 16 ci, ci contiguous
                const int ci = c0 + kk;
                {
                    const int t = t0 + m + off;
                    float x = 0.f;
                    if (ci < Cin && t >= 0 && t < T) x = load_planes(in, ((int64_t)b * T + t) * Cin_p + ci, in_plane, P);
                    Xs[kk][m] = x;
                }
                {
                    const int co = co0 + m;
                    float x = 0.f;
                    if (co < Cout && ci < Cin) x = load_planes(w, ((int64_t)tap * Cout + co) * Cin_p + ci, w_plane, P);
                    Ws[kk][m] = x;
                }
            }
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < KC; kk++) {
                float a[4], x[4];
#pragma unroll
                for (int i = 0; i < 4; i++) a[i] = Xs[kk][ty * 4 + i];
#pragma unroll
                for (int j = 0; j < 4; j++) x[j] = Ws[kk][tx * 4 + j];
#pragma unroll
                for (int i = 0; i < 4; i++)
#pragma unroll
                    for (int j = 0; j < 4; j++) acc[i][j] = fmaf(a[i], x[j], acc[i][j]);
            }
            __syncthreads();
        }
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int t = t0 + ty * 4 + i;
        if (t >= T) continue;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int co = co0 + tx * 4 + j;
            if (co >= Cout) continue;
            const int64_t o = ((int64_t)b * T + t) * Cout + co;
            float r = acc[i][j] + (bias ? bias[(int64_t)b * bias_bstride + co] : 0.f);
            if (residual) r += residual[o];
            if (accumulate) r += out[o];
            out[o] = r;
            if (gn_sums) {       // debug backend: plain per-element fp64 atomics
                double* gs = gn_sums + ((int64_t)b * gn_groups + co / (Cout / gn_groups)) * 2;
                atomicAdd(gs, (double)r);
                atomicAdd(gs + 1, (double)r * (double)r);
            }
        }
    }
}

// dW[co][ci][tap] = sum_{b,t} dy[b,t,co] * u[b,t+off(tap),ci]
template <int P>
__global__ void __launch_bounds__(256) conv_wgrad_simt_kernel(
    const bf16* __restrict__ dy, const bf16* __restrict__ u, int B, int Cout, int Cin, int T, int taps, int dil,
    int transposed, float* __restrict__ dw) {
    pdl_wait();
    __shared__ float As[KC][TM + 4];
    __shared__ float Bs[KC][TN + 4];
    const int ci0 = blockIdx.x * TN, co0 = blockIdx.y * TM, tap = blockIdx.z;
    const int off = (tap - (taps - 1) / 2) * dil;
    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    const int Cout_p = (Cout + 7) & ~7, Cin_p = (Cin + 7) & ~7;
    const int64_t dy_plane = (int64_t)B * T * Cout_p, u_plane = (int64_t)B * T * Cin_p;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = 0.f;
    for (int b = 0; b < B; b++) {
        for (int tt = 0; tt < T; tt += KC) {
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const int idx = tid * 4 + e;
                const int kk = idx >> 6, m = idx & 63;       // 16 frames x 64 channels, channels contiguous
                const int t = tt + kk;
                {
                    const int co = co0 + m;
                    float x = 0.f;
                    if (co < Cout && t < T) x = load_planes(dy, ((int64_t)b * T + t) * Cout_p + co, dy_plane, P);
                    As[kk][m] = x;
                }
                {
                    const int ci = ci0 + m, ts = t + off;
                    float x = 0.f;
                    if (ci < Cin && t < T && ts >= 0 && ts < T)
                        x = load_planes(u, ((int64_t)b * T + ts) * Cin_p + ci, u_plane, P);
                    Bs[kk][m] = x;
                }
            }
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < KC; kk++) {
                float a[4], x[4];
#pragma unroll
                for (int i = 0; i < 4; i++) a[i] = As[kk][ty * 4 + i];
#pragma unroll
                for (int j = 0; j < 4; j++) x[j] = Bs[kk][tx * 4 + j];
#pragma unroll
                for (int i = 0; i < 4; i++)
#pragma unroll
                    for (int j = 0; j < 4; j++) acc[i][j] = fmaf(a[i], x[j], acc[i][j]);
            }
            __syncthreads();
        }
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int co = co0 + ty * 4 + i;
        if (co >= Cout) continue;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int ci = ci0 + tx * 4 + j;
            if (ci >= Cin) continue;
            const int64_t o = transposed ? (((int64_t)ci * Cout + co) * taps + (taps - 1 - tap))
                                         : (((int64_t)co * Cin + ci) * taps + tap);
            dw[o] = acc[i][j];
        }
    }
}

// VQ lookup: 16 frames per block, every thread scans codes tid, tid+256, ...
constexpr int VQ_FR = 16;
__global__ void __launch_bounds__(256) vq_lookup_simt_kernel(const float* __restrict__ zn, const float* __restrict__ z_sq,
                                                             const float* __restrict__ e_hat,
                                                             const float* __restrict__ e_sq, int64_t N, int D,
                                                             int K, int64_t* __restrict__ idx) {
    pdl_wait();
    extern __shared__ float sm[];
    float* zs = sm;                                   // [VQ_FR][D]
    float* bd = sm + VQ_FR * D;                       // [VQ_FR][256] best distance
    int* bi = (int*)(bd + VQ_FR * 256);               // [VQ_FR][256] best index
    const int64_t n0 = (int64_t)blockIdx.x * VQ_FR;
    const int tid = threadIdx.x;
    for (int e = tid; e < VQ_FR * D; e += 256) {
        const int64_t n = n0 + e / D;
        zs[e] = n < N ? zn[n0 * D + e] : 0.f;          // frames are contiguous rows of D floats
    }
    __syncthreads();
    float best[VQ_FR];
    int bestk[VQ_FR];
    float zq[VQ_FR];
#pragma unroll
    for (int f = 0; f < VQ_FR; f++) {
        best[f] = INFINITY;
        bestk[f] = 0x7fffffff;
        zq[f] = (n0 + f < N) ? z_sq[n0 + f] : 0.f;
    }
    for (int k = tid; k < K; k += 256) {
        float dot[VQ_FR];
#pragma unroll
        for (int f = 0; f < VQ_FR; f++) dot[f] = 0.f;
        const float* er = e_hat + (int64_t)k * D;
        for (int d = 0; d < D; d++) {
            const float ev = er[d];
#pragma unroll
            for (int f = 0; f < VQ_FR; f++) dot[f] = fmaf(zs[f * D + d], ev, dot[f]);
        }
        const float es = e_sq[k];
#pragma unroll
        for (int f = 0; f < VQ_FR; f++) {
            const float dist = fmaf(-2.f, dot[f], zq[f] + es);
            if (dist < best[f]) {
                best[f] = dist;
                bestk[f] = k;
            }
        }
    }
#pragma unroll
    for (int f = 0; f < VQ_FR; f++) {
        bd[f * 256 + tid] = best[f];
        bi[f * 256 + tid] = bestk[f];
    }
    __syncthreads();
    // one warp per 2 frames: reduce 256 candidates, lowest index wins ties
    const int w = tid >> 5, lane = tid & 31;
    for (int f = w; f < VQ_FR; f += 8) {
        float d = INFINITY;
        int k = 0x7fffffff;
        for (int c = lane; c < 256; c += 32) {
            const float dc = bd[f * 256 + c];
            const int kc = bi[f * 256 + c];
            if (dc < d || (dc == d && kc < k)) {
                d = dc;
                k = kc;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float d2 = __shfl_xor_sync(0xffffffffu, d, o);
            const int k2 = __shfl_xor_sync(0xffffffffu, k, o);
            if (d2 < d || (d2 == d && k2 < k)) {
                d = d2;
                k = k2;
            }
        }
        if (lane == 0 && n0 + f < N) idx[n0 + f] = (int64_t)k;
    }
}

int conv_fwd_simt(const void* in, int P, int B, int Cin, int T, const void* w, int taps, int dil, int Cout,
                  const float* bias, int bias_bstride, const float* residual, float* out, int accumulate,
                  double* gn_sums, int gn_groups, cudaStream_t st) {
    dim3 grid(ceil_div(T, TM), ceil_div(Cout, TN), B);
    const int Cin_p = round_up8(Cin);
    if (P == 1)
        vnb::launch_k(conv_fwd_simt_kernel<1>, grid, 256, 0, st, (const bf16*)in, B, Cin, T, (const bf16*)w, taps, dil, Cout, Cin_p,
                                                      bias, bias_bstride, residual, out, accumulate, gn_sums, gn_groups);
    else if (P == 2)
        vnb::launch_k(conv_fwd_simt_kernel<2>, grid, 256, 0, st, (const bf16*)in, B, Cin, T, (const bf16*)w, taps, dil, Cout, Cin_p,
                                                      bias, bias_bstride, residual, out, accumulate, gn_sums, gn_groups);
    else
        vnb::launch_k(conv_fwd_simt_kernel<3>, grid, 256, 0, st, (const bf16*)in, B, Cin, T, (const bf16*)w, taps, dil, Cout, Cin_p,
                                                      bias, bias_bstride, residual, out, accumulate, gn_sums, gn_groups);
    VNB_LAUNCH_CHECK("conv_fwd_simt");
    return VNB_OK;
}

int conv_wgrad_simt(const void* dy, const void* u, int P, int B, int Cout, int Cin, int T, int taps, int dil,
                    int transposed, float* dw, cudaStream_t st) {
    dim3 grid(ceil_div(Cin, TN), ceil_div(Cout, TM), taps);
    if (P == 1)
        vnb::launch_k(conv_wgrad_simt_kernel<1>, grid, 256, 0, st, (const bf16*)dy, (const bf16*)u, B, Cout, Cin, T, taps, dil,
                                                        transposed, dw);
    else if (P == 2)
        vnb::launch_k(conv_wgrad_simt_kernel<2>, grid, 256, 0, st, (const bf16*)dy, (const bf16*)u, B, Cout, Cin, T, taps, dil,
                                                        transposed, dw);
    else
        vnb::launch_k(conv_wgrad_simt_kernel<3>, grid, 256, 0, st, (const bf16*)dy, (const bf16*)u, B, Cout, Cin, T, taps, dil,
                                                        transposed, dw);
    VNB_LAUNCH_CHECK("conv_wgrad_simt");
    return VNB_OK;
}

int vq_lookup_simt(const float* zn, const float* z_sq, const float* e_hat, const float* e_sq, int B, int D, int T,
                   int K, int64_t* idx, cudaStream_t st) {
    const int64_t N = (int64_t)B * T;
    const size_t smem = (size_t)VQ_FR * D * 4 + (size_t)VQ_FR * 256 * 8;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(vq_lookup_simt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) {
            set_error("vq_lookup_simt: smem %zu: %s", smem, cudaGetErrorString(e));
            return VNB_E_CUDA;
        }
    }
    vnb::launch_k(vq_lookup_simt_kernel, ceil_div(N, VQ_FR), 256, smem, st, zn, z_sq, e_hat, e_sq, N, D, K, idx);
    VNB_LAUNCH_CHECK("vq_lookup_simt");
    return VNB_OK;
}

}  // namespace vnb
