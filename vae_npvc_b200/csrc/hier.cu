// Ops that the hierarchical VQ-VAE2 + global-style-token model adds to the single-level path.
// Reference: vae_npvc/model/vqvae2.py:96 (mean over time), :104-114 (upsample + channel concat), :130-143 (upsample);
// vae_npvc/model/layers_gst.py:46-60, 128-147 (one query per utterance against N style tokens, H heads).
// Activations are channels-last (B, T, C) rows.  The strided down-sampling convs need no kernel of their own: on
// channels-last rows a stride-s, kernel-2s conv is a 3-tap stride-1 conv over rows of s frames (functional.py).
#include "common.cuh"

namespace vnb {

// out[b, t, off + c] = src[b, min(t / rep, Ts - 1), c]: nearest-neighbour upsample (every frame repeated rep = T / Ts
// times, the tail filled with the last frame, vqvae2.py:130-143) written into a channel slice of the concatenation.
__global__ void __launch_bounds__(256) upsample_cat_fwd_kernel(const float4* __restrict__ src, int C4, int Ts, int rep,
                                                               float4* __restrict__ out, int Ct4, int off4, int T,
                                                               int64_t total) {
    pdl_wait();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(i % C4);
        const int64_t bt = i / C4;
        const int t = (int)(bt % T);
        const int64_t b = bt / T;
        int ts = t / rep;
        ts = ts < Ts ? ts : Ts - 1;
        out[(b * T + t) * Ct4 + off4 + c] = __ldg(src + (b * Ts + ts) * C4 + c);
    }
}

// dsrc[b, ts, c] = sum over the frames t that read (b, ts, c) of dout[b, t, off + c]
__global__ void __launch_bounds__(256) upsample_cat_bwd_kernel(const float4* __restrict__ dout, int Ct4, int off4, int T,
                                                               int rep, float4* __restrict__ dsrc, int C4, int Ts,
                                                               int64_t total) {
    pdl_wait();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(i % C4);
        const int64_t bt = i / C4;
        const int ts = (int)(bt % Ts);
        const int64_t b = bt / Ts;
        const int t0 = ts * rep;
        const int t1 = ts == Ts - 1 ? T : (t0 + rep < T ? t0 + rep : T);
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int t = t0; t < t1; t++) {
            const float4 v = __ldg(dout + (b * T + t) * Ct4 + off4 + c);
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
        dsrc[i] = s;
    }
}

// out[b, c] = mean_t x[b, t, c]   (block = 32 channels x 8 frame lanes; grid (ceil(C/32), B))
__global__ void __launch_bounds__(256) time_mean_fwd_kernel(const float* __restrict__ x, int C, int T,
                                                            float* __restrict__ out) {
    pdl_wait();
    __shared__ float part[8][33];
    const int cx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + cx, b = blockIdx.y;
    float s = 0.f;
    if (c < C)
        for (int t = ty; t < T; t += 8) s += x[((int64_t)b * T + t) * C + c];
    part[ty][cx] = s;
    __syncthreads();
    if (ty == 0 && c < C) {
        float a = 0.f;
#pragma unroll
        for (int j = 0; j < 8; j++) a += part[j][cx];
        out[(int64_t)b * C + c] = a / (float)T;
    }
}

// dx[b, t, c] = dy[b, c] / T
__global__ void __launch_bounds__(256) time_mean_bwd_kernel(const float* __restrict__ dy, int C, int T,
                                                            float* __restrict__ dx, int64_t total) {
    pdl_wait();
    const float inv = 1.f / (float)T;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        const int64_t b = i / ((int64_t)C * T);
        dx[i] = dy[b * C + c] * inv;
    }
}

// Style-token attention core, one warp per (utterance, head): scores = q_h . K_{n,h} / sqrt(dk), softmax over the N
// tokens, ctx_h = sum_n attn_n V_{n,h}  (layers_gst.py:143-147, :128-134 without mask / dropout).  N <= 32, dk <= 128.
__global__ void __launch_bounds__(128) gst_attn_fwd_kernel(const float* __restrict__ q, const float* __restrict__ K,
                                                           const float* __restrict__ V, int B, int N, int H, int dk,
                                                           float* __restrict__ attn, float* __restrict__ ctx) {
    pdl_wait();
    const int w = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (w >= B * H) return;
    const int b = w / H, h = w - b * H, F = H * dk;
    const float* qh = q + (int64_t)b * F + h * dk;
    const float scale = rsqrtf((float)dk);
    float my = -INFINITY;                       // lane n holds the score of token n
    for (int n = 0; n < N; n++) {
        float s = 0.f;
        for (int d = lane; d < dk; d += 32) s = fmaf(qh[d], K[(int64_t)n * F + h * dk + d], s);
        s = warp_sum(s) * scale;
        if (lane == n) my = s;
    }
    float m = my;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    const float e = lane < N ? expf(my - m) : 0.f;
    const float a = e / warp_sum(e);
    if (lane < N) attn[((int64_t)b * H + h) * N + lane] = a;
    for (int d0 = 0; d0 < dk; d0 += 32) {       // uniform trip count: the shuffles below need the whole warp
        const int d = d0 + lane;
        float c = 0.f;
        for (int n = 0; n < N; n++) {
            const float an = __shfl_sync(0xffffffffu, a, n);
            if (d < dk) c = fmaf(an, V[(int64_t)n * F + h * dk + d], c);
        }
        if (d < dk) ctx[(int64_t)b * F + h * dk + d] = c;
    }
}

// backward, part 1 (one warp per (utterance, head)): da_n = dctx_h . V_{n,h}; ds = attn * (da - sum attn da);
// dq_h = sum_n ds_n K_{n,h} / sqrt(dk); ds stored for part 2
__global__ void __launch_bounds__(128) gst_attn_bwd_q_kernel(const float* __restrict__ dctx, const float* __restrict__ K,
                                                             const float* __restrict__ V, const float* __restrict__ attn,
                                                             int B, int N, int H, int dk, float* __restrict__ ds,
                                                             float* __restrict__ dq) {
    pdl_wait();
    const int w = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (w >= B * H) return;
    const int b = w / H, h = w - b * H, F = H * dk;
    const float* dch = dctx + (int64_t)b * F + h * dk;
    const float scale = rsqrtf((float)dk);
    float da = 0.f;
    for (int n = 0; n < N; n++) {
        float s = 0.f;
        for (int d = lane; d < dk; d += 32) s = fmaf(dch[d], V[(int64_t)n * F + h * dk + d], s);
        s = warp_sum(s);
        if (lane == n) da = s;
    }
    const float a = lane < N ? attn[((int64_t)b * H + h) * N + lane] : 0.f;
    const float dot = warp_sum(a * da);
    const float dsn = a * (da - dot);
    if (lane < N) ds[((int64_t)b * H + h) * N + lane] = dsn;
    for (int d0 = 0; d0 < dk; d0 += 32) {
        const int d = d0 + lane;
        float g = 0.f;
        for (int n = 0; n < N; n++) {
            const float dn = __shfl_sync(0xffffffffu, dsn, n);
            if (d < dk) g = fmaf(dn, K[(int64_t)n * F + h * dk + d], g);
        }
        if (d < dk) dq[(int64_t)b * F + h * dk + d] = g * scale;
    }
}

// backward, part 2 (thread per (token, feature); deterministic sums over the batch):
// dV[n, f] = sum_b attn[b,h,n] dctx[b,f];  dK[n, f] = sum_b ds[b,h,n] q[b,f] / sqrt(dk)
__global__ void __launch_bounds__(128) gst_attn_bwd_kv_kernel(const float* __restrict__ dctx, const float* __restrict__ q,
                                                              const float* __restrict__ attn, const float* __restrict__ ds,
                                                              int B, int N, int H, int dk, float* __restrict__ dK,
                                                              float* __restrict__ dV) {
    pdl_wait();
    const int F = H * dk;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N * F) return;
    const int n = i / F, f = i - n * F, h = f / dk;
    const float scale = rsqrtf((float)dk);
    float gk = 0.f, gv = 0.f;
    for (int b = 0; b < B; b++) {
        const int64_t an = ((int64_t)b * H + h) * N + n;
        gv = fmaf(attn[an], dctx[(int64_t)b * F + f], gv);
        gk = fmaf(ds[an], q[(int64_t)b * F + f], gk);
    }
    dV[i] = gv;
    dK[i] = gk * scale;
}

static inline int grid_for(int64_t n) {
    int64_t g = (n + 255) / 256;
    return (int)(g < 1 ? 1 : (g > 148 * 16 ? 148 * 16 : g));
}

}  // namespace vnb

using namespace vnb;
#define ST ((cudaStream_t)stream)

extern "C" int vnb_upsample_cat_fwd(const float* src, int B, int C, int Ts, float* out, int Ct, int off, int T,
                                    void* stream) {
    VNB_REQUIRE(src && out && B > 0 && C > 0 && Ts > 0 && T >= Ts && off >= 0 && off + C <= Ct,
                "upsample_cat_fwd: bad arguments (upsampling to a shorter length is undefined in the reference)");
    VNB_REQUIRE((C & 3) == 0 && (Ct & 3) == 0 && (off & 3) == 0 && (((uintptr_t)src | (uintptr_t)out) & 15) == 0,
                "upsample_cat_fwd: channel counts / offset must be multiples of 4 and pointers 16-byte aligned");
    const int64_t total = (int64_t)B * T * (C / 4);
    launch_k(upsample_cat_fwd_kernel, grid_for(total), 256, 0, ST, (const float4*)src, C / 4, Ts, T / Ts, (float4*)out, Ct / 4,
             off / 4, T, total);
    VNB_LAUNCH_CHECK("upsample_cat_fwd");
    return VNB_OK;
}

extern "C" int vnb_upsample_cat_bwd(const float* dout, int B, int Ct, int off, int T, float* dsrc, int C, int Ts,
                                    void* stream) {
    VNB_REQUIRE(dout && dsrc && B > 0 && C > 0 && Ts > 0 && T >= Ts && off >= 0 && off + C <= Ct, "upsample_cat_bwd: bad arguments");
    VNB_REQUIRE((C & 3) == 0 && (Ct & 3) == 0 && (off & 3) == 0 && (((uintptr_t)dsrc | (uintptr_t)dout) & 15) == 0,
                "upsample_cat_bwd: channel counts / offset must be multiples of 4 and pointers 16-byte aligned");
    const int64_t total = (int64_t)B * Ts * (C / 4);
    launch_k(upsample_cat_bwd_kernel, grid_for(total), 256, 0, ST, (const float4*)dout, Ct / 4, off / 4, T, T / Ts, (float4*)dsrc,
             C / 4, Ts, total);
    VNB_LAUNCH_CHECK("upsample_cat_bwd");
    return VNB_OK;
}

extern "C" int vnb_time_mean_fwd(const float* x, int B, int C, int T, float* out, void* stream) {
    VNB_REQUIRE(x && out && B > 0 && B <= 65535 && C > 0 && T > 0, "time_mean_fwd: bad arguments");
    launch_k(time_mean_fwd_kernel, dim3(ceil_div(C, 32), B), 256, 0, ST, x, C, T, out);
    VNB_LAUNCH_CHECK("time_mean_fwd");
    return VNB_OK;
}

extern "C" int vnb_time_mean_bwd(const float* dy, int B, int C, int T, float* dx, void* stream) {
    VNB_REQUIRE(dy && dx && B > 0 && C > 0 && T > 0, "time_mean_bwd: bad arguments");
    const int64_t total = (int64_t)B * T * C;
    launch_k(time_mean_bwd_kernel, grid_for(total), 256, 0, ST, dy, C, T, dx, total);
    VNB_LAUNCH_CHECK("time_mean_bwd");
    return VNB_OK;
}

extern "C" int vnb_gst_attn_fwd(const float* q, const float* K, const float* V, int B, int N, int H, int dk, float* attn,
                                float* ctx, void* stream) {
    VNB_REQUIRE(q && K && V && attn && ctx && B > 0 && N > 0 && N <= 32 && H > 0 && dk > 0,
                "gst_attn_fwd: bad arguments (at most 32 style tokens)");
    launch_k(gst_attn_fwd_kernel, ceil_div((int64_t)B * H, 4), 128, 0, ST, q, K, V, B, N, H, dk, attn, ctx);
    VNB_LAUNCH_CHECK("gst_attn_fwd");
    return VNB_OK;
}

extern "C" int vnb_gst_attn_bwd(const float* dctx, const float* q, const float* K, const float* V, const float* attn, int B,
                                int N, int H, int dk, float* ds, float* dq, float* dK, float* dV, void* stream) {
    VNB_REQUIRE(dctx && q && K && V && attn && ds && dq && dK && dV && B > 0 && N > 0 && N <= 32 && H > 0 && dk > 0,
                "gst_attn_bwd: bad arguments (at most 32 style tokens)");
    launch_k(gst_attn_bwd_q_kernel, ceil_div((int64_t)B * H, 4), 128, 0, ST, dctx, K, V, attn, B, N, H, dk, ds, dq);
    VNB_LAUNCH_CHECK("gst_attn_bwd_q");
    launch_k(gst_attn_bwd_kv_kernel, ceil_div((int64_t)N * H * dk, 128), 128, 0, ST, dctx, q, attn, ds, B, N, H, dk, dK, dV);
    VNB_LAUNCH_CHECK("gst_attn_bwd_kv");
    return VNB_OK;
}
