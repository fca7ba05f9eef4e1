// Speaker-embedding gather/scatter, the time-invariant conditioning linear, log-loss and the
// flat-buffer optimiser (global-norm clip + Adam + StepLR).
// Reference: vae_npvc/model/layers.py:42-54 (Conditions), :231 (conv_cond), :283-296 (log_loss);
// vae_npvc/trainer/basic.py:38-46, 62-72 (clip_grad_norm_, Adam(betas=(0.5,0.999)), StepLR).
#include "common.cuh"

namespace vnb {

__global__ void embed_fwd_kernel(const float* __restrict__ table, const int64_t* __restrict__ idx, int n, int dim,
                                 float* __restrict__ out) {
    pdl_wait();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * dim) return;
    const int r = i / dim, d = i - r * dim;
    out[i] = table[idx[r] * dim + d];
}

// deterministic scatter-add: thread per column, sequential over the (few) rows
__global__ void embed_bwd_kernel(const float* __restrict__ dout, const int64_t* __restrict__ idx, int n, int dim,
                                 float* __restrict__ dtable) {
    pdl_wait();
    const int d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= dim) return;
    for (int r = 0; r < n; r++) dtable[idx[r] * dim + d] += dout[r * dim + d];
}

// one warp per output (b,o)
__global__ void __launch_bounds__(256) linear_fwd_kernel(const float* __restrict__ y, const float* __restrict__ W,
                                                         const float* __restrict__ bias,
                                                         const float* __restrict__ bias2, int B, int I, int O,
                                                         float* __restrict__ p) {
    pdl_wait();
    const int64_t w = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (w >= (int64_t)B * O) return;
    const int lane = threadIdx.x & 31;
    const int b = (int)(w / O), o = (int)(w - (int64_t)b * O);
    float s = 0.f;
    for (int i = lane; i < I; i += 32) s = fmaf(W[(int64_t)o * I + i], y[(int64_t)b * I + i], s);
    s = warp_sum(s);
    if (lane == 0) p[w] = s + (bias ? bias[o] : 0.f) + (bias2 ? bias2[o] : 0.f);
}

__global__ void linear_bwd_w_kernel(const float* __restrict__ dp, const float* __restrict__ y, int B, int I, int O,
                                    float* __restrict__ dW) {
    pdl_wait();
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (int64_t)O * I) return;
    const int o = (int)(e / I), i = (int)(e - (int64_t)o * I);
    float s = 0.f;
    for (int b = 0; b < B; b++) s = fmaf(dp[(int64_t)b * O + o], y[(int64_t)b * I + i], s);
    dW[e] = s;
}

// dy[b,i] += sum_o W[o,i]*dp[b,o]  (dy zeroed by the host wrapper unless accumulating).
// grid (ceil(I/32), ceil(O/64), ceil(B/16)), block 32 x 8: a block owns 64 rows of W and 16 samples; lanes run
// over i (coalesced W rows), warp ty takes samples ty and ty+8 of the chunk, so every W element is read once per
// 16 samples instead of once per sample, and the 64 loads of a thread are independent (the per-sample version
// walked O/8 = 128 dependent L2 round trips: 24 us for a 0.5 MB matrix).
__global__ void __launch_bounds__(256) linear_bwd_y_kernel(const float* __restrict__ dp, const float* __restrict__ W,
                                                           int B, int I, int O, float* __restrict__ dy) {
    pdl_wait();
    __shared__ float dps[16][64];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int i = blockIdx.x * 32 + tx, o0 = blockIdx.y * 64, b0 = blockIdx.z * 16;
    for (int e = threadIdx.x; e < 16 * 64; e += 256) {
        const int bb = e >> 6, oo = e & 63;
        dps[bb][oo] = (b0 + bb < B && o0 + oo < O) ? dp[(int64_t)(b0 + bb) * O + o0 + oo] : 0.f;
    }
    __syncthreads();
    if (i >= I) return;
    float s0 = 0.f, s1 = 0.f;
    const int on = O - o0 < 64 ? O - o0 : 64;
    if (on == 64) {
        // 4 batches of 16 independent loads: W is usually cold (DRAM latency), so keep many in flight
#pragma unroll 1
        for (int ob = 0; ob < 64; ob += 16) {
            float w[16];
#pragma unroll
            for (int u = 0; u < 16; u++) w[u] = W[(int64_t)(o0 + ob + u) * I + i];
#pragma unroll
            for (int u = 0; u < 16; u++) {
                s0 = fmaf(w[u], dps[ty][ob + u], s0);
                s1 = fmaf(w[u], dps[ty + 8][ob + u], s1);
            }
        }
    } else {
        for (int oo = 0; oo < on; oo++) {
            const float w = W[(int64_t)(o0 + oo) * I + i];
            s0 = fmaf(w, dps[ty][oo], s0);
            s1 = fmaf(w, dps[ty + 8][oo], s1);
        }
    }
    if (b0 + ty < B) atomicAdd(dy + (int64_t)(b0 + ty) * I + i, s0);
    if (b0 + ty + 8 < B) atomicAdd(dy + (int64_t)(b0 + ty + 8) * I + i, s1);
}

// ---- the conditioning path of ALL gated decoder blocks in one launch each way ---------------------------------------
// Every block applies its own 1x1 conv to the SAME time-invariant speaker embedding (vqvae.py:309-312, layers.py:231):
// P[l][b][o] = sum_i W_l[o][i] * y[b][i] + bias_l[o] + bias2_l[o].  ptrs: L rows of 4 int64 device addresses
// {W_l (O x I), bias_l | 0, bias2_l | 0, dW_l | 0}.  One warp per (l, o): the weight row stays in registers for all B samples.
__global__ void __launch_bounds__(256) linear_fwd_batch_kernel(const float* __restrict__ y, const int64_t* __restrict__ ptrs,
                                                               int L, int B, int I, int O, float* __restrict__ p) {
    pdl_wait();
    const int64_t w = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (w >= (int64_t)L * O) return;
    const int lane = threadIdx.x & 31;
    const int l = (int)(w / O), o = (int)(w - (int64_t)l * O);
    const float* W = reinterpret_cast<const float*>(ptrs[4 * l]) + (int64_t)o * I;
    const float* b1 = reinterpret_cast<const float*>(ptrs[4 * l + 1]);
    const float* b2 = reinterpret_cast<const float*>(ptrs[4 * l + 2]);
    const float bias = (b1 ? b1[o] : 0.f) + (b2 ? b2[o] : 0.f);
    // eight samples per pass: independent accumulators and reductions (the one-sample-at-a-time loop was a chain of B
    // dependent warp reductions: 26 us for 21 MFLOP at B = 16)
    for (int b0 = 0; b0 < B; b0 += 8) {
        float s[8];
#pragma unroll
        for (int u = 0; u < 8; u++) s[u] = 0.f;
        for (int i = lane; i < I; i += 32) {
            const float wv = W[i];
#pragma unroll
            for (int u = 0; u < 8; u++)
                if (b0 + u < B) s[u] = fmaf(wv, y[(int64_t)(b0 + u) * I + i], s[u]);
        }
#pragma unroll
        for (int u = 0; u < 8; u++) s[u] = warp_sum(s[u]);
        if (lane == 0) {
#pragma unroll
            for (int u = 0; u < 8; u++)
                if (b0 + u < B) p[((int64_t)l * B + b0 + u) * O + o] = s[u] + bias;
        }
    }
}

// Second generation (I % 4 == 0): a block owns one layer and 16 output rows; the weight tile (16 x I) and the embeddings of
// 16 utterances sit in shared memory (rows padded by 4 floats: conflict-free float4 reads), thread (oo, bb) computes one
// output with I FMAs.  The warp-per-row kernel above executed 35x the minimal instruction count (4 FMAs per lane and
// sample, then a 5-step shuffle reduction per output): 24.5 us in ncu, issue-bound, for 21 MFLOP.
__global__ void __launch_bounds__(256) linear_fwd_batch_v2_kernel(const float* __restrict__ y, const int64_t* __restrict__ ptrs,
                                                                  int L, int B, int I, int O, float* __restrict__ p) {
    pdl_wait();
    extern __shared__ __align__(16) float lin_s[];
    const int pitch = I + 4, ot_n = (O + 15) / 16;
    float* ws = lin_s;                       // [16][pitch]
    float* ys = lin_s + 16 * pitch;          // [16][pitch]
    const int l = blockIdx.x / ot_n, o0 = (blockIdx.x - l * ot_n) * 16;
    const float* W = reinterpret_cast<const float*>(ptrs[4 * l]);
    const float* b1 = reinterpret_cast<const float*>(ptrs[4 * l + 1]);
    const float* b2 = reinterpret_cast<const float*>(ptrs[4 * l + 2]);
    const int I4 = I >> 2;
    for (int e = threadIdx.x; e < 16 * I4; e += 256) {
        const int r = e / I4, c4 = e - r * I4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (o0 + r < O) v = __ldg(reinterpret_cast<const float4*>(W + (int64_t)(o0 + r) * I) + c4);
        *reinterpret_cast<float4*>(ws + r * pitch + 4 * c4) = v;
    }
    const int oo = threadIdx.x & 15, bb = threadIdx.x >> 4;
    const int o = o0 + oo;
    const float bias = o < O ? (b1 ? b1[o] : 0.f) + (b2 ? b2[o] : 0.f) : 0.f;
    for (int b0 = 0; b0 < B; b0 += 16) {
        __syncthreads();
        for (int e = threadIdx.x; e < 16 * I4; e += 256) {
            const int r = e / I4, c4 = e - r * I4;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (b0 + r < B) v = __ldg(reinterpret_cast<const float4*>(y + (int64_t)(b0 + r) * I) + c4);
            *reinterpret_cast<float4*>(ys + r * pitch + 4 * c4) = v;
        }
        __syncthreads();
        const float4* wr = reinterpret_cast<const float4*>(ws + oo * pitch);
        const float4* yr = reinterpret_cast<const float4*>(ys + bb * pitch);
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll 8
        for (int c4 = 0; c4 < I4; c4++) {
            const float4 w = wr[c4], v = yr[c4];
            s0 = fmaf(w.x, v.x, s0); s1 = fmaf(w.y, v.y, s1); s2 = fmaf(w.z, v.z, s2); s3 = fmaf(w.w, v.w, s3);
        }
        if (o < O && b0 + bb < B) p[((int64_t)l * B + b0 + bb) * O + o] = ((s0 + s1) + (s2 + s3)) + bias;
    }
}

// dW_l[o][i] = sum_b dp[l][b][o] * y[b][i], second generation (I % 4 == 0): a block owns one layer and 16 rows o; y and the
// dp tile of 16 utterances sit in shared memory, thread (i4, og) accumulates rows og and og + 8 as float4 over the samples
// (22 instructions per output against ~100 of the thread-per-element loop with two global loads per FMA).
__global__ void __launch_bounds__(256) linear_bwd_w_batch_v2_kernel(const float* __restrict__ dp, const float* __restrict__ y,
                                                                    const int64_t* __restrict__ ptrs, int L, int B, int I, int O) {
    pdl_wait();
    extern __shared__ __align__(16) float lin_s[];
    const int ot_n = (O + 15) / 16;
    float* ys = lin_s;                       // [16][I]
    float* dps = lin_s + 16 * I;             // [16][16]
    const int l = blockIdx.x / ot_n, o0 = (blockIdx.x - l * ot_n) * 16;
    float* dW = reinterpret_cast<float*>(ptrs[4 * l + 3]);
    if (!dW) return;
    const int I4 = I >> 2, lane = threadIdx.x & 31, og = threadIdx.x >> 5;
    for (int c0 = 0; c0 < I4; c0 += 32) {              // 128 columns per pass
        const int c4 = c0 + lane;
        float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
        for (int b0 = 0; b0 < B; b0 += 16) {
            __syncthreads();
            for (int e = threadIdx.x; e < 16 * I4; e += 256) {
                const int r = e / I4, cc = e - r * I4;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (b0 + r < B) v = __ldg(reinterpret_cast<const float4*>(y + (int64_t)(b0 + r) * I) + cc);
                *reinterpret_cast<float4*>(ys + r * I + 4 * cc) = v;
            }
            {
                const int r = threadIdx.x >> 4, oo = threadIdx.x & 15;
                dps[r * 16 + oo] = (b0 + r < B && o0 + oo < O) ? dp[((int64_t)l * B + b0 + r) * O + o0 + oo] : 0.f;
            }
            __syncthreads();
            if (c4 < I4) {
#pragma unroll
                for (int b = 0; b < 16; b++) {
                    const float4 v = *reinterpret_cast<const float4*>(ys + b * I + 4 * c4);
                    const float d0 = dps[b * 16 + og], d1 = dps[b * 16 + og + 8];
                    a0.x = fmaf(d0, v.x, a0.x); a0.y = fmaf(d0, v.y, a0.y); a0.z = fmaf(d0, v.z, a0.z); a0.w = fmaf(d0, v.w, a0.w);
                    a1.x = fmaf(d1, v.x, a1.x); a1.y = fmaf(d1, v.y, a1.y); a1.z = fmaf(d1, v.z, a1.z); a1.w = fmaf(d1, v.w, a1.w);
                }
            }
        }
        if (c4 < I4) {
            if (o0 + og < O) *(reinterpret_cast<float4*>(dW + (int64_t)(o0 + og) * I) + c4) = a0;
            if (o0 + og + 8 < O) *(reinterpret_cast<float4*>(dW + (int64_t)(o0 + og + 8) * I) + c4) = a1;
        }
    }
}

// backward of the above, both halves in ONE launch (block-role split):
//   blocks [0, nW):  dW_l[o][i] = sum_b dp[l][b][o] * y[b][i]       (written, thread per element)
//   blocks [nW, ..): dy[b][i] += sum_l sum_o W_l[o][i] * dp[l][b][o] (atomicAdd into ZEROED dy; a block owns (l, 64 rows of
//                    W_l, 32 columns i) and up to 16 samples per pass)
__global__ void __launch_bounds__(256) linear_bwd_batch_kernel(const float* __restrict__ dp, const float* __restrict__ y,
                                                               const int64_t* __restrict__ ptrs, int L, int B, int I, int O,
                                                               int nW, float* __restrict__ dy) {
    pdl_wait();
    if ((int)blockIdx.x < nW) {
        const int64_t e = (int64_t)blockIdx.x * 256 + threadIdx.x;
        if (e >= (int64_t)L * O * I) return;
        const int l = (int)(e / ((int64_t)O * I));
        const int64_t r = e - (int64_t)l * O * I;
        const int o = (int)(r / I), i = (int)(r - (int64_t)o * I);
        float* dW = reinterpret_cast<float*>(ptrs[4 * l + 3]);
        if (!dW) return;
        float s = 0.f;
        for (int b = 0; b < B; b++) s = fmaf(dp[((int64_t)l * B + b) * O + o], y[(int64_t)b * I + i], s);
        dW[r] = s;
        return;
    }
    if (!dy) return;
    // a block owns 32 columns i and a RANGE of the L * ceil(O / 64) weight-row tiles: partial sums of up to 16 samples stay
    // in registers across the tiles and are added to dy once per block (one block per tile meant 640 x 512 atomics on
    // 2048 addresses: 47 us)
    __shared__ float dps[16][64];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int it_n = (I + 31) / 32, ot_n = (O + 63) / 64;
    const int blk = (int)blockIdx.x - nW;
    const int it = blk % it_n, chunk = blk / it_n, nchunk = ((int)gridDim.x - nW) / it_n;
    const int ntile = L * ot_n;
    const int tile0 = (int)((int64_t)ntile * chunk / nchunk), tile1 = (int)((int64_t)ntile * (chunk + 1) / nchunk);
    const int i = it * 32 + tx;
    for (int b0 = 0; b0 < B; b0 += 16) {
        float s0 = 0.f, s1 = 0.f;
        for (int tile = tile0; tile < tile1; tile++) {
            const int l = tile / ot_n, o0 = (tile - l * ot_n) * 64;
            const float* W = reinterpret_cast<const float*>(ptrs[4 * l]);
            const int on = O - o0 < 64 ? O - o0 : 64;
            __syncthreads();
            for (int e = threadIdx.x; e < 16 * 64; e += 256) {
                const int bb = e >> 6, oo = e & 63;
                dps[bb][oo] = (b0 + bb < B && oo < on) ? dp[((int64_t)l * B + b0 + bb) * O + o0 + oo] : 0.f;
            }
            __syncthreads();
            if (i < I) {
#pragma unroll 8
                for (int oo = 0; oo < on; oo++) {
                    const float w = W[(int64_t)(o0 + oo) * I + i];
                    s0 = fmaf(w, dps[ty][oo], s0);
                    s1 = fmaf(w, dps[ty + 8][oo], s1);
                }
            }
        }
        if (i < I) {
            if (b0 + ty < B) atomicAdd(dy + (int64_t)(b0 + ty) * I + i, s0);
            if (b0 + ty + 8 < B) atomicAdd(dy + (int64_t)(b0 + ty + 8) * I + i, s1);
        }
    }
}

constexpr float LOG_2PI_F = 1.8378770664093453f;

__global__ void __launch_bounds__(256) logloss_fwd_kernel(const float* __restrict__ xhat, const float* __restrict__ x,
                                                          int64_t n, float scale, float* __restrict__ acc) {
    pdl_wait();
    __shared__ float red[32];
    float s = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float d = xhat[i] - x[i];
        s += 0.5f * (LOG_2PI_F + d * d);
    }
    s = block_sum(s, red);
    if (threadIdx.x == 0) atomicAdd(acc, s * scale);
}

__global__ void logloss_bwd_kernel(const float* __restrict__ xhat, const float* __restrict__ x,
                                   const float* __restrict__ g, int64_t n, float scale, float* __restrict__ dxhat) {
    pdl_wait();
    const float gs = (g ? g[0] : 1.f) * scale;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        dxhat[i] = gs * (xhat[i] - x[i]);
}

__global__ void __launch_bounds__(256) sqnorm_kernel(const float* __restrict__ g, int64_t n, float* __restrict__ acc) {
    pdl_wait();
    __shared__ float red[32];
    float s = 0.f;
    const int64_t n4 = n >> 2;
    const float4* g4 = reinterpret_cast<const float4*>(g);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 x = g4[i];
        s += (x.x * x.x + x.y * x.y) + (x.z * x.z + x.w * x.w);
    }
    for (int64_t i = (n4 << 2) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        s += g[i] * g[i];
    s = block_sum(s, red);
    if (threadIdx.x == 0) atomicAdd(acc, s);
}

__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                   float* __restrict__ m, float* __restrict__ v, int64_t n,
                                                   const float* __restrict__ sqn, float max_norm, float lr,
                                                   double beta1d, double beta2d, float eps,
                                                   const float* __restrict__ state, int step_size, float gamma) {
    pdl_wait();
    __shared__ float sh[3];
    // like torch.optim.Adam: 1 - beta is formed in double, then used in fp32 arithmetic
    const float beta1 = (float)beta1d, beta2 = (float)beta2d;
    const float omb1 = (float)(1.0 - beta1d), omb2 = (float)(1.0 - beta2d);
    if (threadIdx.x == 0) {
        const double t = (double)state[0] + 1.0;                 // Adam step (bias correction)
        const double s = (double)state[1];                       // StepLR steps taken so far
        double lr_t = (double)lr;
        if (step_size > 0) lr_t *= pow((double)gamma, floor(s / (double)step_size));
        const double bc1 = 1.0 - pow(beta1d, t), bc2 = 1.0 - pow(beta2d, t);
        float coef = 1.f;
        if (max_norm > 0.f && sqn) coef = fminf(1.f, max_norm / (sqrtf(sqn[0]) + 1e-6f));
        sh[0] = (float)(lr_t / bc1);
        sh[1] = (float)(1.0 / sqrt(bc2));
        sh[2] = coef;
    }
    __syncthreads();
    const float step = sh[0], rs2 = sh[1], coef = sh[2];
    // 16-byte accesses, 7 streams (4 in, 3 out); the flat buffers are 16-byte aligned (torch allocations)
    const int64_t n4 = n >> 2;
    const float4* g4 = reinterpret_cast<const float4*>(g);
    float4* m4 = reinterpret_cast<float4*>(m);
    float4* v4 = reinterpret_cast<float4*>(v);
    float4* p4 = reinterpret_cast<float4*>(p);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 gg = g4[i];
        float4 mm = m4[i], vv = v4[i], pp = p4[i];
        const float gx[4] = {gg.x * coef, gg.y * coef, gg.z * coef, gg.w * coef};
        float* mp = reinterpret_cast<float*>(&mm);
        float* vp = reinterpret_cast<float*>(&vv);
        float* pq = reinterpret_cast<float*>(&pp);
#pragma unroll
        for (int j = 0; j < 4; j++) {
            mp[j] = beta1 * mp[j] + omb1 * gx[j];
            vp[j] = beta2 * vp[j] + omb2 * gx[j] * gx[j];
            pq[j] -= step * mp[j] / (sqrtf(vp[j]) * rs2 + eps);
        }
        m4[i] = mm;
        v4[i] = vv;
        p4[i] = pp;
    }
    for (int64_t i = (n4 << 2) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float gi = g[i] * coef;
        const float mi = beta1 * m[i] + omb1 * gi;
        const float vi = beta2 * v[i] + omb2 * gi * gi;
        m[i] = mi;
        v[i] = vi;
        p[i] -= step * mi / (sqrtf(vi) * rs2 + eps);
    }
}

__global__ void adam_tick_kernel(float* state) {
    pdl_wait();
    state[0] += 1.f;
    state[1] += 1.f;
}

static inline int flat_grid(int64_t n) {
    int64_t g = (n + 255) / 256;
    const int64_t cap = 148 * 8;
    return (int)(g < cap ? g : cap);
}

}  // namespace vnb
using namespace vnb;

extern "C" int vnb_embed_fwd(const float* table, const int64_t* idx, int n, int dim, float* out, void* stream) {
    VNB_REQUIRE(table && idx && out && n > 0 && dim > 0, "embed_fwd: bad arguments");
    vnb::launch_k(embed_fwd_kernel, ceil_div((int64_t)n * dim, 256), 256, 0, (cudaStream_t)stream, table, idx, n, dim, out);
    VNB_LAUNCH_CHECK("embed_fwd");
    return VNB_OK;
}

extern "C" int vnb_embed_bwd(const float* dout, const int64_t* idx, int n, int dim, float* dtable, void* stream) {
    VNB_REQUIRE(dout && idx && dtable && n > 0 && dim > 0, "embed_bwd: bad arguments");
    vnb::launch_k(embed_bwd_kernel, ceil_div(dim, 128), 128, 0, (cudaStream_t)stream, dout, idx, n, dim, dtable);
    VNB_LAUNCH_CHECK("embed_bwd");
    return VNB_OK;
}

extern "C" int vnb_linear_fwd(const float* y, const float* W, const float* bias, const float* bias2, int B, int I,
                              int O, float* p, void* stream) {
    VNB_REQUIRE(y && W && p && B > 0 && I > 0 && O > 0, "linear_fwd: bad arguments");
    vnb::launch_k(linear_fwd_kernel, ceil_div((int64_t)B * O, 8), 256, 0, (cudaStream_t)stream, y, W, bias, bias2, B, I, O, p);
    VNB_LAUNCH_CHECK("linear_fwd");
    return VNB_OK;
}

extern "C" int vnb_linear_bwd(const float* dp, const float* y, const float* W, int B, int I, int O, float* dW,
                              float* dy, int accumulate_dy, void* stream) {
    VNB_REQUIRE(dp && y && W && B > 0 && I > 0 && O > 0, "linear_bwd: bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    if (dW) {
        vnb::launch_k(linear_bwd_w_kernel, ceil_div((int64_t)O * I, 256), 256, 0, st, dp, y, B, I, O, dW);
        VNB_LAUNCH_CHECK("linear_bwd_w");
    }
    if (dy) {
        if (!accumulate_dy && cudaMemsetAsync(dy, 0, (size_t)B * I * sizeof(float), st) != cudaSuccess) {
            set_error("linear_bwd: memset failed");
            return VNB_E_CUDA;
        }
        vnb::launch_k(linear_bwd_y_kernel, dim3(ceil_div(I, 32), ceil_div(O, 64), ceil_div(B, 16)), 256, 0, st, dp, W, B, I, O, dy);
        VNB_LAUNCH_CHECK("linear_bwd_y");
    }
    return VNB_OK;
}

extern "C" int vnb_linear_fwd_batch(const float* y, const int64_t* ptrs, int L, int B, int I, int O, float* p, void* stream) {
    VNB_REQUIRE(y && ptrs && p && L > 0 && B > 0 && I > 0 && O > 0, "linear_fwd_batch: bad arguments");
    if ((I & 3) == 0 && I <= 1024 && ((uintptr_t)y & 15) == 0 && vnb::option(VNB_OPT_EW_V2) != 0) {
        // (weight rows are 16-byte aligned when I % 4 == 0 and the matrices are: the bank's fp32 arena and torch allocations are)
        vnb::launch_k(linear_fwd_batch_v2_kernel, L * ceil_div(O, 16), 256, (size_t)2 * 16 * (I + 4) * sizeof(float),
                      (cudaStream_t)stream, y, ptrs, L, B, I, O, p);
        VNB_LAUNCH_CHECK("linear_fwd_batch_v2");
        return VNB_OK;
    }
    vnb::launch_k(linear_fwd_batch_kernel, ceil_div((int64_t)L * O, 8), 256, 0, (cudaStream_t)stream, y, ptrs, L, B, I, O, p);
    VNB_LAUNCH_CHECK("linear_fwd_batch");
    return VNB_OK;
}

extern "C" int vnb_linear_bwd_batch(const float* dp, const float* y, const int64_t* ptrs, int L, int B, int I, int O,
                                    float* dy, void* stream) {
    VNB_REQUIRE(dp && y && ptrs && L > 0 && B > 0 && I > 0 && O > 0, "linear_bwd_batch: bad arguments");
    const bool v2 = (I & 3) == 0 && I <= 1024 && ((uintptr_t)y & 15) == 0 && vnb::option(VNB_OPT_EW_V2) != 0;
    if (v2) {
        vnb::launch_k(linear_bwd_w_batch_v2_kernel, L * ceil_div(O, 16), 256, (size_t)(16 * I + 256) * sizeof(float),
                      (cudaStream_t)stream, dp, y, ptrs, L, B, I, O);
        VNB_LAUNCH_CHECK("linear_bwd_w_batch_v2");
        if (!dy) return VNB_OK;
    }
    const int nW = v2 ? 0 : ceil_div((int64_t)L * O * I, 256);
    // dy blocks: ceil(I / 32) column groups x chunks of the L * ceil(O / 64) weight-row tiles, about two blocks per SM
    const int it_n = ceil_div(I, 32), ntile = L * ceil_div(O, 64);
    int nchunk = ceil_div(296, it_n);
    if (nchunk > ntile) nchunk = ntile;
    const int nY = dy ? it_n * nchunk : 0;
    vnb::launch_k(linear_bwd_batch_kernel, nW + nY, 256, 0, (cudaStream_t)stream, dp, y, ptrs, L, B, I, O, nW, dy);
    VNB_LAUNCH_CHECK("linear_bwd_batch");
    return VNB_OK;
}

extern "C" int vnb_logloss_fwd(const float* xhat, const float* x, int64_t n, float scale, float* acc, void* stream) {
    VNB_REQUIRE(xhat && x && acc && n > 0, "logloss_fwd: bad arguments");
    vnb::launch_k(logloss_fwd_kernel, flat_grid(n), 256, 0, (cudaStream_t)stream, xhat, x, n, scale, acc);
    VNB_LAUNCH_CHECK("logloss_fwd");
    return VNB_OK;
}

extern "C" int vnb_logloss_bwd(const float* xhat, const float* x, const float* g, int64_t n, float scale, float* dxhat,
                               void* stream) {
    VNB_REQUIRE(xhat && x && dxhat && n > 0, "logloss_bwd: bad arguments");
    vnb::launch_k(logloss_bwd_kernel, flat_grid(n), 256, 0, (cudaStream_t)stream, xhat, x, g, n, scale, dxhat);
    VNB_LAUNCH_CHECK("logloss_bwd");
    return VNB_OK;
}

extern "C" int vnb_sqnorm(const float* g, int64_t n, float* acc, void* stream) {
    VNB_REQUIRE(g && acc && n > 0 && ((uintptr_t)g & 15) == 0, "sqnorm: bad arguments (g must be 16-byte aligned)");
    vnb::launch_k(sqnorm_kernel, flat_grid(n), 256, 0, (cudaStream_t)stream, g, n, acc);
    VNB_LAUNCH_CHECK("sqnorm");
    return VNB_OK;
}

extern "C" int vnb_adam_step(float* p, const float* g, float* m, float* v, int64_t n, const float* sqn, float max_norm,
                             float lr, double beta1, double beta2, float eps, float* state, int step_size, float gamma,
                             void* stream) {
    VNB_REQUIRE(p && g && m && v && state && n > 0, "adam_step: bad arguments");
    VNB_REQUIRE((((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) == 0, "adam_step: buffers must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    vnb::launch_k(adam_kernel, flat_grid(n), 256, 0, st, p, g, m, v, n, sqn, max_norm, lr, beta1, beta2, eps, state, step_size, gamma);
    VNB_LAUNCH_CHECK("adam_step");
    vnb::launch_k(adam_tick_kernel, 1, 1, 0, st, state);
    VNB_LAUNCH_CHECK("adam_tick");
    return VNB_OK;
}
