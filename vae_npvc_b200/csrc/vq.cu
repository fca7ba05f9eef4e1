// Vector-quantiser kernels around the distance GEMM: codebook / frame preparation, gather + loss
// sums + histogram, scalar finalisation, backward, EMA codebook update, jitter.
// Reference: vae_npvc/model/layers_vq.py:28-33, 79-150 (VectorQuantizer), 192-233, 268-323 (EMA),
// 353-379 (Jitter).  Math: SURVEY.md section 9 C/D/E.
#include "common.cuh"

namespace vnb {

// one block (128 threads) per code row
__global__ void __launch_bounds__(128) vq_prep_codebook_kernel(float* __restrict__ E, int K, int D, int normalize,
                                                               int inplace_renorm, float* __restrict__ e_hat,
                                                               bf16* __restrict__ e_planes,
                                                               float* __restrict__ e_sq) {
    pdl_wait();
    __shared__ float red[32];
    const int k = blockIdx.x;
    float* row = E + (int64_t)k * D;
    const int Dp = (D + 7) & ~7;
    const int64_t plane = (int64_t)K * Dp;
    float ss = 0.f;
    for (int d = threadIdx.x; d < D; d += blockDim.x) ss += row[d] * row[d];
    ss = block_sum(ss, red);
    float s1 = 1.f, s2 = 1.f;
    if (normalize) {
        s1 = inplace_renorm ? 1.f / sqrtf(ss) : 1.f;      // embed_norm(): E *= (1/||E||)
        // second, differentiable normalisation of the (renormalised) rows: e_hat = E / ||E||
        float ss2 = 0.f;
        for (int d = threadIdx.x; d < D; d += blockDim.x) {
            const float v = row[d] * s1;
            ss2 += v * v;
        }
        ss2 = block_sum(ss2, red);
        s2 = sqrtf(ss2);
    }
    float q = 0.f;
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        const float v = row[d] * s1;
        if (normalize && inplace_renorm) row[d] = v;
        const float e = normalize ? v / s2 : v;
        if (e_hat) e_hat[(int64_t)k * D + d] = e;
        if (e_planes) store_planes(e_planes, (int64_t)k * Dp + d, plane, 3, e);
        q += e * e;
    }
    q = block_sum(q, red);
    if (threadIdx.x == 0 && e_sq) e_sq[k] = q;
}

// Frames are rows of the channels-last latent: z (N = B*T, D).  One warp per frame, lanes over d.
constexpr int VQ_WARPS = 8;

__global__ void __launch_bounds__(VQ_WARPS * 32) vq_prep_z_kernel(const float* __restrict__ z, int64_t N, int D, int Dp,
                                                                  int normalize, float* __restrict__ zn,
                                                                  bf16* __restrict__ z_planes, float* __restrict__ z_sq,
                                                                  float* __restrict__ z_norm) {
    pdl_wait();
    const int64_t n = (int64_t)blockIdx.x * VQ_WARPS + (threadIdx.x >> 5);
    if (n >= N) return;
    const int lane = threadIdx.x & 31;
    const float* p = z + n * D;
    float ss = 0.f;
    for (int d = lane; d < D; d += 32) ss += p[d] * p[d];
    ss = warp_sum(ss);
    const float nrm = sqrtf(ss);
    const int64_t plane = N * Dp;
    float q = 0.f;
    for (int d = lane; d < D; d += 32) {
        float v = p[d];
        if (normalize) v = v / nrm;
        if (zn) zn[n * D + d] = v;
        if (z_planes) store_planes(z_planes, n * Dp + d, plane, 3, v);
        q += v * v;
    }
    q = warp_sum(q);
    if (lane == 0) {
        if (z_sq) z_sq[n] = q;
        if (z_norm) z_norm[n] = nrm;
    }
}

__global__ void __launch_bounds__(VQ_WARPS * 32) vq_gather_kernel(const int64_t* __restrict__ idx,
                                                                  const float* __restrict__ e_hat,
                                                                  const float* __restrict__ zn,
                                                                  const float* __restrict__ z, int64_t N, int D,
                                                                  float* __restrict__ q_out, float* __restrict__ sums,
                                                                  int* __restrict__ counts) {
    pdl_wait();
    __shared__ float red[32];
    const int64_t n = (int64_t)blockIdx.x * VQ_WARPS + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    float lq = 0.f, ln = 0.f;
    if (n < N) {
        const int k = (int)idx[n];
        const float* er = e_hat + (int64_t)k * D;
        for (int d = lane; d < D; d += 32) {
            const int64_t o = n * D + d;
            const float q = er[d];
            if (q_out) q_out[o] = q;
            if (zn) {                                   // null: rows only (Model.decode, layers_vq.py:62-77)
                const float a = zn[o];
                lq += (q - a) * (q - a);
                if (z) ln += (a - z[o]) * (a - z[o]);
            }
        }
        if (counts && lane == 0) atomicAdd(&counts[k], 1);
    }
    lq = block_sum(lq, red);
    ln = block_sum(ln, red);
    if (threadIdx.x == 0 && sums) {
        atomicAdd(&sums[0], lq);
        atomicAdd(&sums[1], ln);
    }
}

__global__ void __launch_bounds__(256) vq_finalize_kernel(const float* __restrict__ sums, const int* __restrict__ counts,
                                                          int K, float N, float scale, int normalize, float eps,
                                                          float* __restrict__ scal) {
    pdl_wait();
    __shared__ float red[32];
    float h = 0.f;
    for (int k = threadIdx.x; k < K; k += blockDim.x) {
        const float p = (float)counts[k] / N;
        h += p * logf(p + eps);
    }
    h = block_sum(h, red);
    if (threadIdx.x == 0) {
        scal[0] = sums[0] * scale;                                       // z_qut_loss
        scal[1] = (sums[0] + (normalize ? sums[1] : 0.f)) * scale;       // z_enc_loss (value)
        scal[2] = expf(-h);                                              // perplexity
    }
}

// dz and dE_hat (atomic scatter).  One warp per frame.
__global__ void __launch_bounds__(VQ_WARPS * 32) vq_bwd_kernel(const float* __restrict__ g_out,
                                                               const float* __restrict__ g_lq,
                                                               const float* __restrict__ g_le,
                                                               const int64_t* __restrict__ idx,
                                                               const float* __restrict__ z, const float* __restrict__ zn,
                                                               const float* __restrict__ z_norm,
                                                               const float* __restrict__ e_hat, int64_t N, int D,
                                                               float scale, int normalize, float* __restrict__ dz,
                                                               float* __restrict__ dE_hat) {
    pdl_wait();
    const int64_t n = (int64_t)blockIdx.x * VQ_WARPS + (threadIdx.x >> 5);
    if (n >= N) return;
    const int lane = threadIdx.x & 31;
    const float glq = (g_lq ? g_lq[0] : 0.f) * 2.f * scale;
    const float gle = (g_le ? g_le[0] : 0.f) * 2.f * scale;
    const int k = (int)idx[n];
    const float* er = e_hat + (int64_t)k * D;
    if (!normalize) {
        for (int d = lane; d < D; d += 32) {
            const int64_t o = n * D + d;
            const float q = er[d], a = zn[o];
            dz[o] = (g_out ? g_out[o] : 0.f) + gle * (a - q);
            if (dE_hat && glq != 0.f) atomicAdd(&dE_hat[(int64_t)k * D + d], glq * (q - a));
        }
        return;
    }
    // normalised: dzn = g_out + gle*[(zn-q) + (zn-z)];  dz = (dzn - zn*<zn,dzn>)/||z|| - gle*(zn-z)
    const float inv = 1.f / z_norm[n];
    float dot = 0.f;
    for (int d = lane; d < D; d += 32) {
        const int64_t o = n * D + d;
        const float q = er[d], a = zn[o];
        const float dzn = (g_out ? g_out[o] : 0.f) + gle * ((a - q) + (a - z[o]));
        dot += a * dzn;
        if (dE_hat && glq != 0.f) atomicAdd(&dE_hat[(int64_t)k * D + d], glq * (q - a));
    }
    dot = warp_sum(dot);
    for (int d = lane; d < D; d += 32) {
        const int64_t o = n * D + d;
        const float q = er[d], a = zn[o];
        const float dzn = (g_out ? g_out[o] : 0.f) + gle * ((a - q) + (a - z[o]));
        dz[o] = (dzn - a * dot) * inv - gle * (a - z[o]);
    }
}

__global__ void __launch_bounds__(128) vq_bwd_codebook_kernel(const float* __restrict__ dE_hat,
                                                              const float* __restrict__ E, int D, int normalize,
                                                              float* __restrict__ dE) {
    pdl_wait();
    __shared__ float red[32];
    const int k = blockIdx.x;
    const float* e = E + (int64_t)k * D;
    const float* g = dE_hat + (int64_t)k * D;
    if (!normalize) {
        for (int d = threadIdx.x; d < D; d += blockDim.x) dE[(int64_t)k * D + d] = g[d];
        return;
    }
    float ss = 0.f, dot = 0.f;
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        ss += e[d] * e[d];
        dot += e[d] * g[d];
    }
    ss = block_sum(ss, red);
    dot = block_sum(dot, red);
    const float inv = 1.f / sqrtf(ss);
    for (int d = threadIdx.x; d < D; d += blockDim.x)
        dE[(int64_t)k * D + d] = (g[d] - e[d] * dot / ss) * inv;
}

// EMA: per-code sums of assigned frames (one warp per frame)
__global__ void __launch_bounds__(VQ_WARPS * 32) vq_ema_scatter_kernel(const int64_t* __restrict__ idx,
                                                                       const float* __restrict__ z, int64_t N, int D,
                                                                       float* __restrict__ zsum,
                                                                       int* __restrict__ counts) {
    pdl_wait();
    const int64_t n = (int64_t)blockIdx.x * VQ_WARPS + (threadIdx.x >> 5);
    if (n >= N) return;
    const int lane = threadIdx.x & 31;
    const int k = (int)idx[n];
    for (int d = lane; d < D; d += 32) atomicAdd(&zsum[(int64_t)k * D + d], z[n * D + d]);
    if (lane == 0) atomicAdd(&counts[k], 1);
}

// single block: EMA buffers, alive mask, new codebook, statistics
__global__ void __launch_bounds__(512) vq_ema_update_kernel(const float* __restrict__ zsum,
                                                            const int* __restrict__ counts,
                                                            const float* __restrict__ rand_rows, int K, int D,
                                                            float N, float mu, float threshold,
                                                            float* __restrict__ emb_sum, float* __restrict__ emb_elem,
                                                            float* __restrict__ embeddings,
                                                            float* __restrict__ stats) {
    pdl_wait();
    __shared__ float red[32];
    float ent = 0.f, used = 0.f, alive_n = 0.f, diff = 0.f;
    for (int k = threadIdx.x; k < K; k += blockDim.x) {
        const float c = (float)counts[k];
        const float el = mu * emb_elem[k] + (1.f - mu) * c;
        emb_elem[k] = el;
        const float p = c / N;
        ent += p * logf(p + 1e-8f);
        used += (c >= threshold) ? 1.f : 0.f;
        alive_n += (el >= threshold) ? 1.f : 0.f;
    }
    __syncthreads();
    for (int e = threadIdx.x; e < K * D; e += blockDim.x) {
        const int k = e / D;
        const float s = mu * emb_sum[e] + (1.f - mu) * zsum[e];
        emb_sum[e] = s;
        const float el = emb_elem[k];
        const float nv = (el >= threshold) ? s / el : rand_rows[e];
        const float dlt = nv - embeddings[e];
        diff += dlt * dlt;
        embeddings[e] = nv;
    }
    ent = block_sum(ent, red);
    used = block_sum(used, red);
    alive_n = block_sum(alive_n, red);
    diff = block_sum(diff, red);
    if (threadIdx.x == 0) {
        stats[0] = expf(-ent);
        stats[1] = used;
        stats[2] = alive_n;
        stats[3] = sqrtf(diff) / sqrtf((float)K * (float)D);
    }
}

__global__ void vq_frame_rows_kernel(const float* __restrict__ z, const int64_t* __restrict__ perm, int64_t N, int D,
                                     int K, float* __restrict__ out) {
    pdl_wait();
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)K * D) return;
    const int k = (int)(i / D), d = (int)(i - (int64_t)k * D);
    out[i] = z[(perm[k] % N) * D + d];
}

// channels-last (B, T, D): out[b,t,:] = in[b,src[t],:]  (backward: zero the rows with src[t] != t)
__global__ void jitter_kernel(const float* __restrict__ in, const int64_t* __restrict__ src, int64_t n, int D, int T,
                              int backward, float* __restrict__ out) {
    pdl_wait();
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t r = i / D;
    const int d = (int)(i - r * D);
    const int64_t b = r / T;
    const int t = (int)(r - b * T);
    const int s = (int)src[t];
    out[i] = backward ? (s == t ? in[i] : 0.f) : in[(b * T + s) * D + d];
}

// Jitter's per-frame draws (layers_vq.py:360-378) made on the device, so that a recipe with jitter_p != 0 stays inside
// the captured CUDA graph.  state[0] = seed, state[1] = call counter (incremented here): every replay draws afresh.
// Same distribution as the reference's host loop, including its inverted flag (a frame is replaced with probability
// 1 - p, layers_vq.py:364): first frame -> t+1, last -> t-1, others t-1 / t+1 with probability 1/2.
__device__ __forceinline__ uint64_t mix64(uint64_t z) {          // splitmix64 finaliser
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__global__ void jitter_sources_kernel(int T, float p, int64_t* __restrict__ state, int64_t* __restrict__ src) {
    pdl_wait();
    const uint64_t seed = (uint64_t)state[0], call = (uint64_t)state[1];
    for (int t = threadIdx.x; t < T; t += blockDim.x) {
        const uint64_t h = mix64(mix64(seed ^ (call * 0xD1B54A32D192ED03ull)) + (uint64_t)t);
        const float u1 = (float)(h >> 40) * (1.0f / 16777216.0f);            // [0, 1)
        const bool replace = u1 >= p;                                          // probability 1 - p (reference behaviour)
        int s = t;
        if (replace && T > 1) {
            if (t == 0) s = 1;
            else if (t == T - 1) s = T - 2;
            else s = t + (((h >> 8) & 1) ? 1 : -1);
        }
        src[t] = s;
    }
    __syncthreads();
    if (threadIdx.x == 0) state[1] = (int64_t)(call + 1);
}

}  // namespace vnb
using namespace vnb;

extern "C" int vnb_jitter_sources(int T, float p, int64_t* state, int64_t* src, void* stream) {
    VNB_REQUIRE(state && src && T > 0 && p >= 0.f && p <= 1.f, "jitter_sources: bad arguments");
    vnb::launch_k(jitter_sources_kernel, 1, 256, 0, (cudaStream_t)stream, T, p, state, src);
    VNB_LAUNCH_CHECK("jitter_sources");
    return VNB_OK;
}

extern "C" int vnb_vq_prep_codebook(float* E, int K, int D, int normalize, int inplace_renorm, float* e_hat,
                                    void* e_planes, float* e_sq, void* stream) {
    VNB_REQUIRE(E && K > 0 && D > 0, "vq_prep_codebook: bad arguments");
    vnb::launch_k(vq_prep_codebook_kernel, K, 128, 0, (cudaStream_t)stream, E, K, D, normalize, inplace_renorm, e_hat,
                                                                 (bf16*)e_planes, e_sq);
    VNB_LAUNCH_CHECK("vq_prep_codebook");
    return VNB_OK;
}

extern "C" int vnb_vq_prep_z(const float* z, int B, int D, int T, int normalize, float* zn, void* z_planes,
                             float* z_sq, float* z_norm, void* stream) {
    VNB_REQUIRE(z && B > 0 && D > 0 && T > 0, "vq_prep_z: bad arguments");
    const int64_t N = (int64_t)B * T;
    vnb::launch_k(vq_prep_z_kernel, ceil_div(N, VQ_WARPS), VQ_WARPS * 32, 0, (cudaStream_t)stream, z, N, D, round_up8(D), normalize,
                                                                                        zn, (bf16*)z_planes, z_sq, z_norm);
    VNB_LAUNCH_CHECK("vq_prep_z");
    return VNB_OK;
}

extern "C" int vnb_vq_gather(const int64_t* idx, const float* e_hat, const float* zn, const float* z, int B, int D,
                             int T, int K, float* q_out, float* sums, int* counts, void* stream) {
    VNB_REQUIRE(idx && e_hat && (zn || (!sums && !z)) && B > 0 && D > 0 && T > 0 && K > 0, "vq_gather: bad arguments (zn may be null only for a pure gather)");
    const int64_t N = (int64_t)B * T;
    vnb::launch_k(vq_gather_kernel, ceil_div(N, VQ_WARPS), VQ_WARPS * 32, 0, (cudaStream_t)stream, idx, e_hat, zn, z, N, D, q_out,
                                                                                        sums, counts);
    VNB_LAUNCH_CHECK("vq_gather");
    return VNB_OK;
}

extern "C" int vnb_vq_finalize(const float* sums, const int* counts, int K, int64_t N, float scale, int normalize,
                               float eps, float* scal, void* stream) {
    VNB_REQUIRE(sums && counts && scal && K > 0 && N > 0, "vq_finalize: bad arguments");
    vnb::launch_k(vq_finalize_kernel, 1, 256, 0, (cudaStream_t)stream, sums, counts, K, (float)N, scale, normalize, eps, scal);
    VNB_LAUNCH_CHECK("vq_finalize");
    return VNB_OK;
}

extern "C" int vnb_vq_bwd(const float* g_out, const float* g_lq, const float* g_le, const int64_t* idx, const float* z,
                          const float* zn, const float* z_norm, const float* e_hat, const float* E, int B, int D, int T,
                          int K, float scale, int normalize, float* dz, float* dE_hat, void* stream) {
    (void)E;
    (void)K;
    VNB_REQUIRE(idx && zn && e_hat && dz && B > 0 && D > 0 && T > 0, "vq_bwd: bad arguments");
    VNB_REQUIRE(!normalize || (z && z_norm), "vq_bwd: normalised mode needs z and z_norm");
    const int64_t N = (int64_t)B * T;
    vnb::launch_k(vq_bwd_kernel, ceil_div(N, VQ_WARPS), VQ_WARPS * 32, 0, (cudaStream_t)stream, 
        g_out, g_lq, g_le, idx, z, zn, z_norm, e_hat, N, D, scale, normalize, dz, dE_hat);
    VNB_LAUNCH_CHECK("vq_bwd");
    return VNB_OK;
}

extern "C" int vnb_vq_bwd_codebook(const float* dE_hat, const float* E, int K, int D, int normalize, float* dE,
                                   void* stream) {
    VNB_REQUIRE(dE_hat && E && dE && K > 0 && D > 0, "vq_bwd_codebook: bad arguments");
    vnb::launch_k(vq_bwd_codebook_kernel, K, 128, 0, (cudaStream_t)stream, dE_hat, E, D, normalize, dE);
    VNB_LAUNCH_CHECK("vq_bwd_codebook");
    return VNB_OK;
}

extern "C" int vnb_vq_ema_scatter(const int64_t* idx, const float* z, int B, int D, int T, int K, float* zsum,
                                  int* counts, void* stream) {
    (void)K;
    VNB_REQUIRE(idx && z && zsum && counts && B > 0 && D > 0 && T > 0, "vq_ema_scatter: bad arguments");
    const int64_t N = (int64_t)B * T;
    vnb::launch_k(vq_ema_scatter_kernel, ceil_div(N, VQ_WARPS), VQ_WARPS * 32, 0, (cudaStream_t)stream, idx, z, N, D, zsum, counts);
    VNB_LAUNCH_CHECK("vq_ema_scatter");
    return VNB_OK;
}

extern "C" int vnb_vq_ema_update(const float* zsum, const int* counts, int K, int D, int64_t n_frames, float mu,
                                 float threshold, const float* rand_rows, float* emb_sum, float* emb_elem,
                                 float* embeddings, float* stats, void* stream) {
    VNB_REQUIRE(zsum && counts && rand_rows && emb_sum && emb_elem && embeddings && stats && K > 0 && D > 0 && n_frames > 0,
                "vq_ema_update: bad arguments");
    vnb::launch_k(vq_ema_update_kernel, 1, 512, 0, (cudaStream_t)stream, zsum, counts, rand_rows, K, D, (float)n_frames, mu,
                                                              threshold, emb_sum, emb_elem, embeddings, stats);
    VNB_LAUNCH_CHECK("vq_ema_update");
    return VNB_OK;
}

extern "C" int vnb_vq_frame_rows(const float* z, const int64_t* perm, int B, int D, int T, int K, float* out,
                                 void* stream) {
    VNB_REQUIRE(z && perm && out && B > 0 && D > 0 && T > 0 && K > 0, "vq_frame_rows: bad arguments");
    vnb::launch_k(vq_frame_rows_kernel, ceil_div((int64_t)K * D, 256), 256, 0, (cudaStream_t)stream, z, perm, (int64_t)B * T, D, K, out);
    VNB_LAUNCH_CHECK("vq_frame_rows");
    return VNB_OK;
}

extern "C" int vnb_jitter_fwd(const float* in, const int64_t* src, int B, int D, int T, float* out, void* stream) {
    VNB_REQUIRE(in && src && out && B > 0 && D > 0 && T > 0, "jitter_fwd: bad arguments");
    const int64_t n = (int64_t)B * D * T;
    vnb::launch_k(jitter_kernel, ceil_div(n, 256), 256, 0, (cudaStream_t)stream, in, src, n, D, T, 0, out);
    VNB_LAUNCH_CHECK("jitter_fwd");
    return VNB_OK;
}

extern "C" int vnb_jitter_bwd(const float* dout, const int64_t* src, int B, int D, int T, float* din, void* stream) {
    VNB_REQUIRE(dout && src && din && B > 0 && D > 0 && T > 0, "jitter_bwd: bad arguments");
    const int64_t n = (int64_t)B * D * T;
    vnb::launch_k(jitter_kernel, ceil_div(n, 256), 256, 0, (cudaStream_t)stream, dout, src, n, D, T, 1, din);
    VNB_LAUNCH_CHECK("jitter_bwd");
    return VNB_OK;
}
