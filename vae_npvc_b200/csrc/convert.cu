// Packed variable-length conversion / token extraction.
//
// The reference converts one utterance per forward call (vae_npvc/decoder/basic.py:58-75,
// vae_npvc/bin/extract_bnf.py:47-67) because utterances differ in length.  Here utterances of different
// lengths share one padded (B, Tmax, C) batch; lengths[b] frames of sample b are valid.  Every op of the
// path is per sample, and only two things see the padding:
//   * the k=3 convolutions read frame len[b] when producing frame len[b]-1 - it must read as the conv's zero
//     padding, so the producers' outputs are re-zeroed beyond len[b]            -> vnb_mask_rows
//   * GroupNorm statistics run over all (C, T) of a sample (layers.py:154,201)     -> the sums the conv epilogue
//     accumulated over all Tmax frames get the padded frames' share taken out again (vnb_gn_sums_unpad) and are
//     finalised with n = (C/groups) * len[b]                                      -> vnb_gn_finalize_masked
// Both touch only the padded frames, so the cost is proportional to the padding, not to the batch.
// vnb_csid is extract_bnf.py:60's unique_consecutive() for every utterance of the batch in one launch.
#include "common.cuh"

namespace vnb {

constexpr int MASK_ROWS_PER_CTA = 8;

// grid (ceil(T / MASK_ROWS_PER_CTA), B), block 256.  Zeroes frames t >= lengths[b] of an fp32 activation
// (B, T, C) and of up to three operand-plane buffers (P, B, T, Cp) (whole pitch rows: the padding stays finite).
__global__ void __launch_bounds__(256) mask_rows_kernel(float* __restrict__ x, bf16* __restrict__ pa,
                                                        bf16* __restrict__ pb, bf16* __restrict__ pc, int B, int C,
                                                        int T, int Cp, int P, const int* __restrict__ lengths) {
    pdl_wait();
    const int b = blockIdx.y;
    int len = lengths[b];
    len = len < 0 ? 0 : (len > T ? T : len);
    const int t0 = max(blockIdx.x * MASK_ROWS_PER_CTA, len);
    const int t1 = min((int)(blockIdx.x + 1) * MASK_ROWS_PER_CTA, T);
    if (t0 >= t1) return;
    const int nrow = t1 - t0;
    if (x) {
        float* p = x + ((int64_t)b * T + t0) * C;
        const int n = nrow * C;
        if ((C & 3) == 0) {
            float4* p4 = reinterpret_cast<float4*>(p);
            for (int i = threadIdx.x; i < n / 4; i += blockDim.x) p4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        } else {
            for (int i = threadIdx.x; i < n; i += blockDim.x) p[i] = 0.f;
        }
    }
    bf16* planes[3] = {pa, pb, pc};
    const int64_t plane = (int64_t)B * T * Cp;
    const int nv = nrow * Cp / 8;                                    // Cp % 8 == 0: 16-byte stores
#pragma unroll
    for (int k = 0; k < 3; k++) {
        if (!planes[k]) continue;
        for (int p = 0; p < P; p++) {
            uint4* q = reinterpret_cast<uint4*>(planes[k] + p * plane + ((int64_t)b * T + t0) * Cp);
            for (int i = threadIdx.x; i < nv; i += blockDim.x) q[i] = make_uint4(0u, 0u, 0u, 0u);
        }
    }
}

constexpr int UNPAD_MAX_GROUPS = 8;

// grid (ceil(T / MASK_ROWS_PER_CTA), B), block 256.  gn_sums[b, g, 0..1] -= (sum, sum of squares) of h over the
// frames t >= lengths[b] of this CTA's chunk.  fp64 like the conv epilogue's accumulation, so taking the padded
// share out again is exact to fp64 rounding.
__global__ void __launch_bounds__(256) gn_sums_unpad_kernel(const float* __restrict__ h, int C, int T, int groups,
                                                            const int* __restrict__ lengths,
                                                            double* __restrict__ gn_sums) {
    pdl_wait();
    __shared__ double acc[UNPAD_MAX_GROUPS][2];
    const int b = blockIdx.y;
    int len = lengths[b];
    len = len < 0 ? 0 : (len > T ? T : len);
    const int t0 = max(blockIdx.x * MASK_ROWS_PER_CTA, len);
    const int t1 = min((int)(blockIdx.x + 1) * MASK_ROWS_PER_CTA, T);
    if (t0 >= t1) return;
    if (threadIdx.x < UNPAD_MAX_GROUPS * 2) acc[threadIdx.x >> 1][threadIdx.x & 1] = 0.0;
    __syncthreads();
    const int cg = C / groups;
    const float* p = h + ((int64_t)b * T + t0) * C;
    const int n = (t1 - t0) * C;
    // a thread's channel moves by blockDim mod C per step: keep one (group, sum, sumsq) run in registers and
    // flush it to shared memory when the group changes
    double s = 0.0, ss = 0.0;
    int cur = -1;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int g = (i % C) / cg;
        if (g != cur) {
            if (cur >= 0) {
                atomicAdd(&acc[cur][0], s);
                atomicAdd(&acc[cur][1], ss);
            }
            s = ss = 0.0;
            cur = g;
        }
        const double v = (double)p[i];
        s += v;
        ss += v * v;
    }
    if (cur >= 0) {
        atomicAdd(&acc[cur][0], s);
        atomicAdd(&acc[cur][1], ss);
    }
    __syncthreads();
    if (threadIdx.x < groups * 2) {
        const int g = threadIdx.x >> 1, j = threadIdx.x & 1;
        atomicAdd(gn_sums + ((int64_t)b * groups + g) * 2 + j, -acc[g][j]);
    }
}

// mean / rstd per (sample, group) from fp64 (sum, sum of squares) with n = cg * lengths[b] values
__global__ void gn_finalize_masked_kernel(const double* __restrict__ sums, int B, int groups, int cg, int T,
                                          const int* __restrict__ lengths, float eps, float* __restrict__ mean,
                                          float* __restrict__ rstd) {
    pdl_wait();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * groups) return;
    int len = lengths[i / groups];
    len = len < 0 ? 0 : (len > T ? T : len);
    if (len == 0) {                                   // an empty slot of the batch: keep everything finite
        mean[i] = 0.f;
        rstd[i] = 0.f;
        return;
    }
    const double n = (double)cg * (double)len;
    const double m = sums[2 * i] / n;
    double var = sums[2 * i + 1] / n - m * m;
    if (var < 0.0) var = 0.0;
    mean[i] = (float)m;
    rstd[i] = (float)(1.0 / sqrt(var + (double)eps));
}

// One CTA per utterance: out[b, 0..n_b) = the ids of utterance b with consecutive repeats collapsed
// (torch.unique_consecutive, extract_bnf.py:60), out_len[b] = n_b; out[b, n_b..T) = -1.
// Chunks of blockDim frames: head flags -> block-wide exclusive scan (warp shuffles + one shared array).
__global__ void __launch_bounds__(256) csid_kernel(const int64_t* __restrict__ idx, int T,
                                                   const int* __restrict__ lengths, int64_t* __restrict__ out,
                                                   int* __restrict__ out_len) {
    pdl_wait();
    __shared__ int warp_tot[8];
    __shared__ int base_s;
    const int b = blockIdx.x, lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int len = lengths ? lengths[b] : T;
    len = len < 0 ? 0 : (len > T ? T : len);
    const int64_t* in = idx + (int64_t)b * T;
    int64_t* o = out + (int64_t)b * T;
    if (threadIdx.x == 0) base_s = 0;
    __syncthreads();
    for (int t0 = 0; t0 < len; t0 += blockDim.x) {
        const int t = t0 + threadIdx.x;
        int64_t v = 0;
        int head = 0;
        if (t < len) {
            v = in[t];
            head = (t == 0 || in[t - 1] != v) ? 1 : 0;
        }
        int incl = head;                                             // inclusive scan inside the warp
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int u = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += u;
        }
        if (lane == 31) warp_tot[w] = incl;
        __syncthreads();
        int before = base_s;
        for (int j = 0; j < w; j++) before += warp_tot[j];
        if (head) o[before + incl - 1] = v;
        __syncthreads();
        if (threadIdx.x == 0) {
            int tot = 0;
            for (int j = 0; j < (int)(blockDim.x >> 5); j++) tot += warp_tot[j];
            base_s += tot;
        }
        __syncthreads();
    }
    const int n = base_s;
    for (int t = n + threadIdx.x; t < T; t += blockDim.x) o[t] = -1;
    if (threadIdx.x == 0) out_len[b] = n;
}

}  // namespace vnb
using namespace vnb;

#define ST ((cudaStream_t)stream)

extern "C" int vnb_mask_rows(float* x, void* planes_a, void* planes_b, void* planes_c, int B, int C, int T, int P,
                             const int* lengths, void* stream) {
    VNB_REQUIRE(lengths && (x || planes_a || planes_b || planes_c), "mask_rows: null pointer");
    VNB_REQUIRE(B > 0 && B <= 65535 && C > 0 && T > 0 && (P >= 1 && P <= 3), "mask_rows: bad shape");
    vnb::launch_k(mask_rows_kernel, dim3(ceil_div(T, MASK_ROWS_PER_CTA), B), 256, 0, ST, x, (bf16*)planes_a,
                  (bf16*)planes_b, (bf16*)planes_c, B, C, T, round_up8(C), P, lengths);
    VNB_LAUNCH_CHECK("mask_rows");
    return VNB_OK;
}

extern "C" int vnb_gn_sums_unpad(const float* h, int B, int C, int T, int groups, const int* lengths,
                                 double* gn_sums, void* stream) {
    VNB_REQUIRE(h && lengths && gn_sums, "gn_sums_unpad: null pointer");
    VNB_REQUIRE(B > 0 && B <= 65535 && C > 0 && T > 0 && groups > 0 && groups <= UNPAD_MAX_GROUPS && C % groups == 0,
                "gn_sums_unpad: bad shape");
    vnb::launch_k(gn_sums_unpad_kernel, dim3(ceil_div(T, MASK_ROWS_PER_CTA), B), 256, 0, ST, h, C, T, groups, lengths,
                  gn_sums);
    VNB_LAUNCH_CHECK("gn_sums_unpad");
    return VNB_OK;
}

extern "C" int vnb_gn_finalize_masked(const double* gn_sums, int B, int groups, int channels_per_group, int T,
                                      const int* lengths, float eps, float* mean, float* rstd, void* stream) {
    VNB_REQUIRE(gn_sums && lengths && mean && rstd && B > 0 && groups > 0 && channels_per_group > 0 && T > 0,
                "gn_finalize_masked: bad arguments");
    vnb::launch_k(gn_finalize_masked_kernel, ceil_div((int64_t)B * groups, 128), 128, 0, ST, gn_sums, B, groups,
                  channels_per_group, T, lengths, eps, mean, rstd);
    VNB_LAUNCH_CHECK("gn_finalize_masked");
    return VNB_OK;
}

extern "C" int vnb_csid(const int64_t* idx, int B, int T, const int* lengths, int64_t* out, int* out_len,
                        void* stream) {
    VNB_REQUIRE(idx && out && out_len && B > 0 && T > 0, "csid: bad arguments");
    vnb::launch_k(csid_kernel, B, 256, 0, ST, idx, T, lengths, out, out_len);
    VNB_LAUNCH_CHECK("csid");
    return VNB_OK;
}
