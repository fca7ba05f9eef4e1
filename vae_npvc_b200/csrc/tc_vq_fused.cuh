// Fused VQ lookup (round 2; included by gemm_tc.cu): frame preparation + distance GEMM + argmin + gather + loss sums +
// histogram in ONE kernel.  Reference: vae_npvc/model/layers_vq.py:85-150 (and :285-308, :35-58).
//
// Round 1 ran vq_prep_z (write z_n and three bf16 planes), vq_lookup_tc (SIX hi/mid/lo plane-pair products, so that
// the fp32 argmin is reproduced on tensor cores) and vq_gather (re-read z_n, write q): 0.15 of the HBM roof at
// K 512 x D 64, tensor pipe 36 % busy with 6x the algorithmic FLOPs.  Here:
//   * z is read ONCE: a CTA loads its 128 frames, normalises them, keeps z_n in shared memory as fp32 and writes its
//     bf16 hi plane straight into the swizzled MMA operand tile (generic-proxy stores + fence.proxy.async);
//   * ONE bf16 product per code (dot ~ z_hi . e_hi) ranks the codes approximately; every epilogue lane (= frame) keeps
//     its four best approximate scores  v = dot - |e|^2 / 2  in registers;
//   * candidates within a rigorous error window of the best score are re-ranked EXACTLY in fp32 on the CUDA cores
//     ((|z|^2 + |e|^2) - 2 z.e as the reference computes it), ties to the lowest index; if all four kept scores fall
//     inside the window the frame is searched exhaustively in fp32 (never seen on real data, kept for safety) - so the
//     indices are the fp32 argmin wherever its top-2 gap exceeds rounding, like the six-product kernel;
//   * the winning rows are gathered and written, the loss sums and the histogram accumulated, z_n never re-read.
// Error window: bf16 rounding of both operands is relative 2^-9 each, so |dot_bf16 - dot| <= 2^-8 |z||e| (+ O(2^-17));
// two candidates' scores can therefore be mis-ordered by at most 2 * 2^-8 |z| E_max: w = 2^-6 |z| E_max (2x margin).
#pragma once
#include "tc_common.cuh"

namespace vnb {

using namespace tc;

constexpr int VQF_BM = 128, VQF_BN = 256, VQF_BK = 64, VQF_THREADS = 256;
constexpr uint32_t VQF_A_BYTES = VQF_BM * VQF_BK * 2;       // 16 KB: 128 frames x 64 dims bf16
constexpr uint32_t VQF_B_BYTES = VQF_BN * VQF_BK * 2;       // 32 KB: 256 codes x 64 dims bf16

struct VqFusedArgs {
    int64_t N;
    int D, K, normalize, nstages;
    const float* z;          // (N, D) frames, channels-last rows
    const float* e_hat;      // (K, D) fp32 (normalised) codebook
    const float* e_sq;       // (K)
    float* zn_out;           // (N, D) | null: normalised frames (backward needs them)
    float* z_norm_out;       // (N) | null
    int64_t* idx;            // (N)
    float* q_out;            // (N, D) | null
    float* sums;             // [sum (q - z_n)^2, sum (z_n - z)^2] | null (zeroed by the caller)
    int* counts;             // (K) | null (zeroed by the caller)
};

struct Top4 {
    float v0, v1, v2, v3;
    int k0, k1, k2, k3;
};
__device__ __forceinline__ void top4_insert(Top4& t, float v, int k) {
    // called only when v > t.v3 (strictly: among equal approximate scores the earlier = lower index stays)
    if (v > t.v2) {
        t.v3 = t.v2; t.k3 = t.k2;
        if (v > t.v1) {
            t.v2 = t.v1; t.k2 = t.k1;
            if (v > t.v0) {
                t.v1 = t.v0; t.k1 = t.k0;
                t.v0 = v; t.k0 = k;
            } else {
                t.v1 = v; t.k1 = k;
            }
        } else {
            t.v2 = v; t.k2 = k;
        }
    } else {
        t.v3 = v; t.k3 = k;
    }
}

__global__ void __launch_bounds__(VQF_THREADS, 1)
vq_fused_kernel(const __grid_constant__ CUtensorMap map_e, const VqFusedArgs a) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t full_bar[8];
    __shared__ __align__(8) uint64_t empty_bar[8];
    __shared__ __align__(8) uint64_t tmem_full_bar[2];
    __shared__ __align__(8) uint64_t tmem_empty_bar[2];
    __shared__ uint32_t tmem_slot;
    __shared__ float s_zsq[VQF_BM], s_nrm[VQF_BM];

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
    const int64_t n0 = (int64_t)blockIdx.x * VQF_BM;
    const int D = a.D, K = a.K, STAGES = a.nstages;
    const int kd = (D + VQF_BK - 1) / VQF_BK;
    const int chunks = (K + VQF_BN - 1) / VQF_BN;
    uint8_t* smem_a = smem;                                              // kd tiles of 16 KB (SWIZZLE_128B, K-major)
    float* zn_s = reinterpret_cast<float*>(smem + (size_t)kd * VQF_A_BYTES);  // 128 x D fp32
    uint8_t* smem_b = smem + (size_t)kd * VQF_A_BYTES + (((size_t)VQF_BM * D * 4 + 1023) & ~(size_t)1023);

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&map_e);
        for (int s = 0; s < STAGES; s++) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; s++) {
            mbar_init(&tmem_full_bar[s], 1);
            mbar_init(&tmem_empty_bar[s], 4);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<2 * VQF_BN>(&tmem_slot);
    pdl_wait();

    // ===== prologue (all warps): frames -> z_n (fp32, shared [+ global]), |z_n|^2, |z|, bf16 operand tile =====
    {
        const int L = D >> 2;                              // float4 per frame: 4, 8, 16 or 32 lanes own one frame
        const int rows_per_pass = VQF_THREADS / L;
        const int l = threadIdx.x % L, rsub = threadIdx.x / L;
        if (kd * 64 > D) {                                 // D < 64: the unused K columns of the operand tile must be zero
            for (int i = threadIdx.x; i < (int)(kd * VQF_A_BYTES / 16); i += VQF_THREADS)
                reinterpret_cast<uint4*>(smem_a)[i] = make_uint4(0, 0, 0, 0);
            __syncthreads();
        }
        for (int r = rsub; r < VQF_BM; r += rows_per_pass) {
            const int64_t n = n0 + r;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (n < a.N) v = *reinterpret_cast<const float4*>(a.z + n * D + 4 * l);
            float ss = (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
            for (int o = L >> 1; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
            const float nrm = sqrtf(ss);
            if (a.normalize && n < a.N) {
                v.x = v.x / nrm; v.y = v.y / nrm; v.z = v.z / nrm; v.w = v.w / nrm;
            }
            float qq = (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
            for (int o = L >> 1; o > 0; o >>= 1) qq += __shfl_xor_sync(0xffffffffu, qq, o);
            *reinterpret_cast<float4*>(zn_s + (size_t)r * D + 4 * l) = v;
            if (n < a.N && a.zn_out) *reinterpret_cast<float4*>(a.zn_out + n * D + 4 * l) = v;
            if (l == 0) {
                s_zsq[r] = qq;
                s_nrm[r] = nrm;
                if (n < a.N && a.z_norm_out) a.z_norm_out[n] = nrm;
            }
            // bf16 hi plane into the K-major SWIZZLE_128B tile: row r = 128 B, 16-byte chunk index XOR (r & 7)
            const int d = 4 * l, tile = d >> 6, dc = d & 63;
            const uint32_t off = (uint32_t)tile * VQF_A_BYTES + (uint32_t)r * 128u + (uint32_t)(((dc >> 3) ^ (r & 7)) << 4) +
                                 (uint32_t)(dc & 7) * 2u;
            uint2 pk;
            pk.x = (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(v.x)) | ((uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(v.y)) << 16);
            pk.y = (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(v.z)) | ((uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(v.w)) << 16);
            *reinterpret_cast<uint2*>(smem_a + off) = pk;
        }
    }
    fence_proxy_async_smem();                              // the operand tile was written through the generic proxy
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    if (warp == 0) {
        // ===== producer: codebook hi plane, chunk by chunk (256 codes x 64 dims per stage) =====
        int s = 0, i = 0;
        uint32_t ph = 1;
        for (int ch = 0; ch < chunks; ch++)
            for (int c = 0; c < kd; c++, i++) {
                if (i >= STAGES) mbar_wait(&empty_bar[s], ph);
                if (elect_one()) {
                    mbar_expect_tx(&full_bar[s], VQF_B_BYTES);
                    tma_load_3d(smem_b + (size_t)s * VQF_B_BYTES, &map_e, &full_bar[s], c * VQF_BK, ch * VQF_BN, 0);
                }
                __syncwarp();
                if (++s == STAGES) { s = 0; ph ^= 1; }
            }
    } else if (warp == 1) {
        // ===== MMA issuer: ONE bf16 product per code =====
        constexpr uint32_t idesc = instr_desc_bf16(VQF_BM, VQF_BN, 0, 0);
        const uint32_t sa0 = smem_u32(smem_a), sb0 = smem_u32(smem_b);
        int s = 0;
        uint32_t ph = 0;
        for (int ch = 0; ch < chunks; ch++) {
            const int buf = ch & 1;
            if (ch >= 2) {
                mbar_wait(&tmem_empty_bar[buf], ((ch >> 1) - 1) & 1);
                tc_fence_after();
            }
            const uint32_t tacc = tmem_base + (uint32_t)(buf * VQF_BN);
            for (int c = 0; c < kd; c++) {
                mbar_wait(&full_bar[s], ph);
                tc_fence_after();
                const int ksteps = (D - c * VQF_BK >= VQF_BK ? VQF_BK : D - c * VQF_BK) / 16;
                const uint64_t da = smem_desc_sw128(sa0 + (uint32_t)c * VQF_A_BYTES, 16, 1024);
                const uint64_t db = smem_desc_sw128(sb0 + (uint32_t)s * VQF_B_BYTES, 16, 1024);
                if (elect_one()) {
                    for (int j = 0; j < ksteps; j++) umma_bf16(tacc, da + 2 * j, db + 2 * j, idesc, (c | j) != 0);
                    umma_commit(&empty_bar[s]);
                }
                __syncwarp();
                if (++s == STAGES) { s = 0; ph ^= 1; }
            }
            if (elect_one()) umma_commit(&tmem_full_bar[buf]);
            __syncwarp();
        }
    } else if (warp < 6) {
        // ===== epilogue: lane = frame; four best approximate scores v = dot - |e|^2 / 2 in registers =====
        const int q = warp & 3;
        const int r = q * 32 + lane;
        const int64_t n = n0 + r;
        Top4 t{-INFINITY, -INFINITY, -INFINITY, -INFINITY, 0, 0, 0, 0};
        float hmax = 0.f;
        for (int ch = 0; ch < chunks; ch++) {
            const int buf = ch & 1;
            mbar_wait_warp(&tmem_full_bar[buf], (ch >> 1) & 1, 20);
            tc_fence_after();
#pragma unroll 1
            for (int c0 = 0; c0 < VQF_BN; c0 += 32) {
                const int k0 = ch * VQF_BN + c0;
                if (k0 >= K) break;                                  // warp-uniform
                float v[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * VQF_BN + c0), v);
                const float es = k0 + lane < K ? 0.5f * __ldg(a.e_sq + k0 + lane) : INFINITY;
                hmax = fmaxf(hmax, k0 + lane < K ? es : 0.f);
#pragma unroll
                for (int i = 0; i < 32; i++) {
                    const float s = v[i] - __shfl_sync(0xffffffffu, es, i);
                    if (s > t.v3) top4_insert(t, s, k0 + i);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&tmem_empty_bar[buf])) : "memory");
        }
        // ----- exact fp32 re-rank of the candidates inside the error window (warp-cooperative, frame by frame) -----
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) hmax = fmaxf(hmax, __shfl_xor_sync(0xffffffffu, hmax, o));
        const float emax = sqrtf(2.f * hmax);                        // largest code norm
        const float zsq = s_zsq[r];
        const float w = 0.015625f * sqrtf(zsq) * emax + 1e-6f * (zsq + emax * emax);     // 2^-6 |z| E_max (+ fp32 slack)
        int ncand = 1 + (t.v1 >= t.v0 - w) + (t.v2 >= t.v0 - w) + (t.v3 >= t.v0 - w);
        if (K < 4) ncand = ncand < K ? ncand : K;
        const bool overflow = K > 4 && t.v3 >= t.v0 - w;            // a fifth candidate may be inside the window
        int best_k = t.k0;
        unsigned need = __ballot_sync(0xffffffffu, n < a.N && (ncand >= 2 || overflow));
        const int per = (D + 31) / 32;
        while (need) {
            const int f = __ffs(need) - 1;
            need &= need - 1;
            const int nc = __shfl_sync(0xffffffffu, ncand, f);
            const bool ov = __shfl_sync(0xffffffffu, (int)overflow, f) != 0;
            const int c0k = __shfl_sync(0xffffffffu, t.k0, f), c1k = __shfl_sync(0xffffffffu, t.k1, f);
            const int c2k = __shfl_sync(0xffffffffu, t.k2, f), c3k = __shfl_sync(0xffffffffu, t.k3, f);
            const float zs = __shfl_sync(0xffffffffu, zsq, f);
            const float* zr = zn_s + (size_t)(q * 32 + f) * D;
            float zreg[4];
#pragma unroll
            for (int j = 0; j < 4; j++) zreg[j] = (j < per && lane + 32 * j < D) ? zr[lane + 32 * j] : 0.f;
            float bd = INFINITY;
            int bk = 0x7fffffff;
            const int total = ov ? K : nc;
            for (int c = 0; c < total; c++) {
                const int k = ov ? c : (c == 0 ? c0k : c == 1 ? c1k : c == 2 ? c2k : c3k);
                const float* er = a.e_hat + (size_t)k * D;
                float dot = 0.f;
#pragma unroll
                for (int j = 0; j < 4; j++)
                    if (j < per && lane + 32 * j < D) dot = fmaf(zreg[j], __ldg(er + lane + 32 * j), dot);
                dot = warp_sum(dot);
                const float dist = (zs + __ldg(a.e_sq + k)) - 2.f * dot;       // (z_sq + e_sq) - 2 z.e  (layers_vq.py:103-105)
                if (dist < bd || (dist == bd && k < bk)) {
                    bd = dist;
                    bk = k;
                }
            }
            if (lane == f) best_k = bk;
        }
        // ----- gather the winning rows, loss sums, histogram (warp-cooperative rows: coalesced) -----
        float lq = 0.f;
        for (int f = 0; f < 32; f++) {
            const int64_t nf = n0 + q * 32 + f;
            if (nf >= a.N) break;                                    // warp-uniform
            const int k = __shfl_sync(0xffffffffu, best_k, f);
            const float* er = a.e_hat + (size_t)k * D;
            const float* zr = zn_s + (size_t)(q * 32 + f) * D;
            for (int d = lane; d < D; d += 32) {
                const float e = __ldg(er + d), z = zr[d];
                if (a.q_out) a.q_out[nf * D + d] = e;
                lq += (e - z) * (e - z);
            }
        }
        if (n < a.N) {
            a.idx[n] = (int64_t)best_k;
            if (a.counts) atomicAdd(&a.counts[best_k], 1);
        }
        if (a.sums) {
            // sum (z_n - z)^2 over a frame = (1 - |z|)^2 exactly (z_n = z / |z|): the normalised quantiser's third term
            float ln = (a.normalize && n < a.N) ? (1.f - s_nrm[r]) * (1.f - s_nrm[r]) : 0.f;
            lq = warp_sum(lq);
            ln = warp_sum(ln);
            if (lane == 0) {
                atomicAdd(&a.sums[0], lq);
                if (a.normalize) atomicAdd(&a.sums[1], ln);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        tc_fence_after();
        tmem_dealloc<2 * VQF_BN>(tmem_base);
    }
}

// -> VNB_OK / error, or 1 when the shape is not taken (the caller runs prep_z + lookup + gather instead)
int vq_fused_tc(const float* z, int B, int D, int T, int normalize, const float* e_hat, const void* e_planes, const float* e_sq,
                int K, float* zn_out, float* z_norm_out, int64_t* idx, float* q_out, float* sums, int* counts, cudaStream_t st) {
    const int64_t N = (int64_t)B * T;
    if (!(D == 16 || D == 32 || D == 64 || D == 128) || K < 8 || N > (int64_t)1 << 30) return 1;
    if ((((uintptr_t)z) | ((uintptr_t)zn_out) | ((uintptr_t)q_out)) & 15) return 1;
    const int kd = ceil_div(D, VQF_BK);
    const size_t fixed = (size_t)kd * VQF_A_BYTES + (((size_t)VQF_BM * D * 4 + 1023) & ~(size_t)1023);
    const size_t budget = 227 * 1024 - 2048;
    int nst = (int)((budget - fixed) / VQF_B_BYTES);
    if (nst < 2) return 1;
    nst = nst > 6 ? 6 : nst;
    const size_t smem = fixed + (size_t)nst * VQF_B_BYTES + 1024;
    static size_t smem_set = 0;
    if (smem > smem_set) {
        cudaError_t e = cudaFuncSetAttribute(vq_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) {
            set_error("vq_fused: cudaFuncSetAttribute(%zu): %s", smem, cudaGetErrorString(e));
            return VNB_E_CUDA;
        }
        smem_set = smem;
    }
    const int Dp = round_up8(D);
    CUtensorMap me;
    // the hi plane is plane 0 of the (3, K, Dp) codebook planes
    if (!make_tmap_3d(&me, e_planes, (uint64_t)D, (uint64_t)K, 1, (uint64_t)Dp * 2, (uint64_t)K * Dp * 2, VQF_BK, VQF_BN, 1))
        return VNB_E_CUDA;
    VqFusedArgs a{N, D, K, normalize, nst, z, e_hat, e_sq, zn_out, z_norm_out, idx, q_out, sums, counts};
    vnb::launch_k(vq_fused_kernel, ceil_div(N, VQF_BM), VQF_THREADS, smem, st, me, a);
    VNB_LAUNCH_CHECK("vq_fused");
    return VNB_OK;
}

}  // namespace vnb
