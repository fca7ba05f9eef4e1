// Fused VQ lookup (round 2; included by gemm_tc.cu): frame preparation + distance GEMM + argmin + gather + loss sums +
// histogram in ONE kernel.  Reference: vae_npvc/model/layers_vq.py:85-150 (and :285-308, :35-58).
//
// Round 1 ran vq_prep_z (write z_n and three bf16 planes), vq_lookup_tc (SIX hi/mid/lo plane-pair products, so that
// the fp32 argmin is reproduced on tensor cores) and vq_gather (re-read z_n, write q): 0.15 of the HBM roof at
// K 512 x D 64, tensor pipe 36 % busy with 6x the algorithmic FLOPs.  Here:
//   * z is read ONCE: a CTA loads its 128 frames, normalises them, writes z_n (fp32, for backward) and puts the bf16
//     hi and mid planes straight into the swizzled MMA operand tiles (generic-proxy stores + fence.proxy.async);
//   * THREE bf16 products per code (z_mid.e_hi + z_hi.e_mid + z_hi.e_hi, the two-plane split: 16 significant bits per
//     operand) rank the codes - half the tensor work of the six-product kernel; eight epilogue warps (lane = frame, two
//     warps per TMEM lane quarter, one column half each) reduce every block of 32 scores  v = dot - |e|^2 / 2  to its
//     two best with a min/max TREE (no serial compare chain; the code's index rides in the five lowest mantissa bits)
//     and keep the three best of the whole codebook in registers;
//   * where the runner-up is outside a RIGOROUS error window of the best score, the best IS the fp32 argmin (no exact
//     evaluation needed); where it is inside, the two are re-ranked exactly in fp32 on the CUDA cores ((|z|^2 + |e|^2)
//     - 2 z.e as the reference computes it), ties to the lowest index; where the third is inside too, the frame is
//     searched exhaustively in fp32 - so the indices are the fp32 argmin wherever its top-2 gap exceeds rounding,
//     like the six-product kernel;
//   * the winning rows are gathered and written, the loss sums and the histogram accumulated.
// Error window: hi + mid leaves |x - x_hi - x_mid| <= 2^-18 |x| per element; the dropped product z_mid.e_mid and the two
// residuals are each <= 2^-18 |z||e| (Cauchy-Schwarz), the fp32 accumulation of D <= 128 terms <= 2^-17 |z||e|:
// |dot_computed - dot| <= 2^-16 |z||e|, two scores can be mis-ordered by at most 2^-15 |z| E_max; w = 2^-14 |z| E_max.
// (ONE bf16 product was measured first: its rigorous window, 2^-6 |z| E_max, holds 4+ candidates for a tenth of the
// frames of un-normalised data and every such frame costs an exhaustive fp32 search - 25-100x slower than the six-product
// kernel, profiles/r02h_vq_fused_single_product.txt.  Only a statistical window would make it fast; bit-exactness is the
// contract here.)
#pragma once
#include "tc_common.cuh"

namespace vnb {

using namespace tc;

constexpr int VQF_BM = 128, VQF_BN = 128, VQF_BK = 64;   // 128-code chunks: 2 x 128 TMEM columns, two CTAs per SM
constexpr int VQF_EPI_WARPS = 8, VQF_THREADS = 64 + 32 * VQF_EPI_WARPS;   // TMA warp, MMA warp, 8 epilogue warps
constexpr uint32_t VQF_A_BYTES = VQF_BM * VQF_BK * 2;       // 16 KB: 128 frames x 64 dims bf16
constexpr uint32_t VQF_B_BYTES = VQF_BN * VQF_BK * 2;       // 16 KB: 128 codes x 64 dims bf16 (ONE plane = one stage)
constexpr float VQF_NEG = -3.0e38f;

struct VqFusedArgs {
    int64_t N;
    int D, K, normalize, nstages, dbg;
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

// Running top-3 of the scores a lane (= frame) has seen.  A score carries the index of its code inside its block of
// 32 in its five lowest mantissa bits (a relative perturbation <= 2^-18, inside the error window), so only the block
// number travels beside it - branch-free: five min/max and five compare/select per insert.
struct Top3 {
    float r0, r1, r2;
    int g0, g1;
};
__device__ __forceinline__ void top3_insert(Top3& t, float x, int id) {
    const bool c0 = x > t.r0, c1 = x > t.r1;
    t.r2 = fmaxf(t.r2, fminf(t.r1, x));
    t.g1 = c0 ? t.g0 : (c1 ? id : t.g1);
    t.r1 = fmaxf(t.r1, fminf(t.r0, x));
    t.g0 = c0 ? id : t.g0;
    t.r0 = fmaxf(t.r0, x);
}
__device__ __forceinline__ float tag_score(float s, int i) { return __int_as_float((__float_as_int(s) & ~31) | i); }
__device__ __forceinline__ int tagged_code(float r, int g) { return (g << 5) | (__float_as_int(r) & 31); }

__global__ void __launch_bounds__(VQF_THREADS, 2)
vq_fused_kernel(const __grid_constant__ CUtensorMap map_e, const VqFusedArgs a) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t full_bar[8];
    __shared__ __align__(8) uint64_t empty_bar[8];
    __shared__ __align__(8) uint64_t tmem_full_bar[2];
    __shared__ __align__(8) uint64_t tmem_empty_bar[2];
    __shared__ uint32_t tmem_slot;
    __shared__ float s_zsq[VQF_BM], s_nrm[VQF_BM];
    __shared__ __align__(16) float s_es[2][VQF_BN];          // |e|^2 / 2 of the chunk in flight (double-buffered)
    __shared__ float s_r[3][VQF_BM];                         // the second column half's top-3, handed to the first
    __shared__ int s_g[2][VQF_BM];
    __shared__ int s_best[VQF_BM];
    __shared__ float s_hmax[VQF_EPI_WARPS];

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
    const int64_t n0 = (int64_t)blockIdx.x * VQF_BM;
    const int D = a.D, K = a.K, STAGES = a.nstages;
    const int kd = (D + VQF_BK - 1) / VQF_BK;
    const int chunks = (K + VQF_BN - 1) / VQF_BN;
    uint8_t* smem_a = smem;                                              // [plane hi | mid][kd] tiles of 16 KB (SWIZZLE_128B)
    uint8_t* smem_b = smem + (size_t)2 * kd * VQF_A_BYTES;               // ring of 32 KB stages: one codebook plane tile each
    const float* znp = a.normalize ? a.zn_out : a.z;                     // fp32 frames as the exact re-rank / gather read them
    long long* ts = ((a.dbg == 9 && blockIdx.x == 0) || (a.dbg == 10 && blockIdx.x == gridDim.x - 1)) ? g_tc_ts[1] : nullptr;     // VNB_TC_DBG=9 / 10: clock64 timeline of the first / last CTA

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&map_e);
        for (int s = 0; s < STAGES; s++) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; s++) {
            mbar_init(&tmem_full_bar[s], 1);
            mbar_init(&tmem_empty_bar[s], VQF_EPI_WARPS);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<2 * VQF_BN>(&tmem_slot);
    pdl_wait();
    if (threadIdx.x == 0) TS(0);

    // ===== prologue (all warps): frames -> z_n (fp32, global), |z_n|^2, |z|, bf16 hi / mid operand tiles =====
    {
        const int L = D >> 2;                              // float4 per frame: 4, 8, 16 or 32 lanes own one frame
        const int rows_per_pass = VQF_THREADS / L;
        const int l = threadIdx.x % L, rsub = threadIdx.x / L;
        if (kd * 64 > D) {                                 // D < 64: the unused K columns of the operand tiles must be zero
            for (int i = threadIdx.x; i < (int)(2 * kd * VQF_A_BYTES / 16); i += VQF_THREADS)
                reinterpret_cast<uint4*>(smem_a)[i] = make_uint4(0, 0, 0, 0);
            __syncthreads();
        }
        const int passes = (VQF_BM + rows_per_pass - 1) / rows_per_pass;
        for (int p0 = 0; p0 < passes; p0 += 8) {           // eight row loads in flight per thread before the first is used
            float4 vb[8];
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const int r = rsub + (p0 + u) * rows_per_pass;
                vb[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (p0 + u < passes && r < VQF_BM && n0 + r < a.N) vb[u] = __ldg(reinterpret_cast<const float4*>(a.z + (n0 + r) * D) + l);
            }
#pragma unroll
            for (int u = 0; u < 8; u++) {
                if (p0 + u >= passes) break;               // uniform
                const int r = rsub + (p0 + u) * rows_per_pass;
                const int64_t n = n0 + r;
                const bool ok = r < VQF_BM && n < a.N;
                float4 v = vb[u];
                float ss = (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
                for (int o = L >> 1; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
                const float nrm = sqrtf(ss);
                if (a.normalize && ok) {
                    v.x = v.x / nrm; v.y = v.y / nrm; v.z = v.z / nrm; v.w = v.w / nrm;
                }
                float qq = (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
                for (int o = L >> 1; o > 0; o >>= 1) qq += __shfl_xor_sync(0xffffffffu, qq, o);
                if (r >= VQF_BM) continue;
                if (ok && a.zn_out) *(reinterpret_cast<float4*>(a.zn_out + n * D) + l) = v;
                if (l == 0) {
                    s_zsq[r] = qq;
                    s_nrm[r] = nrm;
                    if (ok && a.z_norm_out) a.z_norm_out[n] = nrm;
                }
                // bf16 planes into the K-major SWIZZLE_128B tiles: row r = 128 B, 16-byte chunk index XOR (r & 7)
                const int d = 4 * l, tile = d >> 6, dc = d & 63;
                const uint32_t off = (uint32_t)tile * VQF_A_BYTES + (uint32_t)r * 128u + (uint32_t)(((dc >> 3) ^ (r & 7)) << 4) +
                                     (uint32_t)(dc & 7) * 2u;
                const float vv[4] = {v.x, v.y, v.z, v.w};
                bf16 hi[4], mid[4], lo_;
#pragma unroll
                for (int j = 0; j < 4; j++) split3(vv[j], hi[j], mid[j], lo_);
                uint2 pk;
                pk.x = (uint32_t)__bfloat16_as_ushort(hi[0]) | ((uint32_t)__bfloat16_as_ushort(hi[1]) << 16);
                pk.y = (uint32_t)__bfloat16_as_ushort(hi[2]) | ((uint32_t)__bfloat16_as_ushort(hi[3]) << 16);
                *reinterpret_cast<uint2*>(smem_a + off) = pk;
                pk.x = (uint32_t)__bfloat16_as_ushort(mid[0]) | ((uint32_t)__bfloat16_as_ushort(mid[1]) << 16);
                pk.y = (uint32_t)__bfloat16_as_ushort(mid[2]) | ((uint32_t)__bfloat16_as_ushort(mid[3]) << 16);
                *reinterpret_cast<uint2*>(smem_a + (size_t)kd * VQF_A_BYTES + off) = pk;
            }
        }
    }
    fence_proxy_async_smem();                              // the operand tiles were written through the generic proxy
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;
    if (threadIdx.x == 0) TS(1);

    if (warp == 0) {
        // ===== producer: per chunk of 128 codes and 64-dim slice: the mid plane tile, then the hi plane tile =====
        int s = 0, i = 0;
        uint32_t ph = 1;
        for (int ch = 0; ch < chunks; ch++)
            for (int c = 0; c < kd; c++)
                for (int pl = 1; pl >= 0; pl--, i++) {
                    if (i >= STAGES) mbar_wait(&empty_bar[s], ph);
                    if (elect_one()) {
                        mbar_expect_tx(&full_bar[s], VQF_B_BYTES);
                        tma_load_3d(smem_b + (size_t)s * VQF_B_BYTES, &map_e, &full_bar[s], c * VQF_BK, ch * VQF_BN, pl);
                    }
                    __syncwarp();
                    if (++s == STAGES) { s = 0; ph ^= 1; }
                }
    } else if (warp == 1) {
        // ===== MMA issuer: z_hi.e_mid on the mid tile, then z_mid.e_hi + z_hi.e_hi on the hi tile =====
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
            if (ch < 4 && lane == 0) TS(2 + 4 * ch);
            for (int c = 0; c < kd; c++) {
                const int ksteps = (D - c * VQF_BK >= VQF_BK ? VQF_BK : D - c * VQF_BK) / 16;
                const uint64_t da_hi = smem_desc_sw128(sa0 + (uint32_t)c * VQF_A_BYTES, 16, 1024);
                const uint64_t da_mid = smem_desc_sw128(sa0 + (uint32_t)(kd + c) * VQF_A_BYTES, 16, 1024);
                for (int step = 0; step < 2; step++) {
                    mbar_wait(&full_bar[s], ph);
                    tc_fence_after();
                    if (ch < 4 && lane == 0 && c == 0) TS(3 + 4 * ch + step);
                    const uint64_t db = smem_desc_sw128(sb0 + (uint32_t)s * VQF_B_BYTES, 16, 1024);
                    if (elect_one()) {
                        if (step == 0) {
                            for (int j = 0; j < ksteps; j++) umma_bf16(tacc, da_hi + 2 * j, db + 2 * j, idesc, (c | j) != 0);
                        } else {
                            for (int j = 0; j < ksteps; j++) umma_bf16(tacc, da_mid + 2 * j, db + 2 * j, idesc, 1);
                            for (int j = 0; j < ksteps; j++) umma_bf16(tacc, da_hi + 2 * j, db + 2 * j, idesc, 1);
                        }
                        umma_commit(&empty_bar[s]);
                    }
                    __syncwarp();
                    if (++s == STAGES) { s = 0; ph ^= 1; }
                }
            }
            if (elect_one()) umma_commit(&tmem_full_bar[buf]);
            __syncwarp();
            if (ch < 4 && lane == 0) TS(5 + 4 * ch);
        }
    } else {
        // ===== epilogue: 8 warps; lane = frame, each warp scans one column half of every chunk =====
        const int ew = warp - 2;
        const int q = warp & 3;                            // the TMEM lane quarter this warp may read
        const int half = ew >> 2;
        const int r = q * 32 + lane;
        const int64_t n = n0 + r;
        Top3 t{VQF_NEG, VQF_NEG, VQF_NEG, 0, 0};
        float hmax = 0.f;
        const bool es_loader = ew < VQF_BN / 32;           // these warps stage |e|^2 / 2 of every chunk
        float e_next = 3.0e38f;
        if (es_loader) {
            const int k = ew * 32 + lane;
            e_next = k < K ? 0.5f * __ldg(a.e_sq + k) : 3.0e38f;
        }
        for (int ch = 0; ch < chunks; ch++) {
            const int buf = ch & 1;
            if (es_loader) s_es[buf][ew * 32 + lane] = e_next;   // codes >= K: score -3e38, never a candidate (K >= 8)
            if (e_next < 1.0e38f) hmax = fmaxf(hmax, e_next);
            if (es_loader && ch + 1 < chunks) {
                const int k = (ch + 1) * VQF_BN + ew * 32 + lane;
                e_next = k < K ? 0.5f * __ldg(a.e_sq + k) : 3.0e38f;
            }
            named_bar_sync(1, 32 * VQF_EPI_WARPS);
            if (ch < 4 && ew == 0 && lane == 0) TS(20 + 4 * ch);
            mbar_wait_warp(&tmem_full_bar[buf], (ch >> 1) & 1, 20);
            tc_fence_after();
            if (ch < 4 && ew == 0 && lane == 0) TS(21 + 4 * ch);
#pragma unroll 1
            for (int b = 0; b < VQF_BN / 64; b++) {
                const int c0 = half * (VQF_BN / 2) + b * 32;
                const int k0 = ch * VQF_BN + c0;
                if (k0 >= K) break;                                  // warp-uniform
                float v[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * VQF_BN + c0), v);
                const float4* es4 = reinterpret_cast<const float4*>(&s_es[buf][c0]);
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    const float4 e = es4[i];                         // broadcast read
                    v[4 * i + 0] -= e.x;
                    v[4 * i + 1] -= e.y;
                    v[4 * i + 2] -= e.z;
                    v[4 * i + 3] -= e.w;
                }
#pragma unroll
                for (int i = 0; i < 32; i++) v[i] = tag_score(v[i], i);
                // the block's two best tagged scores (tree: independent min/max, no serial chain)
                float hi[16], lo[16];
#pragma unroll
                for (int j = 0; j < 16; j++) {
                    hi[j] = fmaxf(v[2 * j], v[2 * j + 1]);
                    lo[j] = fminf(v[2 * j], v[2 * j + 1]);
                }
#pragma unroll
                for (int m = 8; m >= 1; m >>= 1) {
#pragma unroll
                    for (int j = 0; j < m; j++) {
                        const float h = fmaxf(hi[j], hi[j + m]);
                        const float l2 = fmaxf(fminf(hi[j], hi[j + m]), fmaxf(lo[j], lo[j + m]));
                        hi[j] = h;
                        lo[j] = l2;
                    }
                }
                const int id = k0 >> 5;
                top3_insert(t, hi[0], id);
                top3_insert(t, lo[0], id);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&tmem_empty_bar[buf])) : "memory");
            if (ch < 4 && ew == 0 && lane == 0) TS(22 + 4 * ch);
        }
        if (ew == 0 && lane == 0) TS(40);
        // ----- hand the second half's top-3 to the first, largest code norm -----
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) hmax = fmaxf(hmax, __shfl_xor_sync(0xffffffffu, hmax, o));
        if (lane == 0) s_hmax[ew] = hmax;
        if (half == 1) {
            s_r[0][r] = t.r0; s_r[1][r] = t.r1; s_r[2][r] = t.r2;
            s_g[0][r] = t.g0; s_g[1][r] = t.g1;
        }
        named_bar_sync(1, 32 * VQF_EPI_WARPS);
        if (half == 0) {
            top3_insert(t, s_r[0][r], s_g[0][r]);
            top3_insert(t, s_r[1][r], s_g[1][r]);
            t.r2 = fmaxf(t.r2, s_r[2][r]);
            float hm = s_hmax[0];
#pragma unroll
            for (int j = 1; j < VQF_EPI_WARPS; j++) hm = fmaxf(hm, s_hmax[j]);
            const float emax = sqrtf(2.f * hm);                          // largest code norm
            const float zsq = s_zsq[r];
            // 2^-14 |z| E_max (three plane products, see the header) + 2^-16 E_max^2 (the index tag) + fp32 slack
            const float w = 6.103515625e-5f * sqrtf(zsq) * emax + 1.52587890625e-5f * emax * emax + 1e-6f * (zsq + emax * emax);
            const int ka = tagged_code(t.r0, t.g0), kb = tagged_code(t.r1, t.g1);
            const bool two = t.r1 >= t.r0 - w;
            const bool overflow = t.r2 >= t.r0 - w;                      // a third candidate inside the window
            int best_k = ka;
            // ----- exact fp32 re-rank where a second code is inside the window (warp-cooperative, frame by frame) -----
            unsigned need = __ballot_sync(0xffffffffu, n < a.N && two);
            const int per = (D + 31) / 32;
            while (need) {
                const int f = __ffs(need) - 1;
                need &= need - 1;
                const bool ov = __shfl_sync(0xffffffffu, (int)overflow, f) != 0;
                const int c0k = __shfl_sync(0xffffffffu, ka, f), c1k = __shfl_sync(0xffffffffu, kb, f);
                const float zs = __shfl_sync(0xffffffffu, zsq, f);
                const float* zr = znp + (size_t)(n0 + q * 32 + f) * D;
                float zreg[4];
#pragma unroll
                for (int j = 0; j < 4; j++) zreg[j] = (j < per && lane + 32 * j < D) ? zr[lane + 32 * j] : 0.f;
                float bd = INFINITY;
                int bk = 0x7fffffff;
                if (ov) {
                    // a third code inside the window (~1e-5 of the frames): exact search of the whole codebook, the lanes
                    // splitting the codes (a lane's codes ascend, so its ties keep the lowest index)
                    const float4* z4 = reinterpret_cast<const float4*>(zr);
                    for (int k = lane; k < K; k += 32) {
                        const float4* er4 = reinterpret_cast<const float4*>(a.e_hat + (size_t)k * D);
                        float d0 = 0.f, d1 = 0.f, d2 = 0.f, d3 = 0.f;
                        for (int j = 0; j < (D >> 2); j++) {
                            const float4 e = __ldg(er4 + j), zz = z4[j];
                            d0 = fmaf(zz.x, e.x, d0); d1 = fmaf(zz.y, e.y, d1); d2 = fmaf(zz.z, e.z, d2); d3 = fmaf(zz.w, e.w, d3);
                        }
                        const float dist = (zs + __ldg(a.e_sq + k)) - 2.f * ((d0 + d1) + (d2 + d3));
                        if (dist < bd) {
                            bd = dist;
                            bk = k;
                        }
                    }
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) {
                        const float od = __shfl_xor_sync(0xffffffffu, bd, o);
                        const int ok2 = __shfl_xor_sync(0xffffffffu, bk, o);
                        if (od < bd || (od == bd && ok2 < bk)) {
                            bd = od;
                            bk = ok2;
                        }
                    }
                }
                for (int c = 0; c < (ov ? 0 : 2); c++) {
                    const int k = c == 0 ? c0k : c1k;
                    const float* er = a.e_hat + (size_t)k * D;
                    float dot = 0.f;
#pragma unroll
                    for (int j = 0; j < 4; j++)
                        if (j < per && lane + 32 * j < D) dot = fmaf(zreg[j], __ldg(er + lane + 32 * j), dot);
                    dot = warp_sum(dot);
                    const float dist = (zs + __ldg(a.e_sq + k)) - 2.f * dot;   // (z_sq + e_sq) - 2 z.e  (layers_vq.py:103-105)
                    if (dist < bd || (dist == bd && k < bk)) {
                        bd = dist;
                        bk = k;
                    }
                }
                if (lane == f) best_k = bk;
            }
            s_best[r] = best_k;
            if (n < a.N) {
                a.idx[n] = (int64_t)best_k;
                if (a.counts) atomicAdd(&a.counts[best_k], 1);
            }
            if (a.sums && a.normalize) {
                // sum (z_n - z)^2 over a frame = (1 - |z|)^2 exactly (z_n = z / |z|): the normalised quantiser's third term
                float ln = n < a.N ? (1.f - s_nrm[r]) * (1.f - s_nrm[r]) : 0.f;
                ln = warp_sum(ln);
                if (lane == 0) atomicAdd(&a.sums[1], ln);
            }
        }
        named_bar_sync(1, 32 * VQF_EPI_WARPS);
        if (ew == 0 && lane == 0) TS(41);
        // ----- gather the winning rows, loss sum: 16 frames per warp, float4 lanes, loads of four passes in flight -----
        {
            const int L = D >> 2, fpp = 32 / L, sub = lane / L, l = lane % L;
            const int passes = 16 / (fpp > 16 ? 16 : fpp);
            float lq = 0.f;
            for (int p0 = 0; p0 < passes; p0 += 4) {
                float4 e[4], z[4];
                int64_t nf[4];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int fr = ew * 16 + (p0 + u) * fpp + sub;
                    nf[u] = (p0 + u < passes && sub < 16 && fr < ew * 16 + 16) ? n0 + fr : a.N;
                    if (nf[u] < a.N) {
                        const int k = s_best[fr];
                        e[u] = __ldg(reinterpret_cast<const float4*>(a.e_hat + (size_t)k * D) + l);
                        z[u] = *(reinterpret_cast<const float4*>(znp + nf[u] * D) + l);
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    if (nf[u] < a.N) {
                        if (a.q_out) *(reinterpret_cast<float4*>(a.q_out + nf[u] * D) + l) = e[u];
                        const float dx = e[u].x - z[u].x, dy = e[u].y - z[u].y, dz = e[u].z - z[u].z, dw = e[u].w - z[u].w;
                        lq += (dx * dx + dy * dy) + (dz * dz + dw * dw);
                    }
                }
            }
            if (a.sums) {
                lq = warp_sum(lq);
                if (lane == 0) atomicAdd(&a.sums[0], lq);
            }
        }
        if (ew == 0 && lane == 0) TS(43);
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) TS(42);
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
    // K > 8192: the six-product kernel already runs at the tensor roof there and the three-kernel path is as fast
    // (profiles/r02h_vq_sweep.txt: 0.90x at K 16384 x D 128)
    if (!(D == 16 || D == 32 || D == 64 || D == 128) || K < 8 || K > 8192 || N > (int64_t)1 << 30) return 1;
    if ((((uintptr_t)z) | ((uintptr_t)zn_out) | ((uintptr_t)q_out)) & 15) return 1;
    if (normalize && !zn_out) return 1;                    // the exact re-rank reads the normalised frames back (L2)
    const int kd = ceil_div(D, VQF_BK);
    const size_t fixed = (size_t)2 * kd * VQF_A_BYTES;
    const size_t budget = 104 * 1024;                      // + 7 KB static < 113 KB: two CTAs per SM overlap each other's phases
    int nst = (int)((budget - fixed) / VQF_B_BYTES);
    if (nst < 2) return 1;
    nst = nst > 4 ? 4 : nst;
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
    // hi and mid are planes 0 and 1 of the (3, K, Dp) codebook planes
    if (!make_tmap_3d(&me, e_planes, (uint64_t)D, (uint64_t)K, 2, (uint64_t)Dp * 2, (uint64_t)K * Dp * 2, VQF_BK, VQF_BN, 1))
        return VNB_E_CUDA;
    VqFusedArgs a{N, D, K, normalize, nst, option(VNB_OPT_TC_DBG), z, e_hat, e_sq, zn_out, z_norm_out, idx, q_out, sums, counts};
    vnb::launch_k(vq_fused_kernel, ceil_div(N, VQF_BM), VQF_THREADS, smem, st, me, a);
    VNB_LAUNCH_CHECK("vq_fused");
    return VNB_OK;
}

}  // namespace vnb
