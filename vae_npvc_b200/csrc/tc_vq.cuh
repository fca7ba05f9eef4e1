// tcgen05 VQ lookup: distance GEMM with a fused  ||z||^2 + ||e||^2 - 2 z.e  and running-argmin epilogue
// (included by gemm_tc.cu).  Reference: vae_npvc/model/layers_vq.py:103-108 (and :285-290, :46-52).
//
//     dot[n, k] = sum_{pair} sum_d Z[pa][n][d] * E[pb][k][d]        (hi/mid/lo planes: fp32-faithful)
//     idx[n]    = argmin_k ((z_sq[n] + e_sq[k]) - 2 dot[n, k]),  ties -> lowest k
//
// One CTA owns 128 frames (UMMA M, TMEM lanes) and walks the codebook in chunks of 128 codes.  The
// accumulator is double-buffered in TMEM (2 x 128 columns): while the epilogue warps drain chunk i
// (tcgen05.ld, 32 codes at a time) the MMA warp already contracts chunk i+1.  Because a TMEM lane
// IS a frame, each epilogue thread keeps the running (best distance, best index) of its own frame in
// registers - the N x K distance matrix never exists in memory and no cross-lane reduction is needed.
// Both operands are K-major (d contiguous): z planes (3, N, Dp), codebook planes (3, K, Dp).
#pragma once
#include "tc_common.cuh"

namespace vnb {

using namespace tc;

constexpr int VQT_BM = 128;        // frames per CTA
constexpr int VQT_BN = 128;        // codes per accumulator chunk
constexpr int VQT_BK = 64;
constexpr int VQT_STAGES = 4;
constexpr int VQT_THREADS = 192;

struct VqTcArgs {
    int64_t N;
    int D, K;
    const float* z_sq;
    const float* e_sq;
    int64_t* idx;
};

__global__ void __launch_bounds__(VQT_THREADS, 1)
vq_lookup_tc_stream_kernel(const __grid_constant__ CUtensorMap map_z, const __grid_constant__ CUtensorMap map_e,
                    const VqTcArgs a) {
    constexpr uint32_t A_BYTES = VQT_BM * VQT_BK * 2, B_BYTES = VQT_BN * VQT_BK * 2;
    constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t full_bar[VQT_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[VQT_STAGES];
    __shared__ __align__(8) uint64_t tmem_full_bar[2];
    __shared__ __align__(8) uint64_t tmem_empty_bar[2];
    __shared__ uint32_t tmem_slot;

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // warp-uniform for the compiler
    const int64_t n0 = (int64_t)blockIdx.x * VQT_BM;
    const int chunks = (a.K + VQT_BN - 1) / VQT_BN;
    const int kd = (a.D + VQT_BK - 1) / VQT_BK;
    const int per_chunk = 6 * kd;

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&map_z);
        tma_prefetch_desc(&map_e);
        for (int s = 0; s < VQT_STAGES; s++) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; s++) {
            mbar_init(&tmem_full_bar[s], 1);
            mbar_init(&tmem_empty_bar[s], 4);          // one arrival per epilogue warp
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<2 * VQT_BN>(&tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;
    pdl_wait();      // everything above (barriers, TMEM, tensor-map prefetch) overlapped the previous kernel's tail

    // producer and issuer loops are walked by their whole (converged) warp so that counters and descriptors live in
    // uniform registers; one elected lane issues the TMA / tcgen05 instructions (see conv_fwd_tc_kernel)
    if (warp == 0) {
        {
            int kc = 0;
            for (int ch = 0; ch < chunks; ch++)
                for (int pr = 0; pr < 6; pr++)
                    for (int c = 0; c < kd; c++, kc++) {
                        const int s = kc % VQT_STAGES;
                        if (kc >= VQT_STAGES) mbar_wait(&empty_bar[s], ((kc / VQT_STAGES) - 1) & 1);
                        uint8_t* sa = smem + (size_t)s * STAGE_BYTES;
                        if (elect_one()) {
                            mbar_expect_tx(&full_bar[s], STAGE_BYTES);
                            tma_load_3d(sa, &map_z, &full_bar[s], c * VQT_BK, (int)n0, c_pair_a[pr]);
                            tma_load_3d(sa + A_BYTES, &map_e, &full_bar[s], c * VQT_BK, ch * VQT_BN, c_pair_b[pr]);
                        }
                        __syncwarp();
                    }
        }
    } else if (warp == 1) {
        {
            constexpr uint32_t idesc = instr_desc_bf16(VQT_BM, VQT_BN, 0, 0);
            int kc = 0;
            for (int ch = 0; ch < chunks; ch++) {
                const int buf = ch & 1;
                if (ch >= 2) {
                    mbar_wait(&tmem_empty_bar[buf], ((ch >> 1) - 1) & 1);     // epilogue drained this buffer
                    tc_fence_after();
                }
                const uint32_t tacc = tmem_base + (uint32_t)(buf * VQT_BN);
                for (int i = 0; i < per_chunk; i++, kc++) {
                    const int s = kc % VQT_STAGES;
                    mbar_wait(&full_bar[s], (kc / VQT_STAGES) & 1);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + (size_t)s * STAGE_BYTES);
                    const uint32_t sb = sa + A_BYTES;
                    if (elect_one()) {
#pragma unroll
                        for (int j = 0; j < VQT_BK / 16; j++)
                            umma_bf16(tacc, smem_desc_sw128(sa + j * 32, 16, 1024), smem_desc_sw128(sb + j * 32, 16, 1024),
                                      idesc, (i | j) != 0);
                        umma_commit(&empty_bar[s]);
                    }
                    __syncwarp();
                }
                if (elect_one()) umma_commit(&tmem_full_bar[buf]);
                __syncwarp();
            }
        }
    } else {
        // epilogue: lane = frame; running argmin over the codebook in registers
        const int q = warp & 3;
        const int64_t n = n0 + q * 32 + lane;
        const bool row_ok = n < a.N;
        const float zq = row_ok ? a.z_sq[n] : 0.f;
        float best = INFINITY;
        int bestk = 0;
        for (int ch = 0; ch < chunks; ch++) {
            const int buf = ch & 1;
            mbar_wait(&tmem_full_bar[buf], (ch >> 1) & 1);
            tc_fence_after();
#pragma unroll 1
            for (int c0 = 0; c0 < VQT_BN; c0 += 32) {
                float v[32];
                __syncwarp();
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * VQT_BN + c0), v);
                const int k0 = ch * VQT_BN + c0;
#pragma unroll
                for (int i = 0; i < 32; i++) {
                    const int k = k0 + i;
                    if (k < a.K) {
                        // torch: (z_sq + e_sq) - 2*dot  (layers_vq.py:103-105)
                        const float dist = fmaf(-2.f, v[i], zq + __ldg(a.e_sq + k));
                        if (dist < best) {
                            best = dist;
                            bestk = k;
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&tmem_empty_bar[buf])) : "memory");
            }
        }
        if (row_ok) a.idx[n] = (int64_t)bestk;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        tc_fence_after();
        tmem_dealloc<2 * VQT_BN>(tmem_base);
    }
}

static int vq_lookup_tc_stream(const void* z_planes, const float* z_sq, const void* e_planes, const float* e_sq, int B, int D, int T,
                 int K, int64_t* idx, cudaStream_t st) {
    const int64_t N = (int64_t)B * T;
    if (D < 8 || K < 8 || N > (int64_t)1 << 30) return 1;
    constexpr size_t smem = (size_t)VQT_STAGES * (VQT_BM + VQT_BN) * VQT_BK * 2 + 1024;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(vq_lookup_tc_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) {
            set_error("vq_lookup_tc: cudaFuncSetAttribute(%zu): %s", smem, cudaGetErrorString(e));
            return VNB_E_CUDA;
        }
        attr_done = true;
    }
    const int Dp = round_up8(D);
    CUtensorMap mz, me;
    if (!make_tmap_3d(&mz, z_planes, (uint64_t)D, (uint64_t)N, 3, (uint64_t)Dp * 2, (uint64_t)N * Dp * 2, VQT_BK, VQT_BM, 1))
        return VNB_E_CUDA;
    if (!make_tmap_3d(&me, e_planes, (uint64_t)D, (uint64_t)K, 3, (uint64_t)Dp * 2, (uint64_t)K * Dp * 2, VQT_BK, VQT_BN, 1))
        return VNB_E_CUDA;
    VqTcArgs a{N, D, K, z_sq, e_sq, idx};
    vnb::launch_k(vq_lookup_tc_stream_kernel, ceil_div(N, VQT_BM), VQT_THREADS, smem, st, mz, me, a);
    VNB_LAUNCH_CHECK("vq_lookup_tc");
    return VNB_OK;
}


// ---------------------------------------------------------------------------------------------------------------
// Resident-frames variant (D <= 128): the main path.
//   * the CTA's 128 frames (all three planes, 48 KB at D = 64) are loaded ONCE and stay in shared memory;
//   * the codebook is walked in chunks of 256 codes (one 128 x 256 MMA = 128 tensor cycles per ~70-cycle issue;
//     the 128-wide version was issue bound at 22-48 % of the tensor roof), each codebook plane of a chunk is
//     loaded once and used by every frame plane it pairs with:  E2 x Z0 | E1 x (Z1, Z0) | E0 x (Z2, Z1, Z0)
//     (small terms first), i.e. 3 instead of 6 operand loads per chunk and 96 KB instead of 288 KB of L2 traffic;
//   * two 256-column accumulators fill TMEM (512 columns): the epilogue drains one while the MMAs fill the other.
// ---------------------------------------------------------------------------------------------------------------
constexpr int VQR_BN = 256;
constexpr uint32_t VQR_A_BYTES = VQT_BM * VQT_BK * 2;        // 16 KB: 128 frames x 64 dims of one plane
constexpr uint32_t VQR_B_BYTES = VQR_BN * VQT_BK * 2;        // 32 KB: 256 codes x 64 dims of one plane

struct VqResArgs {
    int64_t N;
    int D, K, nstages;
    const float* z_sq;
    const float* e_sq;
    int64_t* idx;
};

__global__ void __launch_bounds__(VQT_THREADS, 1)
vq_lookup_tc_kernel(const __grid_constant__ CUtensorMap map_z, const __grid_constant__ CUtensorMap map_e,
                    const VqResArgs a) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t a_bar;
    __shared__ __align__(8) uint64_t full_bar[8];
    __shared__ __align__(8) uint64_t empty_bar[8];
    __shared__ __align__(8) uint64_t tmem_full_bar[2];
    __shared__ __align__(8) uint64_t tmem_empty_bar[2];
    __shared__ uint32_t tmem_slot;

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // warp-uniform for the compiler
    const int64_t n0 = (int64_t)blockIdx.x * VQT_BM;
    const int chunks = (a.K + VQR_BN - 1) / VQR_BN;
    const int kd = (a.D + VQT_BK - 1) / VQT_BK;
    const int STAGES = a.nstages;
    uint8_t* smem_a = smem;                                       // [plane][d-chunk] 16 KB tiles
    uint8_t* smem_b = smem + (size_t)3 * kd * VQR_A_BYTES;        // ring of 32 KB stages

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&map_z);
        tma_prefetch_desc(&map_e);
        mbar_init(&a_bar, 1);
        for (int s = 0; s < STAGES; s++) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; s++) {
            mbar_init(&tmem_full_bar[s], 1);
            mbar_init(&tmem_empty_bar[s], 4);          // one arrival per epilogue warp
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<2 * VQR_BN>(&tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;
    pdl_wait();      // everything above (barriers, TMEM, tensor-map prefetch) overlapped the previous kernel's tail

    if (warp == 0) {
        {
            // ===== producer (converged warp, elected lane issues): the frames once, then the codebook planes chunk
            // by chunk =====
            if (elect_one()) {
                mbar_expect_tx(&a_bar, (uint32_t)3 * kd * VQR_A_BYTES);
                for (int pl = 0; pl < 3; pl++)
                    for (int c = 0; c < kd; c++)
                        tma_load_3d(smem_a + (size_t)(pl * kd + c) * VQR_A_BYTES, &map_z, &a_bar, c * VQT_BK, (int)n0, pl);
            }
            __syncwarp();
            int s = 0, i = 0;
            uint32_t ph = 1;
            for (int ch = 0; ch < chunks; ch++)
                for (int pb = 2; pb >= 0; pb--)
                    for (int c = 0; c < kd; c++, i++) {
                        if (i >= STAGES) mbar_wait(&empty_bar[s], ph);
                        if (elect_one()) {
                            mbar_expect_tx(&full_bar[s], VQR_B_BYTES);
                            tma_load_3d(smem_b + (size_t)s * VQR_B_BYTES, &map_e, &full_bar[s], c * VQT_BK, ch * VQR_BN, pb);
                        }
                        __syncwarp();
                        if (++s == STAGES) { s = 0; ph ^= 1; }
                    }
        }
    } else if (warp == 1) {
        {
            // ===== MMA issuer (converged warp, elected lane issues) =====
            constexpr uint32_t idesc = instr_desc_bf16(VQT_BM, VQR_BN, 0, 0);
            mbar_wait(&a_bar, 0);
            tc_fence_after();
            const uint32_t sa0 = smem_u32(smem_a), sb0 = smem_u32(smem_b);
            int s = 0;
            uint32_t ph = 0;
            bool ready = __all_sync(0xffffffffu, mbar_test(&full_bar[0], 0));
            const int per_chunk = 3 * kd;
            const int total = chunks * per_chunk;
            int i = 0;
            for (int ch = 0; ch < chunks; ch++) {
                const int buf = ch & 1;
                if (ch >= 2) {
                    mbar_wait(&tmem_empty_bar[buf], ((ch >> 1) - 1) & 1);     // epilogue drained this buffer
                    tc_fence_after();
                }
                const uint32_t tacc = tmem_base + (uint32_t)(buf * VQR_BN);
                uint32_t acc = 0;
                for (int pb = 2; pb >= 0; pb--) {
                    for (int c = 0; c < kd; c++, i++) {
                        if (!ready) mbar_wait(&full_bar[s], ph);
                        tc_fence_after();
                        int s1 = s + 1;
                        uint32_t ph1 = ph;
                        if (s1 == STAGES) { s1 = 0; ph1 ^= 1; }
                        bool next_ready = false;
                        if (i + 1 < total) next_ready = __all_sync(0xffffffffu, mbar_test(&full_bar[s1], ph1));
                        const uint64_t db = smem_desc_sw128(sb0 + (uint32_t)s * VQR_B_BYTES, 16, 1024);
                        // frame planes paired with codebook plane pb: pb = 2 -> {0}; 1 -> {1, 0}; 0 -> {2, 1, 0}
                        if (elect_one()) {
                            uint32_t acc1 = acc;
                            for (int pa = 2 - pb; pa >= 0; pa--) {
                                const uint64_t da = smem_desc_sw128(sa0 + (uint32_t)(pa * kd + c) * VQR_A_BYTES, 16, 1024);
#pragma unroll
                                for (int j = 0; j < VQT_BK / 16; j++) {
                                    umma_bf16(tacc, da + 2 * j, db + 2 * j, idesc, acc1);
                                    acc1 = 1;
                                }
                            }
                            umma_commit(&empty_bar[s]);
                        }
                        __syncwarp();
                        acc = 1;
                        ready = next_ready;
                        s = s1;
                        ph = ph1;
                    }
                }
                if (elect_one()) umma_commit(&tmem_full_bar[buf]);
                __syncwarp();
            }
        }
    } else {
        // ===== epilogue: lane = frame; running argmin over the codebook in registers =====
        const int q = warp & 3;
        const int64_t n = n0 + q * 32 + lane;
        const bool row_ok = n < a.N;
        const float zq = row_ok ? a.z_sq[n] : 0.f;
        float best = INFINITY;
        int bestk = 0;
        for (int ch = 0; ch < chunks; ch++) {
            const int buf = ch & 1;
            mbar_wait_warp(&tmem_full_bar[buf], (ch >> 1) & 1, 20);
            tc_fence_after();
#pragma unroll 1
            for (int c0 = 0; c0 < VQR_BN; c0 += 32) {
                const int k0 = ch * VQR_BN + c0;
                if (k0 >= a.K) break;                                   // warp-uniform
                float v[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * VQR_BN + c0), v);
                const float es = k0 + lane < a.K ? __ldg(a.e_sq + k0 + lane) : INFINITY;
#pragma unroll
                for (int i = 0; i < 32; i++) {
                    // torch: (z_sq + e_sq) - 2*dot  (layers_vq.py:103-105); codes >= K carry +inf
                    const float dist = fmaf(-2.f, v[i], zq + __shfl_sync(0xffffffffu, es, i));
                    if (dist < best) {
                        best = dist;
                        bestk = k0 + i;
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&tmem_empty_bar[buf])) : "memory");
            }
        }
        if (row_ok) a.idx[n] = (int64_t)bestk;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        tc_fence_after();
        tmem_dealloc<2 * VQR_BN>(tmem_base);
    }
}

int vq_lookup_tc(const void* z_planes, const float* z_sq, const void* e_planes, const float* e_sq, int B, int D, int T,
                 int K, int64_t* idx, cudaStream_t st) {
    const int64_t N = (int64_t)B * T;
    if (D < 8 || K < 8 || N > (int64_t)1 << 30) return 1;
    static const int force_stream = getenv("VNB_VQ_STREAM") ? atoi(getenv("VNB_VQ_STREAM")) : 0;
    const int kd = ceil_div(D, VQT_BK);
    const size_t a_bytes = (size_t)3 * kd * VQR_A_BYTES;
    const size_t budget = 227 * 1024 - 2048;
    if (force_stream || a_bytes + 2 * VQR_B_BYTES > budget) return vq_lookup_tc_stream(z_planes, z_sq, e_planes, e_sq, B, D, T, K, idx, st);
    int nst = (int)((budget - a_bytes) / VQR_B_BYTES);
    nst = nst > 6 ? 6 : nst;
    const size_t smem = a_bytes + (size_t)nst * VQR_B_BYTES + 1024;
    static size_t smem_set = 0;
    if (smem > smem_set) {
        cudaError_t e = cudaFuncSetAttribute(vq_lookup_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) {
            set_error("vq_lookup_tc: cudaFuncSetAttribute(%zu): %s", smem, cudaGetErrorString(e));
            return VNB_E_CUDA;
        }
        smem_set = smem;
    }
    const int Dp = round_up8(D);
    CUtensorMap mz, me;
    if (!make_tmap_3d(&mz, z_planes, (uint64_t)D, (uint64_t)N, 3, (uint64_t)Dp * 2, (uint64_t)N * Dp * 2, VQT_BK, VQT_BM, 1))
        return VNB_E_CUDA;
    if (!make_tmap_3d(&me, e_planes, (uint64_t)D, (uint64_t)K, 3, (uint64_t)Dp * 2, (uint64_t)K * Dp * 2, VQT_BK, VQR_BN, 1))
        return VNB_E_CUDA;
    VqResArgs a{N, D, K, nst, z_sq, e_sq, idx};
    vnb::launch_k(vq_lookup_tc_kernel, ceil_div(N, VQT_BM), VQT_THREADS, smem, st, mz, me, a);
    VNB_LAUNCH_CHECK("vq_lookup_tc");
    return VNB_OK;
}

}  // namespace vnb
