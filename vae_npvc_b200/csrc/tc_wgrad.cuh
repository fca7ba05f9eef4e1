// tcgen05 weight-gradient GEMM (included by gemm_tc.cu - one translation unit for all tcgen05 kernels).
//
//     dW[tap][co, ci] = sum_{pair} sum_{b} sum_{t} dY[pa][b][t][co] * U[pb][b][t + off(tap)][ci]
//
// The contraction runs over FRAMES, which are the rows of the channels-last planes, so both operands
// are MN-major tiles: 64 frame rows x 64 channels (128 B, SWIZZLE_128B) per TMA box, the tap shift is
// again a TMA row coordinate and the per-sample zero padding is out-of-bounds fill.
//   A = dY planes (M = 128 output channels = two 64-channel boxes), 3-D map (co, t, plane*b)
//   B = U  planes (N = BN input channels  = BN/64 boxes),           3-D map (ci, t, plane*b)
// grid = (ci tiles, co tiles, taps * ksplit): the (sample, 64-frame block) list is split over ksplit
// CTAs that reduce into dW with fp32 red.global.add (dW zeroed by the host wrapper when ksplit > 1).
// Epilogue: TMEM lane = co, 32 consecutive ci per tcgen05.ld, scattered into the reference parameter
// layout (d0, d1, k) - Conv1d dw[co][ci][tap], ConvTranspose1d dw[ci][co][taps-1-tap].
#pragma once
#include "tc_common.cuh"

namespace vnb {

using namespace tc;

constexpr int WG_BM = 128;
constexpr int WG_KT = 64;          // frames per pipeline stage
constexpr int WG_STAGES = 4;
constexpr int WG_THREADS = 192;
constexpr uint32_t WG_BOX_BYTES = WG_KT * 128;      // one 64-channel x 64-frame box

struct WgradTcArgs {
    int B, Cout, Cin, T, taps, dil, P, transposed, ksplit, atomic;
    float* dw;
};

template <int BN>
__global__ void __launch_bounds__(WG_THREADS, 1)
conv_wgrad_tc_kernel(const __grid_constant__ CUtensorMap map_dy, const __grid_constant__ CUtensorMap map_u,
                     const WgradTcArgs a) {
    constexpr uint32_t A_BYTES = (WG_BM / 64) * WG_BOX_BYTES;    // 16 KB
    constexpr uint32_t B_BYTES = (BN / 64) * WG_BOX_BYTES;
    constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t full_bar[WG_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[WG_STAGES];
    __shared__ __align__(8) uint64_t tmem_full_bar;
    __shared__ uint32_t tmem_slot;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ci0 = blockIdx.x * BN, co0 = blockIdx.y * WG_BM;
    const int tap = blockIdx.z / a.ksplit, split = blockIdx.z % a.ksplit;
    const int off = (tap - (a.taps - 1) / 2) * a.dil;
    const int pairs = a.P == 1 ? 1 : 6;
    const int tblocks = (a.T + WG_KT - 1) / WG_KT;
    const int kblocks = a.B * tblocks;
    const int kb0 = (int)((int64_t)kblocks * split / a.ksplit), kb1 = (int)((int64_t)kblocks * (split + 1) / a.ksplit);
    const int nk = pairs * (kb1 - kb0);

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&map_dy);
        tma_prefetch_desc(&map_u);
        for (int s = 0; s < WG_STAGES; s++) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<BN>(&tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    if (nk == 0) {
        // nothing to contract for this split (more splits than k-blocks): contributes zero
    } else if (warp == 0) {
        if (lane == 0) {
            int kc = 0;
            for (int pr = 0; pr < pairs; pr++) {
                const int pa = a.P == 1 ? 0 : c_pair_a[pr], pb = a.P == 1 ? 0 : c_pair_b[pr];
                for (int kb = kb0; kb < kb1; kb++, kc++) {
                    const int b = kb / tblocks, t0 = (kb - b * tblocks) * WG_KT;
                    const int s = kc % WG_STAGES;
                    if (kc >= WG_STAGES) mbar_wait(&empty_bar[s], ((kc / WG_STAGES) - 1) & 1);
                    uint8_t* sa = smem + (size_t)s * STAGE_BYTES;
                    uint8_t* sb = sa + A_BYTES;
                    mbar_expect_tx(&full_bar[s], STAGE_BYTES);
#pragma unroll
                    for (int m = 0; m < WG_BM / 64; m++)
                        tma_load_3d(sa + m * WG_BOX_BYTES, &map_dy, &full_bar[s], co0 + m * 64, t0, pa * a.B + b);
#pragma unroll
                    for (int n = 0; n < BN / 64; n++)
                        tma_load_3d(sb + n * WG_BOX_BYTES, &map_u, &full_bar[s], ci0 + n * 64, t0 + off, pb * a.B + b);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = instr_desc_bf16(WG_BM, BN, /*a_mn=*/1, /*b_mn=*/1);
            for (int kc = 0; kc < nk; kc++) {
                const int s = kc % WG_STAGES;
                mbar_wait(&full_bar[s], (kc / WG_STAGES) & 1);
                tc_fence_after();
                const uint32_t sa = smem_u32(smem + (size_t)s * STAGE_BYTES);
                const uint32_t sb = sa + A_BYTES;
#pragma unroll
                for (int j = 0; j < WG_KT / 16; j++) {
                    // MN-major: 16 frame rows of 128 B = 2048 B per UMMA_K; 64-channel boxes are
                    // WG_BOX_BYTES apart (LBO); 8-row swizzle atoms 1024 B apart (SBO)
                    const uint64_t da = smem_desc_sw128(sa + j * 2048, WG_BOX_BYTES, 1024);
                    const uint64_t db = smem_desc_sw128(sb + j * 2048, WG_BOX_BYTES, 1024);
                    umma_bf16(tmem_base, da, db, idesc, (kc | j) != 0);
                }
                umma_commit(&empty_bar[s]);
            }
            umma_commit(&tmem_full_bar);
        }
    } else {
        // epilogue: lane = output channel co
        const int q = warp & 3;
        const int co = co0 + q * 32 + lane;
        mbar_wait(&tmem_full_bar, 0);
        tc_fence_after();
        const bool row_ok = co < a.Cout;
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 32) {
            float v[32];
            __syncwarp();
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
            if (!row_ok) continue;
#pragma unroll
            for (int i = 0; i < 32; i++) {
                const int ci = ci0 + c0 + i;
                if (ci >= a.Cin) break;
                const int64_t o = a.transposed ? (((int64_t)ci * a.Cout + co) * a.taps + (a.taps - 1 - tap))
                                               : (((int64_t)co * a.Cin + ci) * a.taps + tap);
                if (a.atomic) atomicAdd(a.dw + o, v[i]);
                else a.dw[o] = v[i];
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        tc_fence_after();
        tmem_dealloc<BN>(tmem_base);
    }
}

template <int BN>
static int launch_wgrad_tc(const void* dy, const void* u, WgradTcArgs a, cudaStream_t st) {
    constexpr size_t smem = (size_t)WG_STAGES * ((WG_BM / 64) + (BN / 64)) * WG_BOX_BYTES + 1024;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(conv_wgrad_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) {
            set_error("conv_wgrad_tc: cudaFuncSetAttribute(%zu): %s", smem, cudaGetErrorString(e));
            return VNB_E_CUDA;
        }
        attr_done = true;
    }
    const int Cout_p = round_up8(a.Cout), Cin_p = round_up8(a.Cin);
    CUtensorMap mdy, mu;
    if (!make_tmap_3d(&mdy, dy, (uint64_t)a.Cout, (uint64_t)a.T, (uint64_t)a.P * a.B, (uint64_t)Cout_p * 2,
                      (uint64_t)a.T * Cout_p * 2, 64, WG_KT, 1))
        return VNB_E_CUDA;
    if (!make_tmap_3d(&mu, u, (uint64_t)a.Cin, (uint64_t)a.T, (uint64_t)a.P * a.B, (uint64_t)Cin_p * 2,
                      (uint64_t)a.T * Cin_p * 2, 64, WG_KT, 1))
        return VNB_E_CUDA;
    const int base = ceil_div(a.Cin, BN) * ceil_div(a.Cout, WG_BM) * a.taps;
    const int kblocks = a.B * ceil_div(a.T, WG_KT);
    static const int force_ks = getenv("VNB_WG_KSPLIT") ? atoi(getenv("VNB_WG_KSPLIT")) : 0;
    int ksplit = force_ks ? force_ks : ceil_div(148, base);
    if (ksplit > kblocks) ksplit = kblocks;
    if (ksplit < 1) ksplit = 1;
    a.ksplit = ksplit;
    a.atomic = ksplit > 1;
    if (a.atomic) {
        cudaError_t e = cudaMemsetAsync(a.dw, 0, (size_t)a.Cout * a.Cin * a.taps * sizeof(float), st);
        if (e != cudaSuccess) {
            set_error("conv_wgrad_tc: memset: %s", cudaGetErrorString(e));
            return VNB_E_CUDA;
        }
    }
    dim3 grid(ceil_div(a.Cin, BN), ceil_div(a.Cout, WG_BM), a.taps * ksplit);
    conv_wgrad_tc_kernel<BN><<<grid, WG_THREADS, smem, st>>>(mdy, mu, a);
    VNB_LAUNCH_CHECK("conv_wgrad_tc");
    return VNB_OK;
}

int conv_wgrad_tc(const void* dy, const void* u, int P, int B, int Cout, int Cin, int T, int taps, int dil,
                  int transposed, float* dw, cudaStream_t st) {
    if (Cout < 8 || Cin < 8 || (int64_t)P * B > 65535 || taps * 148 > 65535) return 1;
    WgradTcArgs a{B, Cout, Cin, T, taps, dil, P, transposed, 1, 0, dw};
    return Cin > 64 ? launch_wgrad_tc<128>(dy, u, a, st) : launch_wgrad_tc<64>(dy, u, a, st);
}

}  // namespace vnb
