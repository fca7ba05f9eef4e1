// tcgen05 / TMEM / TMA GEMM backend (sm_100a).
//
// conv_fwd_tc: stride-1 'same' Conv1d (and its data-gradient) as an implicit GEMM per sample, on the
// channels-last operand planes:
//     D[t, co] = sum_{pair} sum_{tap} sum_{ci} X[pb][b][t + off(tap)][ci] * W[pa][tap][co][ci]
//   A = activation planes (M = 128 frames, K-major, SWIZZLE_128B) via a 3-D TMA map (ci, t, plane*b):
//       the tap shift is the ROW coordinate (TMA wants 16-byte aligned innermost coordinates, rows are
//       free) and the conv zero padding is TMA out-of-bounds fill, per sample
//   B = weight planes (N = BN output channels, K-major, SWIZZLE_128B) via a 3-D map (ci, co, plane*tap)
//   accumulator 128 x BN fp32 in TMEM; epilogue warps tcgen05.ld it (lane = frame, 32 consecutive
//   output channels per load), add bias / residual and store channels-last rows.
//   P = 3 (fp32-faithful) contracts the 6 significant hi/mid/lo plane pairs on the same pipeline.
// Warp roles (192 threads): warp 0 TMA producer, warp 1 MMA issuer + TMEM allocator,
// warps 2..5 epilogue (TMEM lane quarter = warp_id % 4).
#include <stdlib.h>

#include "tc_common.cuh"

namespace vnb {

using namespace tc;

constexpr int TC_BM = 128;         // frames per tile (UMMA M)
constexpr int TC_BK = 64;          // K elements per pipeline stage (one 128-byte swizzle row)
constexpr int TC_STAGES = 4;
constexpr int TC_THREADS = 192;

// plane pairs (a-plane, b-plane) of the 3-way split, smallest terms first
__constant__ int c_pair_a[6] = {1, 2, 0, 1, 0, 0};
__constant__ int c_pair_b[6] = {1, 0, 2, 0, 1, 0};

struct ConvTcArgs {
    int B, Cin, Cout, T, taps, dil, P;
    const float* bias;
    int bias_bstride;
    const float* residual;
    float* out;
    int accumulate;
    double* gn_sums;      // optional (B, gn_groups, 2): fused GroupNorm sum / sum-of-squares
    int gn_groups;
    int dbg;      // VNB_TC_DBG bisect level (0 = normal)
};

__device__ __forceinline__ double warp_sum_f64(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <int BN>
__global__ void __launch_bounds__(TC_THREADS, 1)
conv_fwd_tc_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                   const ConvTcArgs a) {
    constexpr uint32_t A_BYTES = TC_BM * TC_BK * 2;              // 16 KB of activations
    constexpr uint32_t B_BYTES = BN * TC_BK * 2;                 // BN x 128 B of weights
    constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // 1024-byte alignment is required by SWIZZLE_128B
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t full_bar[TC_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[TC_STAGES];
    __shared__ __align__(8) uint64_t tmem_full_bar;
    __shared__ uint32_t tmem_slot;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int t0 = blockIdx.x * TC_BM, co0 = blockIdx.y * BN, b = blockIdx.z;
    const int pairs = a.P == 1 ? 1 : 6;
    const int kc_per = (a.Cin + TC_BK - 1) / TC_BK;
    int nk = pairs * a.taps * kc_per;
    if (a.dbg && nk > TC_STAGES) nk = TC_STAGES;

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&map_w);
        tma_prefetch_desc(&map_x);
        for (int s = 0; s < TC_STAGES; s++) {
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

    if (a.dbg == 1) {
    } else if (warp == 0) {
        if (lane == 0) {
            // ===== TMA producer =====
            int kc = 0;
            for (int pr = 0; pr < pairs && kc < nk; pr++) {
                const int pa = a.P == 1 ? 0 : c_pair_a[pr], pb = a.P == 1 ? 0 : c_pair_b[pr];
                for (int tap = 0; tap < a.taps && kc < nk; tap++) {
                    const int off = (tap - (a.taps - 1) / 2) * a.dil;
                    for (int c = 0; c < kc_per && kc < nk; c++, kc++) {
                        const int s = kc % TC_STAGES;
                        if (kc >= TC_STAGES) mbar_wait(&empty_bar[s], ((kc / TC_STAGES) - 1) & 1);
                        uint8_t* sa = smem + (size_t)s * STAGE_BYTES;
                        mbar_expect_tx(&full_bar[s], STAGE_BYTES);
                        tma_load_3d(sa, &map_x, &full_bar[s], c * TC_BK, t0 + off, pb * a.B + b);
                        tma_load_3d(sa + A_BYTES, &map_w, &full_bar[s], c * TC_BK, co0, pa * a.taps + tap);
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // ===== MMA issuer =====
            constexpr uint32_t idesc = instr_desc_bf16(TC_BM, BN, /*a_mn=*/0, /*b_mn=*/0);
            for (int kc = 0; kc < nk; kc++) {
                const int s = kc % TC_STAGES;
                mbar_wait(&full_bar[s], (kc / TC_STAGES) & 1);
                tc_fence_after();
                const uint32_t sa = smem_u32(smem + (size_t)s * STAGE_BYTES);
                const uint32_t sb = sa + A_BYTES;
                if (a.dbg == 2) continue;
#pragma unroll
                for (int j = 0; j < TC_BK / 16; j++) {
                    // K-major operands: 16 K-elements = 32 bytes along the swizzled 128-byte row;
                    // 8-row groups are 1024 B apart (SBO); LBO is unused for swizzled K-major layouts
                    const uint64_t da = smem_desc_sw128(sa + j * 32, 16, 1024);
                    const uint64_t db = smem_desc_sw128(sb + j * 32, 16, 1024);
                    umma_bf16(tmem_base, da, db, idesc, (kc | j) != 0);
                }
                umma_commit(&empty_bar[s]);           // frees the smem stage when these MMAs retire
            }
            if (a.dbg != 2) umma_commit(&tmem_full_bar);              // accumulator complete
        }
    } else {
        // ===== epilogue: warps 2..5, TMEM lane quarter = warp % 4; lane = frame =====
        const int q = warp & 3;
        const int t = t0 + q * 32 + lane;
        if (a.dbg == 2) goto done;
        mbar_wait(&tmem_full_bar, 0);
        if (a.dbg == 3) goto done;
        tc_fence_after();
        {
            const bool row_ok = t < a.T;
            const size_t row_off = ((size_t)b * a.T + (row_ok ? t : 0)) * a.Cout;
            const float* bias = a.bias ? a.bias + (size_t)b * a.bias_bstride : nullptr;
            const bool vec_ok = (a.Cout & 3) == 0;
            // fused GroupNorm statistics: per-thread fp64 running sums of the current channel group
            // (32-column blocks never straddle a group: host checks (Cout/groups) % 32 == 0)
            const int cg = a.gn_sums ? a.Cout / a.gn_groups : 1;
            double gs0 = 0.0, gs1 = 0.0;
            int gcur = -1;
#pragma unroll 1
            for (int c0 = 0; c0 <= BN; c0 += 32) {
                const int co = co0 + c0;
                if (a.gn_sums) {
                    const int g = (c0 < BN && co < a.Cout) ? co / cg : -1;       // warp-uniform
                    if (g != gcur) {
                        if (gcur >= 0) {
                            gs0 = warp_sum_f64(gs0);
                            gs1 = warp_sum_f64(gs1);
                            if (lane == 0) {
                                double* gp = a.gn_sums + ((size_t)b * a.gn_groups + gcur) * 2;
                                atomicAdd(gp, gs0);
                                atomicAdd(gp + 1, gs1);
                            }
                        }
                        gs0 = gs1 = 0.0;
                        gcur = g;
                    }
                }
                if (c0 == BN) break;
                float v[32];
                __syncwarp();                                                            // tcgen05.ld is .sync.aligned
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);     // warp-collective
                if (!row_ok || co >= a.Cout || a.dbg == 4) continue;
                float* op = a.out + row_off + co;
                const float* rp = a.residual ? a.residual + row_off + co : nullptr;
                float bs0 = 0.f, bs1 = 0.f;
                if (vec_ok) {
#pragma unroll
                    for (int i = 0; i < 32; i += 4) {
                        if (co + i >= a.Cout) break;
                        float4 r = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
                        if (bias) {
                            const float4 x = *reinterpret_cast<const float4*>(bias + co + i);
                            r.x += x.x; r.y += x.y; r.z += x.z; r.w += x.w;
                        }
                        if (rp) {
                            const float4 x = *reinterpret_cast<const float4*>(rp + i);
                            r.x += x.x; r.y += x.y; r.z += x.z; r.w += x.w;
                        }
                        if (a.accumulate) {
                            const float4 x = *reinterpret_cast<const float4*>(op + i);
                            r.x += x.x; r.y += x.y; r.z += x.z; r.w += x.w;
                        }
                        *reinterpret_cast<float4*>(op + i) = r;
                        bs0 += (r.x + r.y) + (r.z + r.w);
                        bs1 += (r.x * r.x + r.y * r.y) + (r.z * r.z + r.w * r.w);
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 32; i++) {
                        if (co + i < a.Cout) {
                            float r = v[i] + (bias ? bias[co + i] : 0.f);
                            if (rp) r += rp[i];
                            if (a.accumulate) r += op[i];
                            op[i] = r;
                            bs0 += r;
                            bs1 += r * r;
                        }
                    }
                }
                gs0 += (double)bs0;
                gs1 += (double)bs1;
            }
        }
    }
done:
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        tc_fence_after();
        tmem_dealloc<BN>(tmem_base);
    }
}

// --------------------------------------------------------------------------------------------
// host
// --------------------------------------------------------------------------------------------
namespace tc {

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeFn get_encode() {
    static EncodeFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (EncodeFn)p;
    }
    return fn;
}

bool make_tmap_3d(CUtensorMap* out, const void* base, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t stride1_bytes,
                  uint64_t stride2_bytes, uint32_t box0, uint32_t box1, uint32_t box2) {
    EncodeFn enc = get_encode();
    if (!enc) {
        set_error("cuTensorMapEncodeTiled driver entry point not available");
        return false;
    }
    cuuint64_t dims[3] = {d0, d1, d2};
    cuuint64_t strides[2] = {stride1_bytes, stride2_bytes};
    cuuint32_t box[3] = {box0, box1, box2};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (%d): dims %llu %llu %llu strides %llu %llu box %u %u %u", (int)r,
                  (unsigned long long)d0, (unsigned long long)d1, (unsigned long long)d2,
                  (unsigned long long)stride1_bytes, (unsigned long long)stride2_bytes, box0, box1, box2);
        return false;
    }
    return true;
}

}  // namespace tc

template <int BN>
static int launch_conv_tc(const void* in, const void* w, int Tdim, int Cin_p, const ConvTcArgs& a, cudaStream_t st) {
    constexpr size_t smem = (size_t)TC_STAGES * (TC_BM * TC_BK * 2 + BN * TC_BK * 2) + 1024;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(conv_fwd_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) {
            set_error("conv_fwd_tc: cudaFuncSetAttribute(%zu): %s", smem, cudaGetErrorString(e));
            return VNB_E_CUDA;
        }
        attr_done = true;
    }
    CUtensorMap mx, mw;
    // activations (P, B, T, Cin_p): dims (ci, t, plane*b); box 64 x 128 x 1
    if (!make_tmap_3d(&mx, in, (uint64_t)a.Cin, (uint64_t)Tdim, (uint64_t)a.P * a.B, (uint64_t)Cin_p * 2,
                      (uint64_t)Tdim * Cin_p * 2, TC_BK, TC_BM, 1))
        return VNB_E_CUDA;
    // weights (P, taps, Cout, Cin_p): dims (ci, co, plane*tap); box 64 x BN x 1
    if (!make_tmap_3d(&mw, w, (uint64_t)a.Cin, (uint64_t)a.Cout, (uint64_t)a.P * a.taps, (uint64_t)Cin_p * 2,
                      (uint64_t)a.Cout * Cin_p * 2, TC_BK, BN, 1))
        return VNB_E_CUDA;
    dim3 grid(ceil_div(a.T, TC_BM), ceil_div(a.Cout, BN), a.B);
    conv_fwd_tc_kernel<BN><<<grid, TC_THREADS, smem, st>>>(mx, mw, a);
    VNB_LAUNCH_CHECK("conv_fwd_tc");
    return VNB_OK;
}

int conv_fwd_tc(const void* in, int P, int B, int Cin, int T, const void* w, int taps, int dil, int Cout,
                const float* bias, int bias_bstride, const float* residual, float* out, int accumulate,
                double* gn_sums, int gn_groups, cudaStream_t st) {
    // shapes this backend takes; everything else goes to the SIMT backend (or errors under VNB_BACKEND_TC)
    if (Cout < 8 || Cin < 8 || B > 65535) return 1;
    if (gn_sums && ((Cout / gn_groups) & 31) != 0) return 1;      // statistics blocks of 32 channels must not straddle groups
    static const int dbg = getenv("VNB_TC_DBG") ? atoi(getenv("VNB_TC_DBG")) : 0;
    ConvTcArgs a{B, Cin, Cout, T, taps, dil, P, bias, bias_bstride, residual, out, accumulate, gn_sums, gn_groups, dbg};
    const int Cin_p = round_up8(Cin);
    // widest N tile that still gives the 148 SMs about one CTA each
    const int64_t mt = (int64_t)ceil_div(T, TC_BM) * B;
    if (Cout > 128 && mt * ceil_div(Cout, 256) >= 112) return launch_conv_tc<256>(in, w, T, Cin_p, a, st);
    if (Cout > 64 && mt * ceil_div(Cout, 128) >= 112) return launch_conv_tc<128>(in, w, T, Cin_p, a, st);
    return launch_conv_tc<64>(in, w, T, Cin_p, a, st);
}

}  // namespace vnb
// debug: number of bounded mbarrier waits that timed out since load (synchronises the device)
extern "C" int vnb_debug_tc_timeouts(void) {
    int v = -1;
    if (cudaMemcpyFromSymbol(&v, vnb::tc::g_tc_timeouts, sizeof(int)) != cudaSuccess) return -1;
    return v;
}
#include "tc_wgrad.cuh"
#include "tc_vq.cuh"
