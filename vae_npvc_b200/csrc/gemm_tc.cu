// tcgen05 / TMEM / TMA GEMM backend (sm_100a).
//
// conv_fwd_tc: stride-1 'same' Conv1d (and its data-gradient) as an implicit GEMM per sample
//     D[co, t] = sum_{pair} sum_{tap} sum_{ci} W[pa][tap][co][ci] * X[pb][b][ci][t + off(tap)]
//   A = weight planes   (K-major,  SWIZZLE_128B)  via a 3-D TMA map (ci, co, plane*tap)
//   B = activation planes (N-major: the reference's NCT layout as it is) via a 3-D TMA map
//       (t, ci, plane*b); the tap shift is a TMA coordinate, the zero padding is TMA OOB fill
//   accumulator 128 x BN fp32 in TMEM; epilogue warps tcgen05.ld it, add bias / residual and store.
//   P = 3 (fp32-faithful) contracts the 6 significant hi/mid/lo plane pairs on the same pipeline.
// Warp roles (192 threads): warp 0 TMA producer, warp 1 MMA issuer + TMEM allocator,
// warps 2..5 epilogue (TMEM lane quarter = warp_id % 4).
#include "tc_common.cuh"

namespace vnb {

using namespace tc;

constexpr int TC_BM = 128;         // co rows per tile (UMMA M)
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
};

template <int BN>
__global__ void __launch_bounds__(TC_THREADS, 1)
conv_fwd_tc_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_x,
                   const ConvTcArgs a) {
    constexpr uint32_t A_BYTES = TC_BM * TC_BK * 2;              // 16 KB
    constexpr uint32_t B_CHUNK = 64 * TC_BK * 2;                 // 8 KB per 64 frames
    constexpr uint32_t B_BYTES = (BN / 64) * B_CHUNK;
    constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // 1024-byte alignment is required by SWIZZLE_128B
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t full_bar[TC_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[TC_STAGES];
    __shared__ __align__(8) uint64_t tmem_full_bar;
    __shared__ uint32_t tmem_slot;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int t0 = blockIdx.x * BN, co0 = blockIdx.y * TC_BM, b = blockIdx.z;
    const int pairs = a.P == 1 ? 1 : 6;
    const int kc_per = (a.Cin + TC_BK - 1) / TC_BK;
    const int nk = pairs * a.taps * kc_per;

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

    if (warp == 0) {
        if (lane == 0) {
            // ===== TMA producer =====
            int kc = 0;
            for (int pr = 0; pr < pairs; pr++) {
                const int pa = a.P == 1 ? 0 : c_pair_a[pr], pb = a.P == 1 ? 0 : c_pair_b[pr];
                for (int tap = 0; tap < a.taps; tap++) {
                    const int off = (tap - (a.taps - 1) / 2) * a.dil;
                    for (int c = 0; c < kc_per; c++, kc++) {
                        const int s = kc % TC_STAGES;
                        if (kc >= TC_STAGES) mbar_wait(&empty_bar[s], ((kc / TC_STAGES) - 1) & 1);
                        uint8_t* sa = smem + (size_t)s * STAGE_BYTES;
                        uint8_t* sb = sa + A_BYTES;
                        mbar_expect_tx(&full_bar[s], STAGE_BYTES);
                        tma_load_3d(sa, &map_w, &full_bar[s], c * TC_BK, co0, pa * a.taps + tap);
#pragma unroll
                        for (int n = 0; n < BN / 64; n++)
                            tma_load_3d(sb + n * B_CHUNK, &map_x, &full_bar[s], t0 + off + n * 64, c * TC_BK,
                                        pb * a.B + b);
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // ===== MMA issuer =====
            constexpr uint32_t idesc = instr_desc_bf16(TC_BM, BN, /*a_mn=*/0, /*b_mn=*/1);
            for (int kc = 0; kc < nk; kc++) {
                const int s = kc % TC_STAGES;
                mbar_wait(&full_bar[s], (kc / TC_STAGES) & 1);
                tc_fence_after();
                const uint32_t sa = smem_u32(smem + (size_t)s * STAGE_BYTES);
                const uint32_t sb = sa + A_BYTES;
#pragma unroll
                for (int j = 0; j < TC_BK / 16; j++) {
                    // A K-major: 16 K-elements = 32 bytes along the swizzled row; 8-row groups 1024 B apart
                    const uint64_t da = smem_desc_sw128(sa + j * 32, 16, 1024);
                    // B N-major: 16 k-rows of 128 B = 2048 B; 64-frame chunks B_CHUNK apart (LBO); 8-row groups 1024 B (SBO)
                    const uint64_t db = smem_desc_sw128(sb + j * 2048, B_CHUNK, 1024);
                    umma_bf16(tmem_base, da, db, idesc, (kc | j) != 0);
                }
                umma_commit(&empty_bar[s]);           // frees the smem stage when these MMAs retire
            }
            umma_commit(&tmem_full_bar);              // accumulator complete
        }
    } else {
        // ===== epilogue: warps 2..5, TMEM lane quarter = warp % 4 =====
        const int q = warp & 3;
        const int co = co0 + q * 32 + lane;
        mbar_wait(&tmem_full_bar, 0);
        tc_fence_after();
        const bool row_ok = co < a.Cout;
        const float bv = (a.bias && row_ok) ? a.bias[(size_t)b * a.bias_bstride + co] : 0.f;
        const size_t row_off = ((size_t)b * a.Cout + (row_ok ? co : 0)) * a.T;
        const bool vec_ok = (a.T & 3) == 0;
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 32) {
            float v[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);     // warp-collective
            const int t = t0 + c0;
            if (!row_ok || t >= a.T) continue;
            float* op = a.out + row_off + t;
            const float* rp = a.residual ? a.residual + row_off + t : nullptr;
            if (vec_ok && t + 32 <= a.T) {
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                    float4 r = make_float4(v[i] + bv, v[i + 1] + bv, v[i + 2] + bv, v[i + 3] + bv);
                    if (rp) {
                        const float4 x = *reinterpret_cast<const float4*>(rp + i);
                        r.x += x.x; r.y += x.y; r.z += x.z; r.w += x.w;
                    }
                    if (a.accumulate) {
                        const float4 x = *reinterpret_cast<const float4*>(op + i);
                        r.x += x.x; r.y += x.y; r.z += x.z; r.w += x.w;
                    }
                    *reinterpret_cast<float4*>(op + i) = r;
                }
            } else {
#pragma unroll
                for (int i = 0; i < 32; i++) {
                    if (t + i < a.T) {
                        float r = v[i] + bv;
                        if (rp) r += rp[i];
                        if (a.accumulate) r += op[i];
                        op[i] = r;
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
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
static int launch_conv_tc(const CUtensorMap& mw, const CUtensorMap& mx, const ConvTcArgs& a, cudaStream_t st) {
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
    dim3 grid(ceil_div(a.T, BN), ceil_div(a.Cout, TC_BM), a.B);
    conv_fwd_tc_kernel<BN><<<grid, TC_THREADS, smem, st>>>(mw, mx, a);
    VNB_LAUNCH_CHECK("conv_fwd_tc");
    return VNB_OK;
}

int conv_fwd_tc(const void* in, int P, int B, int Cin, int T, int Tp, const void* w, int taps, int dil, int Cout,
                const float* bias, int bias_bstride, const float* residual, float* out, int accumulate,
                cudaStream_t st) {
    // shapes this backend takes; everything else goes to the SIMT backend (or errors under VNB_BACKEND_TC)
    if (Cout < 16 || Cin < 8 || (int64_t)P * B > 65535) return 1;
    const int Cin_p = round_up8(Cin);
    CUtensorMap mw, mx;
    // weights (P, taps, Cout, Cin_p): dims (ci, co, plane*tap); box 64 x 128 x 1
    if (!make_tmap_3d(&mw, w, (uint64_t)Cin, (uint64_t)Cout, (uint64_t)P * taps, (uint64_t)Cin_p * 2,
                      (uint64_t)Cout * Cin_p * 2, TC_BK, TC_BM, 1))
        return VNB_E_CUDA;
    // activations (P, B, Cin, Tp): dims (t, ci, plane*b); box 64 x 64 x 1
    if (!make_tmap_3d(&mx, in, (uint64_t)T, (uint64_t)Cin, (uint64_t)P * B, (uint64_t)Tp * 2, (uint64_t)Cin * Tp * 2,
                      64, TC_BK, 1))
        return VNB_E_CUDA;
    ConvTcArgs a{B, Cin, Cout, T, taps, dil, P, bias, bias_bstride, residual, out, accumulate};
    return T > 64 ? launch_conv_tc<128>(mw, mx, a, st) : launch_conv_tc<64>(mw, mx, a, st);
}

int conv_wgrad_tc(const void*, const void*, int, int, int, int, int, int, int, int, int, float*, cudaStream_t) {
    return 1;
}
int vq_lookup_tc(const void*, const float*, const void*, const float*, int, int, int, int, int, int64_t*,
                 cudaStream_t) {
    return 1;
}

}  // namespace vnb
