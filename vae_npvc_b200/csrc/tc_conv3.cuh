// Third-generation tcgen05 conv GEMM (forward and data gradient), included by gemm_tc.cu.
//
//     D[t, co] = sum_{(px, pw) in pairs(P)} sum_{tap} sum_{ci} X[px][b][t + off(tap)][ci] * W[pw][tap][co][ci]
//
// What changed against conv_fwd_tc_kernel (the round-1 kernel, kept for cluster split-K shapes), and why - measured
// on B200 (profiles/r01_conv_tc_timeline.txt, profiles/r01_ncu_conv_fwd_dominant.txt):
//   * the k-loop of a 128 x 128 tile is bound by what ONE SM takes in from L2 (~76 B/cycle: 430 cycles per 32 KB
//     chunk against 256 cycles of MMA), and L2 -> SM traffic was 14.7x the operand bytes.
//     -> CTA PAIRS (cta_group::2): two CTAs on the two SMs of a TPC own two 128-frame tiles and ONE weight tile; each
//        loads its own activations and only HALF of the weight rows, and the leader issues M = 256 MMAs.  Per SM and
//        64-deep chunk 16 KB + 8 KB instead of 16 KB + 16 KB at BN = 128 - and 16 KB + 4 KB instead of 16 + 8 at BN = 64.
//   * the multi-plane operand modes (P = 2: hi + mid, 3 products; P = 3: 6 products) walked the k-loop once per
//     plane PAIR, loading both operands again for every pair (3x / 6x the traffic of P = 1).
//     -> a pipeline stage now holds ALL P planes of both operands of a chunk and every pair product is issued from it:
//        P = 2 loads 2x (not 3x) and P = 3 loads 3x (not 6x) the bytes of P = 1, and a stage carries 12 / 24 MMAs
//        per barrier round trip.
// Same data layout, TMA maps (tap shift = row coordinate, conv padding = out-of-bounds fill), warp roles and
// smem-staged coalesced epilogue (bias / residual / accumulate / fused GroupNorm statistics) as the round-1 kernel.
// No cluster split-K here: the pair already halves the tile count per SM pair; shapes that want split-K stay on the
// round-1 kernel (host cost model in conv_fwd_tc).
#pragma once
#include "tc_common.cuh"

namespace vnb {

using namespace tc;

struct ConvTc3Args {
    int B, Cin, Cout, T, taps, dil, P;
    int mt_total;         // B * ceil(T / 128): 128-frame tiles
    int kch, nstages;
    const float* bias;
    int bias_bstride;
    const float* residual;
    float* out;
    int accumulate;
    double* gn_sums;
    int gn_groups;
    int dbg;
    int fast_epi;
};

template <int BN, bool CTA2>
__global__ void __launch_bounds__(TC_THREADS, 1)
conv_fwd_tc3_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                    const ConvTc3Args a) {
    constexpr int BROWS = CTA2 ? BN / 2 : BN;                     // weight rows THIS CTA loads
    constexpr uint32_t A_BYTES = TC_BM * TC_BK * 2, B_BYTES = BROWS * TC_BK * 2;
    constexpr int PITCH = BN + 4;
    constexpr int MAX_STAGES = 8;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t full_bar[MAX_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[MAX_STAGES];
    __shared__ __align__(8) uint64_t tmem_full_bar;
    __shared__ uint32_t tmem_slot;

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
    // grid (2 | 1, n tiles, tile pairs | tiles), cluster (2 | 1, 1, 1): blockIdx.x = rank in the pair
    const int rank = CTA2 ? (int)blockIdx.x : 0;
    const int mtile = (int)blockIdx.z * (CTA2 ? 2 : 1) + rank;
    const bool live = mtile < a.mt_total;                         // an odd tile count leaves the last peer idle
    const int mtiles = (a.T + TC_BM - 1) / TC_BM;
    const int b = live ? mtile / mtiles : 0;
    // an idle CTA still loads (the pair's barrier counts its bytes): rows beyond T are out-of-bounds zero fill
    const int t0 = live ? (mtile - b * mtiles) * TC_BM : a.T + TC_BM + a.taps * a.dil;
    const int co0 = blockIdx.y * BN;
    const int P = a.P, kch = a.kch, STAGES = a.nstages;
    const int npair = n_pairs(P), pbase = pair_base(P);
    const uint32_t A_STAGE = (uint32_t)(P * kch) * A_BYTES;
    const uint32_t STAGE_BYTES = (uint32_t)(P * kch) * (A_BYTES + B_BYTES);
    const int kc_per = ((a.Cin + TC_BK - 1) / TC_BK) / kch;
    const int nk = a.taps * kc_per;
    long long* ts = (a.dbg >= 9 && (blockIdx.x | blockIdx.y | blockIdx.z) == 0) ? g_tc_ts[0] : nullptr;
    if (threadIdx.x == 0) TS(0);

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&map_w);
        tma_prefetch_desc(&map_x);
        for (int s = 0; s < STAGES; s++) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&tmem_full_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) {
        if constexpr (CTA2) tmem_alloc2<BN>(&tmem_slot);
        else tmem_alloc<BN>(&tmem_slot);
    }
    tc_fence_before();
    if constexpr (CTA2) cluster_sync_all();       // the peer's barriers must exist before anything signals them
    else __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;
    pdl_wait();
    if (threadIdx.x == 0) TS(1);

    if (warp == 0) {
        // ===== TMA producer (both CTAs of a pair): converged warp, one elected lane issues =====
        const uint32_t fb0 = CTA2 ? map_to_cta(smem_u32(&full_bar[0]), 0) : smem_u32(&full_bar[0]);
        int tap = 0, c = 0, s = 0;
        uint32_t ph = 1;
        for (int i = 0; i < nk; i++) {
            const int off = (tap - (a.taps - 1) / 2) * a.dil;
            if (i >= STAGES) mbar_wait(&empty_bar[s], ph);
            const uint32_t sa = smem_u32(smem + (size_t)s * STAGE_BYTES), sb = sa + A_STAGE;
            if (elect_one()) {
                if (!CTA2 || rank == 0) mbar_expect_tx(&full_bar[s], STAGE_BYTES * (CTA2 ? 2u : 1u));
                const uint32_t fb = fb0 + 8u * (uint32_t)s;
                for (int pl = 0; pl < P; pl++) {
                    if (kch == 1) {
                        tma_load_3d_a<CTA2>(sa + (uint32_t)pl * A_BYTES, &map_x, fb, c * TC_BK, t0 + off, pl * a.B + b);
                        tma_load_3d_a<CTA2>(sb + (uint32_t)pl * B_BYTES, &map_w, fb, c * TC_BK, co0 + rank * BROWS, pl * a.taps + tap);
                    } else {
                        tma_load_4d_a<CTA2>(sa + (uint32_t)(pl * kch) * A_BYTES, &map_x, fb, 0, t0 + off, c * kch, pl * a.B + b);
                        tma_load_4d_a<CTA2>(sb + (uint32_t)(pl * kch) * B_BYTES, &map_w, fb, 0, co0 + rank * BROWS, c * kch, pl * a.taps + tap);
                    }
                }
            }
            __syncwarp();
            if (++c == kc_per) { c = 0; ++tap; }
            if (++s == STAGES) { s = 0; ph ^= 1; }
        }
        if (lane == 0) TS(8);                                    // last load issued
    } else if (warp == 1 && (!CTA2 || rank == 0)) {
        // ===== MMA issuer (the pair's leader): every plane-pair product of a chunk from ONE stage =====
        constexpr uint32_t idesc = instr_desc_bf16(CTA2 ? 256 : 128, BN, 0, 0);
        bool ready = nk > 0 && __all_sync(0xffffffffu, mbar_test(&full_bar[0], 0));
        int s = 0;
        uint32_t ph = 0;
        for (int i = 0; i < nk; i++) {
            if (!ready) mbar_wait(&full_bar[s], ph);
            tc_fence_after();
            if (i == 0 && lane == 0) TS(2);
            int s1 = s + 1;
            uint32_t ph1 = ph;
            if (s1 == STAGES) { s1 = 0; ph1 ^= 1; }
            bool next_ready = false;
            if (i + 1 < nk) next_ready = __all_sync(0xffffffffu, mbar_test(&full_bar[s1], ph1));
            const uint32_t sa = smem_u32(smem + (size_t)s * STAGE_BYTES), sb = sa + A_STAGE;
            if (elect_one()) {
                for (int pr = 0; pr < npair; pr++) {
                    const int px = P == 1 ? 0 : c_pair_b[pbase + pr], pw = P == 1 ? 0 : c_pair_a[pbase + pr];
                    const uint64_t da = smem_desc_sw128(sa + (uint32_t)(px * kch) * A_BYTES, 16, 1024);
                    const uint64_t db = smem_desc_sw128(sb + (uint32_t)(pw * kch) * B_BYTES, 16, 1024);
                    for (int ch = 0; ch < kch; ch++) {
#pragma unroll
                        for (int j = 0; j < TC_BK / 16; j++) {
                            const uint64_t xa = da + (uint64_t)(ch * (int)(A_BYTES / 16) + 2 * j);
                            const uint64_t xb = db + (uint64_t)(ch * (int)(B_BYTES / 16) + 2 * j);
                            if constexpr (CTA2) umma_bf16_2cta(tmem_base, xa, xb, idesc, (i | pr | ch | j) != 0);
                            else umma_bf16(tmem_base, xa, xb, idesc, (i | pr | ch | j) != 0);
                        }
                    }
                }
                if constexpr (CTA2) umma_commit_2cta(&empty_bar[s], 3);       // frees the stage in BOTH CTAs
                else umma_commit(&empty_bar[s]);
            }
            __syncwarp();
            ready = next_ready;
            s = s1;
            ph = ph1;
        }
        if (lane == 0) TS(9);                                    // last MMA issued
        if (elect_one()) {
            if constexpr (CTA2) umma_commit_2cta(&tmem_full_bar, 3);
            else umma_commit(&tmem_full_bar);
        }
        __syncwarp();
    }

    // ===== epilogue (every CTA for its own 128 x BN accumulator): TMEM -> shared staging -> coalesced store =====
    float* stage = reinterpret_cast<float*>(smem);
    const bool fast_epi = a.fast_epi && live && !a.residual && !a.accumulate && (a.Cout & 3) == 0 && co0 + BN <= a.Cout &&
                          t0 + TC_BM <= a.T && (!a.gn_sums || ((a.Cout / a.gn_groups) % BN) == 0) &&
                          ((((uintptr_t)a.bias) | ((uintptr_t)a.out) | ((uintptr_t)a.bias_bstride * 4)) & 15) == 0;
    if (fast_epi) {
        mbar_wait_warp(&tmem_full_bar, 0, 32);
        tc_fence_after();
        if (threadIdx.x == 64) TS(3);
        const float* bias = a.bias ? a.bias + (size_t)b * a.bias_bstride + co0 : nullptr;
        double* gp = a.gn_sums ? a.gn_sums + ((size_t)b * a.gn_groups + co0 / (a.Cout / a.gn_groups)) * 2 : nullptr;
        epilogue_regs_bulk<BN>(tmem_base, stage, bias, a.out + ((size_t)b * a.T + t0) * a.Cout + co0, a.Cout, gp, warp, lane);
    } else {
        const int q = warp & 3, h = warp >> 2;
        mbar_wait_warp(&tmem_full_bar, 0, 32);
        tc_fence_after();
        if (threadIdx.x == 64) TS(3);
        if (live) {
            float* srow = stage + (size_t)(q * 32 + lane) * PITCH;
#pragma unroll 1
            for (int c0 = h * (BN / 2); c0 < (h + 1) * (BN / 2); c0 += 32) {
                float v[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
#pragma unroll
                for (int i = 0; i < 32; i += 4)
                    *reinterpret_cast<float4*>(srow + c0 + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
            }
        }
        tc_fence_before();
        __syncthreads();
        if (threadIdx.x == 64) TS(4);
        if (live) {
            constexpr int TPR = BN / 4;
            constexpr int RPP = TC_THREADS / TPR;
            const int tid = threadIdx.x;
            const int cv = (tid % TPR) * 4, rsub = tid / TPR;
            const int co = co0 + cv;
            const int Cout = a.Cout, Tn = a.T;
            const bool vec_ok = (Cout & 3) == 0;
            const bool col_ok = co < Cout;
            const float* __restrict__ resid = a.residual;
            float* __restrict__ outp = a.out;
            const bool accum = a.accumulate != 0;
            const float* bias = a.bias ? a.bias + (size_t)b * a.bias_bstride : nullptr;
            float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
            if (bias && col_ok) {
                if (vec_ok) bv = *reinterpret_cast<const float4*>(bias + co);
                else {
                    bv.x = bias[co];
                    if (co + 1 < Cout) bv.y = bias[co + 1];
                    if (co + 2 < Cout) bv.z = bias[co + 2];
                    if (co + 3 < Cout) bv.w = bias[co + 3];
                }
            }
            const uint32_t own_base = smem_u32(stage + cv);
            const int cg = a.gn_sums ? Cout / a.gn_groups : 1;
            const int g = (a.gn_sums && col_ok) ? co / cg : -1;
            float fs0 = 0.f, fs1 = 0.f;
            const size_t row0 = (size_t)b * Tn + t0;
            const bool full = vec_ok && co0 + BN <= Cout && t0 + TC_BM <= Tn;
            if (full) {
                constexpr int iters = TC_BM / RPP;
                const size_t o0 = (row0 + rsub) * Cout + co;
                const size_t ostep = (size_t)RPP * Cout;
                constexpr uint32_t rstep = RPP * PITCH * 4;
                const uint32_t own0 = own_base + (uint32_t)rsub * (PITCH * 4);
                if (!resid && !accum) {
#pragma unroll 4
                    for (int it = 0; it < iters; it++) {
                        const float4 x = ld_shared_f4(own0 + (uint32_t)it * rstep);
                        const float4 v = make_float4(x.x + bv.x, x.y + bv.y, x.z + bv.z, x.w + bv.w);
                        *reinterpret_cast<float4*>(outp + o0 + it * ostep) = v;
                        fs0 += (v.x + v.y) + (v.z + v.w);
                        fs1 += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
                    }
                } else {
#pragma unroll 2
                    for (int it = 0; it < iters; it++) {
                        float4 r1 = make_float4(0.f, 0.f, 0.f, 0.f), u1 = r1;
                        if (resid) r1 = *reinterpret_cast<const float4*>(resid + o0 + it * ostep);
                        if (accum) u1 = *reinterpret_cast<const float4*>(outp + o0 + it * ostep);
                        const float4 x = ld_shared_f4(own0 + (uint32_t)it * rstep);
                        float4 v;
                        v.x = (bv.x + x.x) + (r1.x + u1.x);
                        v.y = (bv.y + x.y) + (r1.y + u1.y);
                        v.z = (bv.z + x.z) + (r1.z + u1.z);
                        v.w = (bv.w + x.w) + (r1.w + u1.w);
                        *reinterpret_cast<float4*>(outp + o0 + it * ostep) = v;
                        fs0 += (v.x + v.y) + (v.z + v.w);
                        fs1 += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
                    }
                }
            } else {
#pragma unroll 1
                for (int rl = rsub; rl < TC_BM; rl += RPP) {
                    if (!col_ok || t0 + rl >= Tn) continue;
                    const float4 x = ld_shared_f4(own_base + (uint32_t)rl * (PITCH * 4));
                    const float vals[4] = {bv.x + x.x, bv.y + x.y, bv.z + x.z, bv.w + x.w};
                    const size_t o = (row0 + rl) * Cout + co;
#pragma unroll
                    for (int i = 0; i < 4; i++) {
                        if (co + i < Cout) {
                            float y = vals[i];
                            if (resid) y += resid[o + i];
                            if (accum) y += outp[o + i];
                            outp[o + i] = y;
                            fs0 += y;
                            fs1 += y * y;
                        }
                    }
                }
            }
            if (a.gn_sums) {
                double gs0 = (double)fs0, gs1 = (double)fs1;
                const int g0 = __shfl_sync(0xffffffffu, g, 0);
                const bool uniform = __all_sync(0xffffffffu, g == g0 || g < 0);
                if (uniform) {
                    gs0 = warp_sum_f64(gs0);
                    gs1 = warp_sum_f64(gs1);
                    if (lane == 0 && g0 >= 0) {
                        double* gp = a.gn_sums + ((size_t)b * a.gn_groups + g0) * 2;
                        atomicAdd(gp, gs0);
                        atomicAdd(gp + 1, gs1);
                    }
                } else if (g >= 0) {
                    double* gp = a.gn_sums + ((size_t)b * a.gn_groups + g) * 2;
                    atomicAdd(gp, gs0);
                    atomicAdd(gp + 1, gs1);
                }
            }
        }
    }
    if (threadIdx.x == 64) TS(6);
    tc_fence_before();
    if constexpr (CTA2) cluster_sync_all();       // the leader's MMAs read the peer's shared memory until tmem_full fired
    else __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        if constexpr (CTA2) tmem_dealloc2<BN>(tmem_base);
        else tmem_dealloc<BN>(tmem_base);
    }
    if (threadIdx.x == 32) TS(7);
}

// ---- host ----------------------------------------------------------------------------------------------------
struct ConvTc3Plan {
    int bn, cta2, kch, stages;
    size_t smem;
    int64_t cost;
};

static inline bool conv3_plan_one(int bn, int cta2, int P, int Cin, int Cout, int taps, int64_t mt, ConvTc3Plan* out) {
    const int brows = cta2 ? bn / 2 : bn;
    int kch = (P == 1 && Cin % 128 == 0) ? 2 : 1;
    const size_t chunk = (size_t)P * (TC_BM * TC_BK * 2 + brows * TC_BK * 2);
    const size_t staging = (size_t)TC_BM * (bn + 4) * 4;
    size_t budget = 227 * 1024 - 1024 - 1024;                    // alignment slack + static shared (barriers)
    const int cap_kb = option(VNB_OPT_TC3_SMEM_KB);
    if (cap_kb > 0 && (size_t)cap_kb * 1024 < budget && (size_t)cap_kb * 1024 >= staging + 1024) budget = (size_t)cap_kb * 1024 - 1024;
    if (3 * chunk * kch > budget) kch = 1;
    if (2 * chunk > budget) return false;
    int stages = (int)(budget / (chunk * kch));
    if (stages > 8) stages = 8;
    if (stages < 2) return false;
    const size_t ring = (size_t)stages * chunk * kch;
    const size_t smem = (ring > staging ? ring : staging) + 1024;
    if (smem > 227 * 1024 - 1024) return false;
    // cycles per 64-deep chunk: the larger of the MMA time (pairs x 4 MMAs x N/2 cycles) and the L2 intake of this
    // SM (~76 B/cycle measured), + the barrier round trip that is not hidden when a stage holds one chunk
    const int64_t mma = (int64_t)n_pairs(P) * 4 * (bn / 2);
    const int64_t load = (int64_t)(chunk / 76);
    const int64_t per_chunk = (mma > load ? mma : load) + (kch == 1 && P == 1 ? 40 : 15);
    const int64_t fixed = bn == 64 ? 4700 : bn == 128 ? 6300 : 9000;
    const int64_t tiles = (cta2 ? (mt + 1) / 2 : mt) * ceil_div(Cout, bn);
    const int64_t ctas = tiles * (cta2 ? 2 : 1);
    const int64_t waves = (ctas + 147) / 148;
    const int64_t nchunks = (int64_t)taps * ceil_div(Cin, TC_BK);
    out->bn = bn; out->cta2 = cta2; out->kch = kch; out->stages = stages; out->smem = smem;
    out->cost = waves * (nchunks * per_chunk + fixed + (cta2 ? 500 : 0));
    return true;
}

template <int BN, bool CTA2>
static int launch_conv_tc3(const void* in, const void* w, const ConvTc3Plan& pl, ConvTc3Args a, cudaStream_t st) {
    static size_t smem_set = 0;
    if (pl.smem > smem_set) {
        cudaError_t e = cudaFuncSetAttribute(conv_fwd_tc3_kernel<BN, CTA2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem);
        if (e != cudaSuccess) {
            set_error("conv_fwd_tc3: cudaFuncSetAttribute(%zu): %s", pl.smem, cudaGetErrorString(e));
            return VNB_E_CUDA;
        }
        smem_set = pl.smem;
    }
    constexpr int BROWS = CTA2 ? BN / 2 : BN;
    const int Cin_p = round_up8(a.Cin);
    CUtensorMap mx, mw;
    if (pl.kch == 1) {
        if (!make_tmap_3d(&mx, in, (uint64_t)a.Cin, (uint64_t)a.T, (uint64_t)a.P * a.B, (uint64_t)Cin_p * 2,
                          (uint64_t)a.T * Cin_p * 2, TC_BK, TC_BM, 1))
            return VNB_E_CUDA;
        if (!make_tmap_3d(&mw, w, (uint64_t)a.Cin, (uint64_t)a.Cout, (uint64_t)a.P * a.taps, (uint64_t)Cin_p * 2,
                          (uint64_t)a.Cout * Cin_p * 2, TC_BK, BROWS, 1))
            return VNB_E_CUDA;
    } else {
        const uint64_t dx[4] = {64, (uint64_t)a.T, (uint64_t)a.Cin / 64, (uint64_t)a.P * a.B};
        const uint64_t sx[3] = {(uint64_t)Cin_p * 2, 128, (uint64_t)a.T * Cin_p * 2};
        const uint32_t bx[4] = {64, TC_BM, (uint32_t)pl.kch, 1};
        if (!make_tmap(&mx, in, 4, dx, sx, bx)) return VNB_E_CUDA;
        const uint64_t dw[4] = {64, (uint64_t)a.Cout, (uint64_t)a.Cin / 64, (uint64_t)a.P * a.taps};
        const uint64_t sw[3] = {(uint64_t)Cin_p * 2, 128, (uint64_t)a.Cout * Cin_p * 2};
        const uint32_t bw[4] = {64, BROWS, (uint32_t)pl.kch, 1};
        if (!make_tmap(&mw, w, 4, dw, sw, bw)) return VNB_E_CUDA;
    }
    a.kch = pl.kch;
    a.nstages = pl.stages;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(CTA2 ? 2 : 1, ceil_div(a.Cout, BN), CTA2 ? (a.mt_total + 1) / 2 : a.mt_total);
    cfg.blockDim = dim3(TC_THREADS);
    cfg.dynamicSmemBytes = pl.smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CTA2 ? 2 : 1;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 2 : 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, conv_fwd_tc3_kernel<BN, CTA2>, mx, mw, a);
    if (e != cudaSuccess) {
        set_error("conv_fwd_tc3: launch failed (BN %d, pair %d): %s", BN, (int)CTA2, cudaGetErrorString(e));
        return VNB_E_CUDA;
    }
    count_launch();
    return VNB_OK;
}

// -> VNB_OK / error, or 1 when the shape is left to the round-1 kernel.  `v2_cost`: the round-1 kernel's modelled
// cycles for the same shape (INT64_MAX = unknown), so that cluster split-K shapes stay there.
static int conv_fwd_tc3(const void* in, int P, int B, int Cin, int T, const void* w, int taps, int dil, int Cout,
                        const float* bias, int bias_bstride, const float* residual, float* out, int accumulate,
                        double* gn_sums, int gn_groups, int64_t v2_cost, cudaStream_t st) {
    const int mode = option(VNB_OPT_GEMM_V3);      // 0 off, 1 cost model, 2 always
    const int force_bn = option(VNB_OPT_TC3_BN), force_cta2 = option(VNB_OPT_TC3_PAIR), force_st = option(VNB_OPT_TC3_STAGES);
    if (mode == 0) return 1;
    const int64_t mt = (int64_t)ceil_div(T, TC_BM) * B;
    if (mt > 60000) return 1;
    ConvTc3Plan best{};
    best.cost = INT64_MAX;
    for (int bn = 64; bn <= 256; bn *= 2) {
        if (force_bn ? bn != force_bn : (bn > 64 && Cout <= bn / 2)) continue;
        for (int c2 = 0; c2 <= 1; c2++) {
            if (force_cta2 >= 0 && c2 != force_cta2) continue;
            ConvTc3Plan pl;
            if (conv3_plan_one(bn, c2, P, Cin, Cout, taps, mt, &pl) && pl.cost < best.cost) best = pl;
        }
    }
    if (best.cost == INT64_MAX) return 1;
    if (mode == 1) {
        // Measured on B200 at 16 x 128 frames (profiles/r02b_gemm_sweep_timelines.txt): the k-loop is bound by one SM's
        // L2 intake (~74 B/cycle) for every variant.  The pair halves the weight intake (441 -> 367 cycles per chunk at
        // BN = 128) but pays ~2000 cycles of cluster launch + 2-CTA TMEM allocation, a wash at this size; all planes per
        // stage wins where the tile grid fills the chip WITHOUT split-K (decoder conv_in forward, P = 2: 24.0 -> 21.3 us,
        // P = 3: 41.6 -> 32.6 us; res/skip P = 3: 17.0 -> 14.3 us) and loses to the round-1 kernel's cluster split-K
        // where it does not (N = 512: 17.4 vs 20.3 us).  Policy: single-CTA BN = 128 with every plane per stage for the
        // multi-plane modes when >= 100 tiles exist; everything else stays on the round-1 kernel.
        (void)v2_cost;
        // P = 3 (the strict fp32-faithful mode) stays on the round-1 kernel as well: it contracts the plane pairs
        // smallest-terms-first over the WHOLE reduction, which keeps the fp32 accumulation error at 5e-7 (forward
        // features vs float64); interleaving the pairs per stage measured 2.5e-6
        if (P != 2 && !force_bn && force_cta2 < 0) return 1;
        if (!force_bn && force_cta2 < 0) {
            ConvTc3Plan pl;
            if (mt * ceil_div(Cout, 128) < 100 || Cout <= 64 || !conv3_plan_one(128, 0, P, Cin, Cout, taps, mt, &pl)) return 1;
            best = pl;
        }
    }
    if (force_st >= 2 && force_st < best.stages) best.stages = force_st;
    ConvTc3Args a{B, Cin, Cout, T, taps, dil, P, (int)mt, 1, 2, bias, bias_bstride, residual, out, accumulate, gn_sums, gn_groups,
                  option(VNB_OPT_TC_DBG), option(VNB_OPT_FAST_EPI)};
#define VNB_TC3_CASE(BN_, C2_) \
    if (best.bn == BN_ && best.cta2 == C2_) return launch_conv_tc3<BN_, C2_ != 0>(in, w, best, a, st);
    VNB_TC3_CASE(64, 0)
    VNB_TC3_CASE(64, 1)
    VNB_TC3_CASE(128, 0)
    VNB_TC3_CASE(128, 1)
    VNB_TC3_CASE(256, 0)
    VNB_TC3_CASE(256, 1)
#undef VNB_TC3_CASE
    return 1;
}

}  // namespace vnb
