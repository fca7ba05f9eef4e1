// tcgen05 weight-gradient GEMM (included by gemm_tc.cu - one translation unit for all tcgen05 kernels).
//
//     dW[tap][co, ci] = sum_{pair} sum_{b} sum_{t} dY[pa][b][t][co] * U[pb][b][t + off(tap)][ci]
//
// The contraction runs over FRAMES, which are the rows of the channels-last planes, so both operands
// are MN-major tiles: 128 frame rows x 64 channels (128 B, SWIZZLE_128B) per box, the tap shift is
// again a TMA row coordinate and the per-sample zero padding is out-of-bounds fill.
//   A = "row" planes (M = 128 channels = two 64-channel boxes), B = "column" planes (N = BN channels).
//   For Conv1d the rows are dY channels (co) and the columns U channels (ci): parameter (co, ci, k).
//   For ConvTranspose1d the host SWAPS the operands (rows = ci from U, columns = co from dY), so that the
//   parameter layout (ci, co, k) is again row-major in (row, column) and the stores stay coalesced.
//   One TMA instruction per operand and stage: a 4-D map (64 ch, frame, 64-channel group, plane*sample)
//   with box {64, 128, groups, 1} lands the groups as consecutive canonical 16 KB boxes (the group dimension
//   has a SMALLER stride than the frame dimension; cuTensorMapEncodeTiled accepts that - tools/tma4d_probe.cu).
//   Channel counts that are not a multiple of 64 use one 3-D box per group instead.
// grid = (ksplit, row tiles * column tiles, taps), cluster (ksplit, 1, 1): the (sample, 64-frame block) list is
// split over the CTAs of a cluster, which exchange their partial tiles through distributed shared memory with
// bulk copies exactly like conv_fwd_tc_kernel (no atomics, no memset, deterministic).
// Epilogue: 4 consecutive columns per thread, BN/4 threads per row, written into the reference parameter
// layout  dw[(row * ncols + col) * taps + tap_index].
#pragma once
#include "tc_common.cuh"

namespace vnb {

using namespace tc;

constexpr int WG_BM = 128;
constexpr int WG_KT = 128;         // frames per pipeline stage (8 MMAs per barrier round trip)
constexpr int WG_THREADS = 256;
constexpr uint32_t WG_BOX_BYTES = WG_KT * 128;      // one 64-channel x 128-frame box

struct WgradTcArgs {
    int B, nrows, ncols, T, taps, P;
    int dil, swapped;
    int a4d, b4d;              // operand maps are 4-D (one TMA per stage) or 3-D (one per 64-channel group)
    int nstages;
    uint32_t recv_off;
    float* dw;
    // frames per pipeline stage (128, or 64 when `multi`) and, round 2, `multi`: a stage holds BOTH planes of both
    // operands of the two-plane mode (P = 2) and the three products hi*hi + hi*mid + mid*hi are issued from it - the
    // operands are loaded twice instead of three times (the k-loop is bound by the SM's L2 intake, profiles/
    // r02b_gemm_sweep_timelines.txt).  P = 3 keeps one pass per plane pair: smallest terms first over the whole
    // reduction is what keeps that mode fp32-faithful.
    int kt, multi;
    int dbg;                   // VNB_TC_DBG = 9: clock64 timeline of CTA (0,0,0) into g_tc_ts[0]
    // round 2, `fuse` (kernel template FUSE, BN = 192): the three taps of a k = 3 conv are three 64-column groups of ONE
    // accumulator - the row operand is loaded once for all taps, the column operand once per tap with its frame shift,
    // and a CTA owns 64 columns x 3 taps = 192 CONSECUTIVE floats of every parameter row (the tap-per-CTA grid wrote
    // every third float: 14 k of 57 k cycles in the store loop, profiles/r02m_wgrad_timelines.txt)
    int fuse;
    uint32_t stage_bytes;      // fuse: bytes of a pipeline stage (64 frame-planes of both operands)
};

template <int BN> struct WgradTcCfg {
    static constexpr int MAX_STAGES = 8;
    static constexpr uint32_t A_BYTES = (WG_BM / 64) * WG_BOX_BYTES;
    static constexpr uint32_t B_BYTES = (BN / 64) * WG_BOX_BYTES;
    static constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int PITCH = BN + 4;
    static constexpr uint32_t STAGING_BYTES = (uint32_t)WG_BM * PITCH * 4;
    static constexpr uint32_t SMEM_BUDGET = 227 * 1024 - 2048;
    static uint32_t recv_bytes(int ks) { return (uint32_t)(ks - 1) * (WG_BM / ks) * PITCH * 4; }
    // stage bytes: STAGE_BYTES (128 frames, one plane) for the one-pass-per-pair path, 2 planes x 64 frames = the same
    // number of bytes for the all-planes-per-stage path - every formula below holds for both
    static bool fits(int ks) {
        if (recv_bytes(ks) + 2 * STAGE_BYTES > SMEM_BUDGET) return false;
        return recv_offset(ks) + recv_bytes(ks) + 1024 <= 227 * 1024 - 1024;
    }
    static int stages(int ks) {
        int n = (int)((SMEM_BUDGET - recv_bytes(ks)) / STAGE_BYTES);
        return n > MAX_STAGES ? MAX_STAGES : n;
    }
    static uint32_t recv_offset(int ks) {
        uint32_t ring = (uint32_t)stages(ks) * STAGE_BYTES;
        uint32_t o = ring > STAGING_BYTES ? ring : STAGING_BYTES;
        return (o + 1023u) & ~1023u;
    }
};

template <int BN, bool FUSE>
__global__ void __launch_bounds__(WG_THREADS, 1)
conv_wgrad_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                     const WgradTcArgs a) {
    using Cfg = WgradTcCfg<BN>;
    constexpr int MAX_STAGES = Cfg::MAX_STAGES;
    constexpr uint32_t A_BYTES = Cfg::A_BYTES, STAGE_BYTES = Cfg::STAGE_BYTES;
    constexpr int PITCH = Cfg::PITCH;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t full_bar[MAX_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[MAX_STAGES];
    __shared__ __align__(8) uint64_t tmem_full_bar;
    __shared__ __align__(8) uint64_t recv_bar;
    __shared__ uint32_t tmem_slot;

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // warp-uniform for the compiler
    const int STAGES = a.nstages;
    const int ks = gridDim.x, kr = blockIdx.x;
    constexpr int NCOL = FUSE ? 64 : BN;                 // parameter columns a CTA owns
    const int ntiles = (a.ncols + NCOL - 1) / NCOL;
    const int m0 = (blockIdx.y / ntiles) * WG_BM, n0 = (blockIdx.y % ntiles) * NCOL;
    const int tap = blockIdx.z;
    const int off = (tap - (a.taps - 1) / 2) * a.dil;
    // FUSE: the row operand never shifts; output tap index g reads the column operand shifted by (g - 1) * dil -
    // for Conv1d that is U[t + off(g)], for the swapped ConvTranspose1d operands the substitution t' = t + off turns
    // sum_t U[t + off][ci] dY[t][co] (parameter tap taps-1-g) into sum_t' U[t'][ci] dY[t' - off][co]: the same rule
    const int off_a = FUSE ? 0 : (a.swapped ? off : 0), off_b = a.swapped ? 0 : off;
    const int tap_idx = a.swapped ? a.taps - 1 - tap : tap;
    const int multi = a.multi, kt = a.kt;
    const int pairs = multi ? 1 : n_pairs(a.P);
    const int tblocks = (a.T + kt - 1) / kt;
    const int kblocks = a.B * tblocks;
    const int kb0 = (int)((int64_t)kblocks * kr / ks), kb1 = (int)((int64_t)kblocks * (kr + 1) / ks);
    const int nk = pairs * (kb1 - kb0);
    // bytes of ONE plane of an operand in a stage (kt frames); with `multi` the P planes follow one another
    const uint32_t box_bytes = (uint32_t)kt * 128u;
    const uint32_t a_plane = (WG_BM / 64) * box_bytes, b_plane = (BN / 64) * box_bytes;
    const uint32_t a_all = multi ? (uint32_t)a.P * a_plane : a_plane;
    const uint32_t stage_stride = FUSE ? a.stage_bytes : STAGE_BYTES;
    long long* ts = (a.dbg >= 9 && (blockIdx.x | blockIdx.y | blockIdx.z) == 0) ? g_tc_ts[0] : nullptr;
    if (threadIdx.x == 0) TS(0);

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&map_a);
        tma_prefetch_desc(&map_b);
        for (int s = 0; s < STAGES; s++) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&tmem_full_bar, 1);
        mbar_init(&recv_bar, 1);
        fence_barrier_init();
        if (ks > 1) mbar_expect_tx(&recv_bar, (uint32_t)(ks - 1) * (WG_BM / ks) * PITCH * 4);
    }
    if (warp == 1) tmem_alloc<(BN == 192 ? 256 : BN)>(&tmem_slot);
    tc_fence_before();
    if (ks > 1) cluster_sync_all();
    else __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;
    pdl_wait();      // everything above (barriers, TMEM, tensor-map prefetch) overlapped the previous kernel's tail
    if (threadIdx.x == 0) TS(1);

    if (warp == 0) {
        if (nk > 0) {
            // ===== TMA producer: the converged warp walks the ring, one elected lane issues (see conv_fwd_tc_kernel);
            // no divisions in the loop =====
            int pr = 0;
            int bb = kb0 / tblocks, tb = kb0 - bb * tblocks;
            int s = 0;
            uint32_t ph = 1;
            for (int i = 0; i < nk; i++) {
                const int pa = a.P == 1 ? 0 : c_pair_a[pair_base(a.P) + pr], pb = a.P == 1 ? 0 : c_pair_b[pair_base(a.P) + pr];
                // the row operand takes the a-plane of the pair when it is dY, the b-plane when it is U (swapped)
                const int pl_a = a.swapped ? pb : pa, pl_b = a.swapped ? pa : pb;
                const int t0 = tb * kt;
                if (i >= STAGES) mbar_wait(&empty_bar[s], ph);
                uint8_t* sa = smem + (size_t)s * stage_stride;
                uint8_t* sb = sa + a_all;
                if (elect_one()) {
                    mbar_expect_tx(&full_bar[s], stage_stride);
                    const int npl = multi ? a.P : 1;
                    for (int q = 0; q < npl; q++) {
                        const int qa = multi ? q : pl_a, qb = multi ? q : pl_b;
                        uint8_t* da_ = sa + (uint32_t)q * a_plane;
                        uint8_t* db_ = sb + (uint32_t)q * b_plane;
                        if (a.a4d) {
                            tma_load_4d(da_, &map_a, &full_bar[s], 0, t0 + off_a, m0 / 64, qa * a.B + bb);
                        } else {
#pragma unroll
                            for (int m = 0; m < WG_BM / 64; m++)
                                tma_load_3d(da_ + m * box_bytes, &map_a, &full_bar[s], m0 + m * 64, t0 + off_a, qa * a.B + bb);
                        }
                        if constexpr (FUSE) {
#pragma unroll
                            for (int g = 0; g < 3; g++)
                                tma_load_4d(db_ + g * box_bytes, &map_b, &full_bar[s], 0, t0 + (g - 1) * a.dil, n0 / 64, qb * a.B + bb);
                        } else if (a.b4d) {
                            tma_load_4d(db_, &map_b, &full_bar[s], 0, t0 + off_b, n0 / 64, qb * a.B + bb);
                        } else {
#pragma unroll
                            for (int n = 0; n < BN / 64; n++)
                                tma_load_3d(db_ + n * box_bytes, &map_b, &full_bar[s], n0 + n * 64, t0 + off_b, qb * a.B + bb);
                        }
                    }
                }
                __syncwarp();
                if (++tb == tblocks) { tb = 0; ++bb; }
                if (bb * tblocks + tb == kb1) { bb = kb0 / tblocks; tb = kb0 - bb * tblocks; ++pr; }   // next plane pair
                if (++s == STAGES) { s = 0; ph ^= 1; }
            }
            if (lane == 0) TS(8);
        }
    } else if (warp == 1) {
        {
            // ===== MMA issuer (converged warp, elected lane) =====
            constexpr uint32_t idesc = instr_desc_bf16(WG_BM, BN, /*a_mn=*/1, /*b_mn=*/1);
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
                const uint32_t sa = smem_u32(smem + (size_t)s * stage_stride);
                // MN-major: 16 frame rows of 128 B = 2048 B per UMMA_K (+128 in 16-byte descriptor units);
                // 64-channel boxes are WG_BOX_BYTES apart (LBO); 8-row swizzle atoms 1024 B apart (SBO)
                if (elect_one()) {
                    if (multi) {
                        // both planes of both operands are in the stage: every significant product from it
                        const int np = n_pairs(a.P), pb0 = pair_base(a.P);
                        for (int pr = 0; pr < np; pr++) {
                            const int pa = c_pair_a[pb0 + pr], pb = c_pair_b[pb0 + pr];
                            const int qa = a.swapped ? pb : pa, qb = a.swapped ? pa : pb;
                            const uint64_t da = smem_desc_sw128(sa + (uint32_t)qa * a_plane, box_bytes, 1024);
                            const uint64_t db = smem_desc_sw128(sa + a_all + (uint32_t)qb * b_plane, box_bytes, 1024);
                            for (int j = 0; j < kt / 16; j++) umma_bf16(tmem_base, da + 128 * j, db + 128 * j, idesc, (i | pr | j) != 0);
                        }
                    } else {
                        const uint64_t da = smem_desc_sw128(sa, box_bytes, 1024), db = smem_desc_sw128(sa + a_all, box_bytes, 1024);
                        for (int j = 0; j < kt / 16; j++) umma_bf16(tmem_base, da + 128 * j, db + 128 * j, idesc, (i | j) != 0);
                    }
                    umma_commit(&empty_bar[s]);
                }
                __syncwarp();
                ready = next_ready;
                s = s1;
                ph = ph1;
            }
            if (elect_one()) umma_commit(&tmem_full_bar);
            __syncwarp();
            if (lane == 0) TS(9);
        }
    }

    // ===== exchange + coalesced store: same scheme as conv_fwd_tc_kernel (see there) =====
    float* stage = reinterpret_cast<float*>(smem);
    const float* recv = reinterpret_cast<const float*>(smem + a.recv_off);
    const int rows_own = WG_BM / ks, r_base = kr * rows_own;
    {
        const int q = warp & 3, h = warp >> 2;
        mbar_wait_warp(&tmem_full_bar, 0, 32);
        tc_fence_after();
        if (threadIdx.x == 64) TS(3);
        float* srow = stage + (size_t)(q * 32 + lane) * PITCH;
#pragma unroll 1
        for (int c0 = h * (BN / 2); c0 < (h + 1) * (BN / 2); c0 += 32) {
            float v[32];
            if (nk > 0) {
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
            } else {
#pragma unroll
                for (int i = 0; i < 32; i++) v[i] = 0.f;
            }
            if constexpr (FUSE) {
                // accumulator column g * 64 + c  ->  position c * 3 + g of the parameter row (64 columns x 3 taps)
                const int g = c0 >> 6, cb = c0 & 63;
#pragma unroll
                for (int i = 0; i < 32; i++) srow[(cb + i) * 3 + g] = v[i];
            } else {
#pragma unroll
                for (int i = 0; i < 32; i += 4)
                    *reinterpret_cast<float4*>(srow + c0 + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
            }
        }
        tc_fence_before();
        fence_proxy_async_smem();
        __syncthreads();
        if (threadIdx.x == 64) TS(4);
        if (threadIdx.x == 64) {
            const uint32_t slot_bytes = (uint32_t)rows_own * PITCH * 4;
            for (int p = 0; p < ks; p++) {
                if (p == kr) continue;
                const int slot = kr < p ? kr : kr - 1;
                bulk_copy_to_cta(map_to_cta(smem_u32(recv) + (uint32_t)slot * slot_bytes, p),
                                 smem_u32(stage + (size_t)p * rows_own * PITCH), slot_bytes,
                                 map_to_cta(smem_u32(&recv_bar), p));
            }
        }
        if (ks > 1) mbar_wait_warp(&recv_bar, 0, 32);
        if (threadIdx.x == 64) TS(5);

        if constexpr (FUSE) {
            // the drain above left every staged row in PARAMETER order (position c * 3 + g): a row of this tile is 192
            // consecutive floats of the parameter.  Warp per row, float4 per lane (48 per row), partial sums in sender order.
            const int ncols = a.ncols, nrows = a.nrows;
            const uint32_t own_b = smem_u32(stage + (size_t)r_base * PITCH), recv_b = smem_u32(recv);
            const uint32_t slot_stride = (uint32_t)rows_own * PITCH * 4;
            for (int rl = warp; rl < rows_own; rl += WG_THREADS / 32) {
                const int row = m0 + r_base + rl;
                if (row >= nrows) break;
                float* o = a.dw + ((size_t)row * ncols + n0) * 3;
                const uint32_t roff = (uint32_t)rl * (PITCH * 4);
#pragma unroll
                for (int v4 = lane; v4 < 48; v4 += 32) {
                    float4 v = ld_shared_f4((kr == 0 ? own_b : recv_b) + roff + (uint32_t)v4 * 16);
#pragma unroll 1
                    for (int p = 1; p < ks; p++) {
                        const uint32_t base = p == kr ? own_b : recv_b + (uint32_t)(p < kr ? p : p - 1) * slot_stride;
                        const float4 x = ld_shared_f4(base + roff + (uint32_t)v4 * 16);
                        v.x += x.x; v.y += x.y; v.z += x.z; v.w += x.w;
                    }
                    if (a.dbg != 31) *reinterpret_cast<float4*>(o + 4 * v4) = v;
                }
            }
        } else {
        constexpr int TPR = BN / 4;
        constexpr int RPP = WG_THREADS / TPR;
        const int tid = threadIdx.x;
        const int cv = (tid % TPR) * 4, rsub = tid / TPR;
        const int col = n0 + cv;
        const int ncols = a.ncols, nrows = a.nrows, taps = a.taps;
        const bool col_ok = col < ncols;
        const bool vec_ok = taps == 1 && (ncols & 3) == 0;
        float* __restrict__ dw = a.dw;
        const uint32_t own_base = smem_u32(stage + (size_t)r_base * PITCH + cv);
        const uint32_t recv_base = smem_u32(recv + cv);
        const uint32_t slot_stride = (uint32_t)rows_own * PITCH * 4;
        // full tiles: no per-element predicates (see conv_fwd_tc_kernel: the generic loop is issue-bound)
        const bool full = m0 + WG_BM <= nrows && n0 + BN <= ncols && (rows_own % RPP) == 0;
        if (full) {
            const int iters = rows_own / RPP;
            const uint32_t rstep = RPP * PITCH * 4;
            const uint32_t own0 = own_base + (uint32_t)rsub * (PITCH * 4), recv0 = recv_base + (uint32_t)rsub * (PITCH * 4);
            float* o = dw + ((size_t)(m0 + r_base + rsub) * ncols + col) * taps + tap_idx;
            const size_t ostep = (size_t)RPP * ncols * taps;
#pragma unroll 2
            for (int it = 0; it < iters; it++, o += ostep) {
                float4 v = ld_shared_f4((kr == 0 ? own0 : recv0) + (uint32_t)it * rstep);
#pragma unroll 1
                for (int p = 1; p < ks; p++) {
                    const uint32_t base = p == kr ? own0 : recv0 + (uint32_t)(p < kr ? p : p - 1) * slot_stride;
                    const float4 x = ld_shared_f4(base + (uint32_t)it * rstep);
                    v.x += x.x; v.y += x.y; v.z += x.z; v.w += x.w;
                }
                if (vec_ok) {
                    *reinterpret_cast<float4*>(o) = v;
                } else {
                    o[0] = v.x;
                    o[taps] = v.y;
                    o[2 * taps] = v.z;
                    o[3 * taps] = v.w;
                }
            }
        } else
#pragma unroll 1
        for (int rl = rsub; rl < rows_own; rl += 2 * RPP) {
            const int rl2 = rl + RPP;
            const int row1 = m0 + r_base + rl, row2 = m0 + r_base + rl2;
            const bool ok1 = col_ok && row1 < nrows;
            const bool ok2 = col_ok && rl2 < rows_own && row2 < nrows;
            float4 acc1 = make_float4(0.f, 0.f, 0.f, 0.f), acc2 = acc1;
#pragma unroll 1
            for (int p = 0; p < ks; p++) {
                const uint32_t base = p == kr ? own_base : recv_base + (uint32_t)(p < kr ? p : p - 1) * slot_stride;
                const float4 x1 = ok1 ? ld_shared_f4(base + (uint32_t)rl * (PITCH * 4)) : make_float4(0.f, 0.f, 0.f, 0.f);
                const float4 x2 = ok2 ? ld_shared_f4(base + (uint32_t)rl2 * (PITCH * 4)) : make_float4(0.f, 0.f, 0.f, 0.f);
                acc1.x += x1.x; acc1.y += x1.y; acc1.z += x1.z; acc1.w += x1.w;
                acc2.x += x2.x; acc2.y += x2.y; acc2.z += x2.z; acc2.w += x2.w;
            }
#pragma unroll 1
            for (int u = 0; u < 2; u++) {
                if (!(u ? ok2 : ok1)) continue;
                const float4 acc = u ? acc2 : acc1;
                float* o = dw + ((size_t)(u ? row2 : row1) * ncols + col) * taps + tap_idx;
                if (vec_ok) {
                    *reinterpret_cast<float4*>(o) = acc;
                } else {
                    o[0] = acc.x;
                    if (col + 1 < ncols) o[taps] = acc.y;
                    if (col + 2 < ncols) o[2 * taps] = acc.z;
                    if (col + 3 < ncols) o[3 * taps] = acc.w;
                }
            }
        }
        }
    }
    if (threadIdx.x == 64) TS(6);
    if (ks > 1) cluster_sync_all();
    __syncthreads();
    if (threadIdx.x == 0) TS(7);
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<(BN == 192 ? 256 : BN)>(tmem_base);
    }
}

// 4-D (C % 64 == 0) or 3-D MN-major operand map of planes (P, B, T, Cp); `groups` = 64-channel boxes per load
static bool make_wgrad_map(CUtensorMap* m, const void* base, int C, int T, int PB, int groups, int* is4d, int kt = WG_KT) {
    const int Cp = round_up8(C);
    if (C % 64 == 0) {
        const uint64_t dims[4] = {64, (uint64_t)T, (uint64_t)C / 64, (uint64_t)PB};
        const uint64_t strides[3] = {(uint64_t)Cp * 2, 128, (uint64_t)T * Cp * 2};
        const uint32_t box[4] = {64, (uint32_t)kt, (uint32_t)groups, 1};
        *is4d = 1;
        return make_tmap(m, base, 4, dims, strides, box);
    }
    *is4d = 0;
    return make_tmap_3d(m, base, (uint64_t)C, (uint64_t)T, (uint64_t)PB, (uint64_t)Cp * 2, (uint64_t)T * Cp * 2, 64, (uint32_t)kt, 1);
}

// shared memory plan of the tap-fused variant (BN = 192): stages of `stage_bytes`, the staging tile reuses the ring
struct WgradFusePlan {
    int nstages;
    uint32_t recv_off, recv_bytes, smem;
    bool ok;
};
static WgradFusePlan wgrad_fuse_plan(int ks, uint32_t stage_bytes) {
    using Cfg = WgradTcCfg<192>;
    WgradFusePlan p{};
    p.recv_bytes = Cfg::recv_bytes(ks);
    if (p.recv_bytes + 2 * stage_bytes > Cfg::SMEM_BUDGET) return p;
    int n = (int)((Cfg::SMEM_BUDGET - p.recv_bytes) / stage_bytes);
    p.nstages = n > Cfg::MAX_STAGES ? Cfg::MAX_STAGES : n;
    const uint32_t ring = (uint32_t)p.nstages * stage_bytes;
    p.recv_off = ((ring > Cfg::STAGING_BYTES ? ring : Cfg::STAGING_BYTES) + 1023u) & ~1023u;
    p.smem = p.recv_off + p.recv_bytes + 1024;
    p.ok = p.smem <= 227 * 1024 - 1024;
    return p;
}

template <int BN, bool FUSE = false>
static int launch_wgrad_tc(const void* rows_pl, const void* cols_pl, int ksplit, WgradTcArgs a, cudaStream_t st) {
    using Cfg = WgradTcCfg<BN>;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(conv_wgrad_tc_kernel<BN, FUSE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 1024);
        if (e != cudaSuccess) {
            set_error("conv_wgrad_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
            return VNB_E_CUDA;
        }
        attr_done = true;
    }
    CUtensorMap ma, mb;
    if (!make_wgrad_map(&ma, rows_pl, a.nrows, a.T, a.P * a.B, WG_BM / 64, &a.a4d, a.kt)) return VNB_E_CUDA;
    if (!make_wgrad_map(&mb, cols_pl, a.ncols, a.T, a.P * a.B, FUSE ? 1 : BN / 64, &a.b4d, a.kt)) return VNB_E_CUDA;
    cudaLaunchConfig_t cfg = {};
    if (FUSE) {
        const WgradFusePlan pl = wgrad_fuse_plan(ksplit, a.stage_bytes);
        if (!pl.ok || !a.b4d) { set_error("conv_wgrad_tc: tap-fused plan does not fit (cluster %d)", ksplit); return VNB_E_ARG; }
        a.nstages = pl.nstages;
        a.recv_off = pl.recv_off;
        cfg.gridDim = dim3(ksplit, ceil_div(a.nrows, WG_BM) * (a.ncols / 64), 1);
        cfg.dynamicSmemBytes = pl.smem;
    } else {
        a.nstages = Cfg::stages(ksplit);
        a.recv_off = Cfg::recv_offset(ksplit);
        cfg.gridDim = dim3(ksplit, ceil_div(a.nrows, WG_BM) * ceil_div(a.ncols, BN), a.taps);
        cfg.dynamicSmemBytes = (size_t)a.recv_off + Cfg::recv_bytes(ksplit) + 1024;
    }
    cfg.blockDim = dim3(WG_THREADS);
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = ksplit;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 2 : 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, conv_wgrad_tc_kernel<BN, FUSE>, ma, mb, a);
    if (e != cudaSuccess) {
        set_error("conv_wgrad_tc: launch failed (BN %d, cluster %d): %s", BN, ksplit, cudaGetErrorString(e));
        return VNB_E_CUDA;
    }
    count_launch();
    return VNB_OK;
}

int conv_wgrad_tc(const void* dy, const void* u, int P, int B, int Cout, int Cin, int T, int taps, int dil,
                  int transposed, float* dw, cudaStream_t st) {
    if (Cout < 8 || Cin < 8 || (int64_t)P * B > 65535 || taps > 65535) return 1;
    // ConvTranspose1d parameters are (ci, co, k): swap the operands so that rows = ci, columns = co
    const int nrows = transposed ? Cin : Cout, ncols = transposed ? Cout : Cin;
    const void* rows_pl = transposed ? u : dy;
    const void* cols_pl = transposed ? dy : u;
    // all planes per stage for the two-plane mode (VNB_OPT_GEMM_V3 = 0 keeps the round-1 pass per plane pair)
    const int multi = (P == 2 && option(VNB_OPT_GEMM_V3) != 0) ? 1 : 0;
    WgradTcArgs a{B, nrows, ncols, T, taps, P, dil, transposed ? 1 : 0, 0, 0, 0, 0, dw, multi ? WG_KT / 2 : WG_KT, multi, option(VNB_OPT_TC_DBG), 0, 0};
    static const int force_bn = getenv("VNB_WG_BN") ? atoi(getenv("VNB_WG_BN")) : 0;
    static const int force_ks = getenv("VNB_WG_KSPLIT") ? atoi(getenv("VNB_WG_KSPLIT")) : 0;
    // default OFF: 15 % faster per launch, no change of the single-GPU step, but the 2-GPU step went from 3.20 to 3.30 ms with
    // it (whatever its CTA budget: profiles/r02m_wgrad_timelines.txt) - the overlapped all-reduce loses more than the
    // weight gradients gain.  VNB_WG_FUSE=1 / vnb_set_option(VNB_OPT_WG_FUSE, 1) switches it on.
    const int fuse_env = option(VNB_OPT_WG_FUSE);
    if (taps == 3 && ncols % 64 == 0 && fuse_env && option(VNB_OPT_GEMM_V3) != 0 && (force_bn == 0 || force_bn == 192)) {
        // tap-fused tiles: 128 rows x 64 columns x 3 taps; 64 frame-planes per stage (40 KB)
        const int npl = multi ? P : 1;
        a.fuse = 1;
        a.kt = 64 / npl >= 16 ? (64 / npl) / 16 * 16 : 16;
        a.stage_bytes = (uint32_t)npl * (uint32_t)a.kt * 128u * (WG_BM / 64 + 3);
        const int kb = multi ? B * ceil_div(T, a.kt) : B * ceil_div(T, a.kt) * n_pairs(P);
        const int64_t tiles = (int64_t)ceil_div(nrows, WG_BM) * (ncols / 64);
        int ksf = 1;
        static const int max_cta = getenv("VNB_WG_FUSE_MAXCTA") ? atoi(getenv("VNB_WG_FUSE_MAXCTA")) : 148;
        while (ksf < 8 && tiles * ksf * 2 <= max_cta && ksf * 2 <= kb && wgrad_fuse_plan(ksf * 2, a.stage_bytes).ok) ksf *= 2;
        if (force_ks) ksf = force_ks;
        return launch_wgrad_tc<192, true>(rows_pl, cols_pl, ksf, a, st);
    }
    // stages walked by a tile: (sample, frame block) x plane pairs - or, all planes per stage, half-size frame blocks
    const int kblocks = multi ? B * ceil_div(T, WG_KT / 2) : B * ceil_div(T, WG_KT) * n_pairs(P);
    int bn = 64, ks = 1;
    int64_t best = INT64_MAX;
    for (int cand = 64; cand <= 256; cand *= 2) {
        if (cand > 64 && ncols <= cand / 2) break;
        const int64_t tiles = (int64_t)ceil_div(nrows, WG_BM) * ceil_div(ncols, cand) * taps;
        for (int k = 1; k <= 8; k *= 2) {
            const bool ok = cand == 256 ? WgradTcCfg<256>::fits(k) : cand == 128 ? WgradTcCfg<128>::fits(k) : WgradTcCfg<64>::fits(k);
            if (k > kblocks || !ok) break;
            const int64_t waves = (tiles * k + 147) / 148;
            // same cost model as conv_fwd_tc (cycles): a 128-frame stage is two of its 64-deep chunks (L2-intake
            // bound: ~745 / 860 / 1380 cycles at BN = 64 / 128 / 256), same fixed parts
            static const int64_t per_stage[3] = {745, 860, 1380};
            static const int64_t fixed[3][4] = {{4700, 8600, 8900, 9500}, {6300, 10600, 11600, 12400}, {9000, 14000, 15000, 16000}};
            const int ci = cand == 64 ? 0 : cand == 128 ? 1 : 2, ki = k == 1 ? 0 : k == 2 ? 1 : k == 4 ? 2 : 3;
            const int64_t cost = waves * ((int64_t)ceil_div(kblocks, k) * per_stage[ci] + fixed[ci][ki]);
            if (cost < best) { best = cost; bn = cand; ks = k; }
        }
    }
    if (force_bn) bn = force_bn;
    if (force_ks) ks = force_ks;
    if (bn == 256) return launch_wgrad_tc<256>(rows_pl, cols_pl, ks, a, st);
    if (bn == 128) return launch_wgrad_tc<128>(rows_pl, cols_pl, ks, a, st);
    return launch_wgrad_tc<64>(rows_pl, cols_pl, ks, a, st);
}

}  // namespace vnb
