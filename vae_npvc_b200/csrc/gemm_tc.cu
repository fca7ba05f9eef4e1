// tcgen05 / TMEM / TMA GEMM backend (sm_100a).
//
// conv_fwd_tc: stride-1 'same' Conv1d (and its data-gradient) as an implicit GEMM per sample, on the
// channels-last operand planes:
//     D[t, co] = sum_{pair} sum_{tap} sum_{ci} X[pb][b][t + off(tap)][ci] * W[pa][tap][co][ci]
//   A = activation planes (M = 128 frames, K-major, SWIZZLE_128B) via a 3-D TMA map (ci, t, plane*b):
//       the tap shift is the ROW coordinate (TMA wants 16-byte aligned innermost coordinates, rows are
//       free) and the conv zero padding is TMA out-of-bounds fill, per sample
//   B = weight planes (N = BN output channels, K-major, SWIZZLE_128B) via a 3-D map (ci, co, plane*tap)
//   accumulator 128 x BN fp32 in TMEM.
//   P = 3 (fp32-faithful) contracts the 6 significant hi/mid/lo plane pairs on the same pipeline.
//
// What the VNB_TC_DBG=9 in-kernel timeline showed on B200 and how the kernel is shaped by it:
//   * one thread pays ~70 cycles to issue a tcgen05.mma, ~190 for an mbarrier wait and ~85 for a commit, whatever
//     the tile; a 128 x 64 MMA executes in 32 cycles, so narrow tiles are ISSUE bound (700 cycles per 64-deep
//     k-step, measured, tensor pipe 13 % busy).  Hence the widest N tile that fits (up to 256: 128 cycles of
//     tensor work per issue) ...
//   * ... and, because 2048 frames x 512 channels are then only 32 tiles, SPLIT-K over a thread-block
//     cluster: the CTAs of a cluster (1, 2, 4 or 8) take consecutive slices of the (pair, tap, ci) loop, park their
//     partial accumulators in their own shared memory, and after one cluster barrier CTA r reduces rows
//     [128 r / ks, 128 (r+1) / ks) of the tile from all peers through distributed shared memory
//   * the same staging makes the epilogue coalesced: bias / residual / accumulate / GroupNorm statistics and
//     the global stores run with 4 consecutive channels per thread and a whole row per (half-)warp, instead
//     of one 2 KB-strided row per thread (1200 cycles per 32 columns before)
// Warp roles (192 threads): warp 0 TMA producer, warp 1 MMA issuer + TMEM allocator,
// warps 2..5 TMEM -> shared staging (TMEM lane quarter = warp_id % 4) and the reduction / store.
#include <stdint.h>
#include <stdlib.h>

#include <type_traits>

#include "tc_common.cuh"

namespace vnb {

using namespace tc;

constexpr int TC_BM = 128;         // frames per tile (UMMA M)
constexpr int TC_BK = 64;          // K elements per pipeline stage (one 128-byte swizzle row)
constexpr int TC_THREADS = 256;     // 8 warps: with one warp per scheduler every epilogue latency is exposed

// plane pairs (a-plane, b-plane), smallest terms first: entries 0..5 = the 6 significant products of the 3-way split
// (P = 3), entries 6..8 = the 3 products of the 2-way split (P = 2: mid*hi, hi*mid, hi*hi)
__constant__ int c_pair_a[9] = {1, 2, 0, 1, 0, 0, 1, 0, 0};
__constant__ int c_pair_b[9] = {1, 0, 2, 0, 1, 0, 0, 1, 0};
__host__ __device__ constexpr int n_pairs(int P) { return P == 1 ? 1 : P == 2 ? 3 : 6; }
__host__ __device__ constexpr int pair_base(int P) { return P == 2 ? 6 : 0; }

// debug timeline (VNB_TC_DBG=9): clock64 stamps of CTA (0,0,0), read by vnb_debug_tc_timeline
__device__ long long g_tc_ts[2][64];
#define TS(i) do { if (ts) ts[i] = clock64(); } while (0)

struct ConvTcArgs {
    int B, Cin, Cout, T, taps, dil, P;
    const float* bias;
    int bias_bstride;
    const float* residual;
    float* out;
    int accumulate;
    double* gn_sums;      // optional (B, gn_groups, 2): fused GroupNorm sum / sum-of-squares
    int gn_groups;
    int dbg;              // VNB_TC_DBG (0 = normal, 9 = record the timeline)
    int nstages;          // depth of the TMA ring (<= 8)
    int kch;              // 64-deep k-chunks per pipeline stage (1, or 2 with 4-D tensor maps when Cin % 128 == 0)
    uint32_t recv_off;    // byte offset of the receive slots in dynamic shared memory
    int fast_epi;         // VNB_OPT_FAST_EPI: register epilogue + bulk row stores for full tiles
    // optional activation backward folded into the epilogue (vnb_conv_fwd_mask): out = mask_scale * acc *
    // act'(mask * mask_scale) [+ residual / accumulate], mask (B, T, Cout) fp32 = the activation's input
    const float* mask;
    int mask_act;
    float mask_scale;
};

__device__ __forceinline__ double warp_sum_f64(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Round 2 - register epilogue for FULL tiles without residual / accumulate (100 of the step's 110 conv launches): the
// bias add and the fused GroupNorm statistics happen on the values as they leave TMEM (a thread owns one frame row and
// 32 consecutive channels: the bias is a warp-uniform broadcast load, the statistics two running sums), the staged tile
// is then handed to the copy engine row by row (cp.async.bulk shared -> global) instead of being re-read and stored by
// the threads.  The store loop this replaces was 2.1k (3.1k with statistics) of a 14-17k cycle kernel
// (profiles/r02b_gemm_sweep_timelines.txt: "staged" -> "stored").
template <int BN>
__device__ __forceinline__ void epilogue_regs_bulk(uint32_t tmem_base, float* stage, const float* __restrict__ bias,
                                                   float* __restrict__ out_tile, int Cout, double* gn_pair, int warp, int lane) {
    constexpr int PITCH = BN + 4;
    const int q = warp & 3, h = warp >> 2;
    float* srow = stage + (size_t)(q * 32 + lane) * PITCH;
    float fs0 = 0.f, fs1 = 0.f;
#pragma unroll 1
    for (int c0 = h * (BN / 2); c0 < (h + 1) * (BN / 2); c0 += 32) {
        float v[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
        if (bias) {
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
                const float4 bb = __ldg(reinterpret_cast<const float4*>(bias + c0 + i));
                v[i] += bb.x; v[i + 1] += bb.y; v[i + 2] += bb.z; v[i + 3] += bb.w;
            }
        }
        if (gn_pair) {
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
                fs0 += (v[i] + v[i + 1]) + (v[i + 2] + v[i + 3]);
                fs1 += (v[i] * v[i] + v[i + 1] * v[i + 1]) + (v[i + 2] * v[i + 2] + v[i + 3] * v[i + 3]);
            }
        }
#pragma unroll
        for (int i = 0; i < 32; i += 4)
            *reinterpret_cast<float4*>(srow + c0 + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
    }
    if (gn_pair) {
        double gs0 = warp_sum_f64((double)fs0), gs1 = warp_sum_f64((double)fs1);
        if (lane == 0) {
            atomicAdd(gn_pair, gs0);
            atomicAdd(gn_pair + 1, gs1);
        }
    }
    tc_fence_before();
    fence_proxy_async_smem();                          // generic-proxy stores -> visible to the bulk-copy engine
    __syncthreads();
    if (threadIdx.x < TC_BM) {
        bulk_store_row(out_tile + (size_t)threadIdx.x * Cout, smem_u32(stage + (size_t)threadIdx.x * PITCH), BN * 4);
        bulk_store_commit_wait_read();
    }
}

// Shared-memory plan (dynamic, after 1 KB alignment):  [ TMA ring: nstages x (16 KB + BN*128 B) | receive slots ]
// The staged 128 x BN fp32 accumulator (row pitch BN+4 floats) reuses the ring; the (ks-1) receive slots of
// rows_own x PITCH floats sit behind max(ring, staging) and are never touched by the main loop.
template <int BN> struct ConvTcCfg {
    static constexpr int MAX_STAGES = 8;
    static constexpr uint32_t A_BYTES = TC_BM * TC_BK * 2;
    static constexpr uint32_t B_BYTES = BN * TC_BK * 2;
    static constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int PITCH = BN + 4;                       // floats per staged accumulator row (16 B pad)
    static constexpr uint32_t STAGING_BYTES = (uint32_t)TC_BM * PITCH * 4;
    static constexpr uint32_t SMEM_BUDGET = 227 * 1024 - 2048;  // minus alignment slack and static barriers
    static uint32_t recv_bytes(int ks) { return (uint32_t)(ks - 1) * (TC_BM / ks) * PITCH * 4; }
    static int stages(int ks, int kch) {
        int n = (int)((SMEM_BUDGET - recv_bytes(ks)) / (STAGE_BYTES * kch));
        return n > MAX_STAGES ? MAX_STAGES : n;
    }
    static uint32_t recv_offset(int ks, int kch) {
        uint32_t ring = (uint32_t)stages(ks, kch) * STAGE_BYTES * kch;
        uint32_t o = ring > STAGING_BYTES ? ring : STAGING_BYTES;
        return (o + 1023u) & ~1023u;
    }
};

template <int BN>
__global__ void __launch_bounds__(TC_THREADS, 1)
conv_fwd_tc_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                   const ConvTcArgs a) {
    using Cfg = ConvTcCfg<BN>;
    constexpr int MAX_STAGES = Cfg::MAX_STAGES;
    const int STAGES = a.nstages;
    constexpr uint32_t A_BYTES = Cfg::A_BYTES, B_BYTES = Cfg::B_BYTES;
    const int kch = a.kch;
    const uint32_t STAGE_BYTES = (uint32_t)kch * Cfg::STAGE_BYTES;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // 1024-byte alignment is required by SWIZZLE_128B (same offset in every CTA of the cluster)
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t full_bar[MAX_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[MAX_STAGES];
    __shared__ __align__(8) uint64_t tmem_full_bar;
    __shared__ __align__(8) uint64_t recv_bar;        // completes when every peer's partial rows have landed here
    __shared__ uint32_t tmem_slot;

    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // warp-uniform for the compiler
    // grid (ksplit, n tiles, samples * frame tiles), cluster (ksplit, 1, 1): blockIdx.x = k-slice = cluster rank
    const int ks = gridDim.x, kr = blockIdx.x;
    const int mtiles = (a.T + TC_BM - 1) / TC_BM;
    const int b = blockIdx.z / mtiles, t0 = (blockIdx.z - b * mtiles) * TC_BM, co0 = blockIdx.y * BN;
    const int pairs = n_pairs(a.P);
    // k runs over (plane pair, tap, group of kch 64-channel chunks): one pipeline stage each
    const int kc_per = ((a.Cin + TC_BK - 1) / TC_BK) / kch;
    const int nkc = pairs * a.taps * kc_per;
    const int kc0 = (int)((int64_t)nkc * kr / ks), kc1 = (int)((int64_t)nkc * (kr + 1) / ks);
    const int nk = kc1 - kc0;                                  // may be 0 (more slices than chunks)
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
        mbar_init(&recv_bar, 1);
        fence_barrier_init();
        if (ks > 1) mbar_expect_tx(&recv_bar, (uint32_t)(ks - 1) * (TC_BM / ks) * Cfg::PITCH * 4);
    }
    if (warp == 1) tmem_alloc<BN>(&tmem_slot);
    tc_fence_before();
    if (ks > 1) cluster_sync_all();   // peers complete_tx on recv_bar: it must be initialised cluster-wide first
    else __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;
    pdl_wait();      // everything above (barriers, TMEM, tensor-map prefetch) overlapped the previous kernel's tail
    if (threadIdx.x == 0) TS(1);

    if (warp == 0) {
        {
            // ===== TMA producer: the whole warp walks the ring (uniform registers), one elected lane issues =====
            const int per_pair = a.taps * kc_per;
            int pr = kc0 / per_pair, rem = kc0 - pr * per_pair;
            int tap = rem / kc_per, c = rem - tap * kc_per;
            int s = 0;
            uint32_t ph = 1;                          // ((i / STAGES) - 1) & 1, kept incrementally
            for (int i = 0; i < nk; i++) {
                const int pa = a.P == 1 ? 0 : c_pair_a[pair_base(a.P) + pr], pb = a.P == 1 ? 0 : c_pair_b[pair_base(a.P) + pr];
                const int off = (tap - (a.taps - 1) / 2) * a.dil;
                if (i >= STAGES) mbar_wait(&empty_bar[s], ph);
                uint8_t* sa = smem + (size_t)s * STAGE_BYTES;
                if (elect_one()) {
                    mbar_expect_tx(&full_bar[s], STAGE_BYTES);
                    if (kch == 1) {
                        tma_load_3d(sa, &map_x, &full_bar[s], c * TC_BK, t0 + off, pb * a.B + b);
                        tma_load_3d(sa + A_BYTES, &map_w, &full_bar[s], c * TC_BK, co0, pa * a.taps + tap);
                    } else {
                        // 4-D maps (64 ch, row, 64-channel chunk, plane): ONE instruction lands kch canonical tiles
                        tma_load_4d(sa, &map_x, &full_bar[s], 0, t0 + off, c * kch, pb * a.B + b);
                        tma_load_4d(sa + kch * A_BYTES, &map_w, &full_bar[s], 0, co0, c * kch, pa * a.taps + tap);
                    }
                }
                __syncwarp();
                if (++c == kc_per) { c = 0; if (++tap == a.taps) { tap = 0; ++pr; } }
                if (++s == STAGES) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp == 1) {
        {
            // ===== MMA issuer: the whole warp walks the pipeline, one elected lane issues =====
            constexpr uint32_t idesc = instr_desc_bf16(TC_BM, BN, /*a_mn=*/0, /*b_mn=*/0);
            // the NEXT stage's barrier is probed (non-blocking) before this stage's MMAs are issued, so the
            // ~190-cycle barrier round trip overlaps the tensor work instead of preceding it
            bool ready = nk > 0 && __all_sync(0xffffffffu, mbar_test(&full_bar[0], 0));
            int s = 0;
            uint32_t ph = 0;                          // (i / STAGES) & 1, kept incrementally
            for (int i = 0; i < nk; i++) {
                if (!ready) mbar_wait(&full_bar[s], ph);
                tc_fence_after();
                if (i == 0 && lane == 0) TS(2);
                int s1 = s + 1;
                uint32_t ph1 = ph;
                if (s1 == STAGES) { s1 = 0; ph1 ^= 1; }
                bool next_ready = false;
                if (i + 1 < nk) next_ready = __all_sync(0xffffffffu, mbar_test(&full_bar[s1], ph1));
                const uint32_t sa = smem_u32(smem + (size_t)s * STAGE_BYTES);
                // K-major operands: 16 K-elements = 32 bytes along the swizzled 128-byte row (+2 in the
                // descriptor's 16-byte address units); 8-row groups are 1024 B apart (SBO); LBO unused
                const uint64_t da = smem_desc_sw128(sa, 16, 1024), db = smem_desc_sw128(sa + kch * A_BYTES, 16, 1024);
                if (elect_one()) {
                    for (int ch = 0; ch < kch; ch++) {
#pragma unroll
                        for (int j = 0; j < TC_BK / 16; j++)
                            umma_bf16(tmem_base, da + (uint64_t)(ch * (A_BYTES / 16) + 2 * j), db + (uint64_t)(ch * (B_BYTES / 16) + 2 * j),
                                      idesc, (i | ch | j) != 0);
                    }
                    umma_commit(&empty_bar[s]);       // frees the smem stage when these MMAs retire
                }
                __syncwarp();
                ready = next_ready;
                s = s1;
                ph = ph1;
            }
            if (elect_one()) umma_commit(&tmem_full_bar);   // accumulator complete (immediately, when nk == 0)
            __syncwarp();
        }
    }

    // ===== exchange.  Each CTA stages its partial 128 x BN accumulator in its OWN shared memory (the TMA ring is
    // idle once tmem_full fires: every load was consumed by an MMA that has retired).  CTA r of the cluster owns
    // rows [128 r / ks, 128 (r+1) / ks) of the tile: one thread sends each peer ITS rows with a bulk
    // shared::cta -> shared::cluster copy (TMA engine, one instruction per peer) that completes on the peer's
    // recv_bar.  Per-thread DSMEM loads / stores managed only ~16 B/cycle/SM here (measured: 8-10k cycles for
    // the same exchange). =====
    constexpr int PITCH = Cfg::PITCH;
    float* stage = reinterpret_cast<float*>(smem);
    const float* recv = reinterpret_cast<const float*>(smem + a.recv_off);
    const int rows_own = TC_BM / ks, r_base = kr * rows_own;
    // full tile, one k-slice, nothing to add from memory: register epilogue + bulk row stores (see epilogue_regs_bulk)
    const bool fast_epi = a.fast_epi && ks == 1 && nk > 0 && !a.residual && !a.accumulate && !a.mask && (a.Cout & 3) == 0 &&
                          co0 + BN <= a.Cout && t0 + TC_BM <= a.T && (!a.gn_sums || ((a.Cout / a.gn_groups) % BN) == 0) &&
                          ((((uintptr_t)a.bias) | ((uintptr_t)a.out) | ((uintptr_t)a.bias_bstride * 4)) & 15) == 0;
    if (fast_epi) {
        mbar_wait_warp(&tmem_full_bar, 0, 32);
        tc_fence_after();
        if (threadIdx.x == 64) TS(3);
        const float* bias = a.bias ? a.bias + (size_t)b * a.bias_bstride + co0 : nullptr;
        double* gp = a.gn_sums ? a.gn_sums + ((size_t)b * a.gn_groups + co0 / (a.Cout / a.gn_groups)) * 2 : nullptr;
        epilogue_regs_bulk<BN>(tmem_base, stage, bias, a.out + ((size_t)b * a.T + t0) * a.Cout + co0, a.Cout, gp, warp, lane);
        if (threadIdx.x == 64) TS(6);
    } else {
        // all 8 warps: TMEM lane quarter = warp % 4 (hardware rule), column half = warp / 4
        const int q = warp & 3, h = warp >> 2;
        mbar_wait_warp(&tmem_full_bar, 0, 32);        // (lanes 1..31 of warps 0/1 arrive here while lane 0 still issues)
        tc_fence_after();
        if (threadIdx.x == 64) TS(3);
        float* srow = stage + (size_t)(q * 32 + lane) * PITCH;
#pragma unroll 1
        for (int c0 = h * (BN / 2); c0 < (h + 1) * (BN / 2); c0 += 32) {
            float v[32];
            if (nk > 0) {
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);     // warp-collective
            } else {
#pragma unroll
                for (int i = 0; i < 32; i++) v[i] = 0.f;                                 // empty k-slice
            }
#pragma unroll
            for (int i = 0; i < 32; i += 4)
                *reinterpret_cast<float4*>(srow + c0 + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
        }
        tc_fence_before();
        fence_proxy_async_smem();                     // generic-proxy stores -> visible to the bulk-copy engine
        __syncthreads();
        if (threadIdx.x == 64) {
            TS(4);
            const uint32_t slot_bytes = (uint32_t)rows_own * PITCH * 4;
            for (int p = 0; p < ks; p++) {
                if (p == kr) continue;
                const int slot = kr < p ? kr : kr - 1;                                   // my slot at peer p
                bulk_copy_to_cta(map_to_cta(smem_u32(recv) + (uint32_t)slot * slot_bytes, p),
                                 smem_u32(stage + (size_t)p * rows_own * PITCH), slot_bytes,
                                 map_to_cta(smem_u32(&recv_bar), p));
            }
        }
        if (ks > 1) mbar_wait_warp(&recv_bar, 0, 32);
        if (threadIdx.x == 64) TS(5);

        // ===== reduction over the k-slices (fixed order: sender rank) + fused epilogue, coalesced: 4 consecutive
        // channels per thread, BN/4 threads per row, two rows per iteration with their loads issued together.
        // Kept small on purpose: this code runs once per CTA out of a cold instruction cache (a fully unrolled
        // version spent 2/3 of its time fetching itself), and with two warps per scheduler nothing else hides
        // a dependent load.
        constexpr int TPR = BN / 4;                   // threads per output row
        constexpr int RPP = TC_THREADS / TPR;         // rows per pass of the 256 threads
        const int tid = threadIdx.x;
        const int cv = (tid % TPR) * 4, rsub = tid / TPR;
        const int co = co0 + cv;
        const int Cout = a.Cout, Tn = a.T;
        const bool vec_ok = (Cout & 3) == 0;
        const bool col_ok = co < Cout;
        const float* __restrict__ resid = a.residual;
        const float* __restrict__ maskp = a.mask;
        const int mact = a.mask_act;
        const float mscale = a.mask_scale;
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
        // shared::cta addresses of this thread's column in row 0 of every sender's block of MY rows
        const uint32_t own_base = smem_u32(stage + (size_t)r_base * PITCH + cv);
        const uint32_t recv_base = smem_u32(recv + cv);
        const uint32_t slot_stride = (uint32_t)rows_own * PITCH * 4;
        // fused GroupNorm statistics of the FINAL values: a thread's 4 channels share a group (host checks
        // (Cout / groups) % 32 == 0); fp32 over the thread's <= 32 rows, fp64 from there on
        const int cg = a.gn_sums ? Cout / a.gn_groups : 1;
        const int g = (a.gn_sums && col_ok) ? co / cg : -1;
        float fs0 = 0.f, fs1 = 0.f;
        const size_t row0 = (size_t)b * Tn + t0 + r_base;
        // full tiles (every tile of the training shapes): no per-element predicates, ~20 instructions per row.  With
        // 2 warps per scheduler this loop is issue-bound; the generic loop below costs ~70 instructions per row
        // (measured 4.4k cycles for a 128 x 128 tile, against ~1.1k for the TMEM drain before it)
        const bool full = vec_ok && co0 + BN <= Cout && t0 + TC_BM <= Tn && (rows_own % RPP) == 0;
        if (full) {
            const int iters = rows_own / RPP;
            const size_t o0 = (row0 + rsub) * Cout + co;
            const size_t ostep = (size_t)RPP * Cout;
            const uint32_t rstep = RPP * PITCH * 4;
            const uint32_t own0 = own_base + (uint32_t)rsub * (PITCH * 4), recv0 = recv_base + (uint32_t)rsub * (PITCH * 4);
            if (ks == 1 && !resid && !accum && !maskp) {
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
                    float4 v = bv;
                    float4 r1 = make_float4(0.f, 0.f, 0.f, 0.f), u1 = r1;
                    float4 m1 = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (resid) r1 = *reinterpret_cast<const float4*>(resid + o0 + it * ostep);
                    if (accum) u1 = *reinterpret_cast<const float4*>(outp + o0 + it * ostep);
                    if (maskp) m1 = *reinterpret_cast<const float4*>(maskp + o0 + it * ostep);
#pragma unroll 1
                    for (int p = 0; p < ks; p++) {
                        const uint32_t base = p == kr ? own0 : recv0 + (uint32_t)(p < kr ? p : p - 1) * slot_stride;
                        const float4 x = ld_shared_f4(base + (uint32_t)it * rstep);
                        v.x += x.x; v.y += x.y; v.z += x.z; v.w += x.w;
                    }
                    if (maskp) {
                        v.x *= mscale * act_grad(m1.x * mscale, mact); v.y *= mscale * act_grad(m1.y * mscale, mact);
                        v.z *= mscale * act_grad(m1.z * mscale, mact); v.w *= mscale * act_grad(m1.w * mscale, mact);
                    }
                    v.x += r1.x + u1.x; v.y += r1.y + u1.y; v.z += r1.z + u1.z; v.w += r1.w + u1.w;
                    *reinterpret_cast<float4*>(outp + o0 + it * ostep) = v;
                    fs0 += (v.x + v.y) + (v.z + v.w);
                    fs1 += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
                }
            }
        } else
#pragma unroll 1
        for (int rl = rsub; rl < rows_own; rl += 2 * RPP) {
            const int rl2 = rl + RPP;
            const bool ok1 = col_ok && t0 + r_base + rl < Tn;
            const bool ok2 = col_ok && rl2 < rows_own && t0 + r_base + rl2 < Tn;
            float4 acc1 = bv, acc2 = bv;
#pragma unroll 1
            for (int p = 0; p < ks; p++) {
                const uint32_t base = p == kr ? own_base : recv_base + (uint32_t)(p < kr ? p : p - 1) * slot_stride;
                const float4 x1 = ok1 ? ld_shared_f4(base + (uint32_t)rl * (PITCH * 4)) : make_float4(0.f, 0.f, 0.f, 0.f);
                const float4 x2 = ok2 ? ld_shared_f4(base + (uint32_t)rl2 * (PITCH * 4)) : make_float4(0.f, 0.f, 0.f, 0.f);
                acc1.x += x1.x; acc1.y += x1.y; acc1.z += x1.z; acc1.w += x1.w;
                acc2.x += x2.x; acc2.y += x2.y; acc2.z += x2.z; acc2.w += x2.w;
            }
            const size_t o1 = (row0 + rl) * Cout + co, o2 = (row0 + rl2) * Cout + co;
            if (vec_ok) {
                float4 r1 = make_float4(0.f, 0.f, 0.f, 0.f), r2 = r1, u1 = r1, u2 = r1;
                if (resid && ok1) r1 = *reinterpret_cast<const float4*>(resid + o1);
                if (resid && ok2) r2 = *reinterpret_cast<const float4*>(resid + o2);
                if (accum && ok1) u1 = *reinterpret_cast<const float4*>(outp + o1);
                if (accum && ok2) u2 = *reinterpret_cast<const float4*>(outp + o2);
                if (maskp) {
                    float4 m1 = make_float4(0.f, 0.f, 0.f, 0.f), m2 = m1;
                    if (ok1) m1 = *reinterpret_cast<const float4*>(maskp + o1);
                    if (ok2) m2 = *reinterpret_cast<const float4*>(maskp + o2);
                    acc1.x *= mscale * act_grad(m1.x * mscale, mact); acc1.y *= mscale * act_grad(m1.y * mscale, mact);
                    acc1.z *= mscale * act_grad(m1.z * mscale, mact); acc1.w *= mscale * act_grad(m1.w * mscale, mact);
                    acc2.x *= mscale * act_grad(m2.x * mscale, mact); acc2.y *= mscale * act_grad(m2.y * mscale, mact);
                    acc2.z *= mscale * act_grad(m2.z * mscale, mact); acc2.w *= mscale * act_grad(m2.w * mscale, mact);
                }
                acc1.x += r1.x + u1.x; acc1.y += r1.y + u1.y; acc1.z += r1.z + u1.z; acc1.w += r1.w + u1.w;
                acc2.x += r2.x + u2.x; acc2.y += r2.y + u2.y; acc2.z += r2.z + u2.z; acc2.w += r2.w + u2.w;
                if (ok1) {
                    if (a.dbg != 31) *reinterpret_cast<float4*>(outp + o1) = acc1;
                    fs0 += (acc1.x + acc1.y) + (acc1.z + acc1.w);
                    fs1 += (acc1.x * acc1.x + acc1.y * acc1.y) + (acc1.z * acc1.z + acc1.w * acc1.w);
                }
                if (ok2) {
                    if (a.dbg != 31) *reinterpret_cast<float4*>(outp + o2) = acc2;
                    fs0 += (acc2.x + acc2.y) + (acc2.z + acc2.w);
                    fs1 += (acc2.x * acc2.x + acc2.y * acc2.y) + (acc2.z * acc2.z + acc2.w * acc2.w);
                }
            } else {
#pragma unroll 1
                for (int u = 0; u < 2; u++) {
                    if (!(u ? ok2 : ok1)) continue;
                    const float4 acc = u ? acc2 : acc1;
                    const size_t o = u ? o2 : o1;
                    const float vals[4] = {acc.x, acc.y, acc.z, acc.w};
#pragma unroll
                    for (int i = 0; i < 4; i++) {
                        if (co + i < Cout) {
                            float x = vals[i];
                            if (maskp) x *= mscale * act_grad(maskp[o + i] * mscale, mact);
                            if (resid) x += resid[o + i];
                            if (accum) x += outp[o + i];
                            outp[o + i] = x;
                            fs0 += x;
                            fs1 += x * x;
                        }
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
        if (threadIdx.x == 64) TS(6);
    }
    // a CTA's outgoing copies read ITS shared memory: nobody may exit before every receiver has its rows
    if (ks > 1) cluster_sync_all();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<BN>(tmem_base);
    }
    if (threadIdx.x == 32) TS(7);
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

bool make_tmap(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
               const uint32_t* box) {
    EncodeFn enc = get_encode();
    if (!enc) {
        set_error("cuTensorMapEncodeTiled driver entry point not available");
        return false;
    }
    cuuint64_t d[5], st[4];
    cuuint32_t bx[5], estr[5] = {1, 1, 1, 1, 1};
    for (int i = 0; i < rank; i++) {
        d[i] = dims[i];
        bx[i] = box[i];
        if (i > 0) st[i - 1] = strides_bytes[i - 1];
    }
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), d, st, bx, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (%d): rank %d dims %llu %llu %llu box %u %u %u", (int)r, rank,
                  (unsigned long long)dims[0], (unsigned long long)dims[1], (unsigned long long)(rank > 2 ? dims[2] : 0),
                  box[0], box[1], rank > 2 ? box[2] : 0);
        return false;
    }
    return true;
}

bool make_tmap_3d(CUtensorMap* out, const void* base, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t stride1_bytes,
                  uint64_t stride2_bytes, uint32_t box0, uint32_t box1, uint32_t box2) {
    const uint64_t dims[3] = {d0, d1, d2}, strides[2] = {stride1_bytes, stride2_bytes};
    const uint32_t box[3] = {box0, box1, box2};
    return make_tmap(out, base, 3, dims, strides, box);
}

}  // namespace tc

template <int BN>
static int launch_conv_tc(const void* in, const void* w, int Cin_p, int ksplit, ConvTcArgs a, cudaStream_t st) {
    using Cfg = ConvTcCfg<BN>;
    constexpr size_t smem_max = 227 * 1024;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(conv_fwd_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_max - 1024);
        if (e != cudaSuccess) {
            set_error("conv_fwd_tc: cudaFuncSetAttribute(%zu): %s", smem_max, cudaGetErrorString(e));
            return VNB_E_CUDA;
        }
        attr_done = true;
    }
    CUtensorMap mx, mw;
    if (a.kch == 1) {
        // activations (P, B, T, Cin_p): dims (ci, t, plane*b); box 64 x 128 x 1
        if (!make_tmap_3d(&mx, in, (uint64_t)a.Cin, (uint64_t)a.T, (uint64_t)a.P * a.B, (uint64_t)Cin_p * 2,
                          (uint64_t)a.T * Cin_p * 2, TC_BK, TC_BM, 1))
            return VNB_E_CUDA;
        // weights (P, taps, Cout, Cin_p): dims (ci, co, plane*tap); box 64 x BN x 1
        if (!make_tmap_3d(&mw, w, (uint64_t)a.Cin, (uint64_t)a.Cout, (uint64_t)a.P * a.taps, (uint64_t)Cin_p * 2,
                          (uint64_t)a.Cout * Cin_p * 2, TC_BK, BN, 1))
            return VNB_E_CUDA;
    } else {
        // Cin % 128 == 0: dims (64 ch, row, 64-channel chunk, plane) - the chunk dimension has the SMALLER stride
        const uint64_t dx[4] = {64, (uint64_t)a.T, (uint64_t)a.Cin / 64, (uint64_t)a.P * a.B};
        const uint64_t sx[3] = {(uint64_t)Cin_p * 2, 128, (uint64_t)a.T * Cin_p * 2};
        const uint32_t bx[4] = {64, TC_BM, (uint32_t)a.kch, 1};
        if (!make_tmap(&mx, in, 4, dx, sx, bx)) return VNB_E_CUDA;
        const uint64_t dw[4] = {64, (uint64_t)a.Cout, (uint64_t)a.Cin / 64, (uint64_t)a.P * a.taps};
        const uint64_t sw[3] = {(uint64_t)Cin_p * 2, 128, (uint64_t)a.Cout * Cin_p * 2};
        const uint32_t bw[4] = {64, BN, (uint32_t)a.kch, 1};
        if (!make_tmap(&mw, w, 4, dw, sw, bw)) return VNB_E_CUDA;
    }
    a.nstages = Cfg::stages(ksplit, a.kch);
    a.recv_off = Cfg::recv_offset(ksplit, a.kch);
    const size_t smem = (size_t)a.recv_off + Cfg::recv_bytes(ksplit) + 1024;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(ksplit, ceil_div(a.Cout, BN), a.B * ceil_div(a.T, TC_BM));
    cfg.blockDim = dim3(TC_THREADS);
    cfg.dynamicSmemBytes = smem;
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
    cudaError_t e = cudaLaunchKernelEx(&cfg, conv_fwd_tc_kernel<BN>, mx, mw, a);
    if (e != cudaSuccess) {
        set_error("conv_fwd_tc: launch failed (BN %d, cluster %d): %s", BN, ksplit, cudaGetErrorString(e));
        return VNB_E_CUDA;
    }
    count_launch();
    return VNB_OK;
}

}  // namespace vnb
#include "tc_conv3.cuh"
namespace vnb {

int conv_fwd_tc(const void* in, int P, int B, int Cin, int T, const void* w, int taps, int dil, int Cout,
                const float* bias, int bias_bstride, const float* residual, float* out, int accumulate,
                double* gn_sums, int gn_groups, cudaStream_t st, const float* mask, int mask_act, float mask_scale) {
    // shapes this backend takes; everything else goes to the SIMT backend (or errors under VNB_BACKEND_TC)
    const int64_t mt = (int64_t)ceil_div(T, TC_BM) * B;
    if (Cout < 8 || Cin < 8 || mt > 65535) return 1;
    if (gn_sums && ((Cout / gn_groups) & 31) != 0) return 1;      // a thread's 4 channels must not straddle groups
    const int dbg = option(VNB_OPT_TC_DBG);
    static const int force_bn = getenv("VNB_TC_BN") ? atoi(getenv("VNB_TC_BN")) : 0;
    static const int force_ks = getenv("VNB_TC_KSPLIT") ? atoi(getenv("VNB_TC_KSPLIT")) : 0;
    ConvTcArgs a{B, Cin, Cout, T, taps, dil, P, bias, bias_bstride, residual, out, accumulate, gn_sums, gn_groups, dbg, 0, 1, 0,
                 option(VNB_OPT_FAST_EPI), mask, mask_act, mask_scale};
    const int Cin_p = round_up8(Cin);
    // N tile and cluster split-K: as many CTAs as fit the 148 SMs in one wave (a tcgen05.mma issue costs the same
    // ~70 cycles for a 32-cycle N=64 tile as for a 128-cycle N=256 one, so wide tiles are preferred on a tie),
    // with the receive slots of the split-K exchange fitting beside the TMA ring
    static const int force_kch = getenv("VNB_TC_KCH") ? atoi(getenv("VNB_TC_KCH")) : 0;
    // two 64-deep chunks per pipeline stage when the channel count allows: one barrier round trip and two TMA
    // instructions per 8 MMAs instead of per 4
    int kch = (Cin % 128 == 0) ? 2 : 1;
    if (force_kch == 1) kch = 1;
    a.kch = kch;
    const int nkc = n_pairs(P) * taps * (ceil_div(Cin, TC_BK) / kch);
    auto fits = [kch](int bn_, int ks_) {
        const uint32_t pitch = bn_ + 4, stage_b = (TC_BM * TC_BK * 2 + bn_ * TC_BK * 2) * kch;
        const uint32_t recv = (uint32_t)(ks_ - 1) * (TC_BM / ks_) * pitch * 4, budget = 227 * 1024 - 2048;
        if (recv + 2 * stage_b > budget) return false;
        uint32_t ring = (budget - recv) / stage_b;
        ring = (ring > 8 ? 8 : ring) * stage_b;
        const uint32_t staging = TC_BM * pitch * 4;
        const uint32_t off = ((ring > staging ? ring : staging) + 1023u) & ~1023u;
        return off + recv + 1024 <= 227 * 1024 - 1024;
    };
    // cost model in SM cycles, fitted to the VNB_TC_DBG=9 timelines of the final kernel (profiles/
    // r01_conv_tc_timeline.txt, sweep over BN x ksplit).  With the converged issuing warp the k-loop is bound by what
    // one SM can take in from L2 (~65-75 B/cycle): per 64-channel chunk ~370 cycles at BN=64 (24 KB), ~430 at BN=128
    // (32 KB), ~690 at BN=256; ~20 % more when a stage holds one chunk.  Fixed part (launch, TMEM, first load,
    // TMEM drain, store): 4.7k / 6.3k cycles for BN = 64 / 128, and ~4k more when a cluster exchanges partials.
    int bn = 64, ks = 1;
    int64_t best = INT64_MAX;
    for (int cand = 64; cand <= 256; cand *= 2) {
        if (cand > 64 && Cout <= cand / 2) break;                 // tile wider than needed
        const int ci = cand == 64 ? 0 : cand == 128 ? 1 : 2;
        static const int64_t per_chunk[2][3] = {{450, 530, 770}, {372, 429, 690}};
        static const int64_t fixed[3][4] = {{4700, 8600, 8900, 9500}, {6300, 10600, 11600, 12400}, {9000, 14000, 15000, 16000}};
        const int64_t tiles = mt * ceil_div(Cout, cand);
        for (int k = 1, ki = 0; k <= 8; k *= 2, ki++) {
            if (k > nkc || !fits(cand, k)) break;
            const int64_t waves = (tiles * k + 147) / 148;
            const int64_t cost = waves * ((int64_t)ceil_div(nkc, k) * kch * per_chunk[kch == 1 ? 0 : 1][ci] + fixed[ci][ki]);
            if (cost < best) { best = cost; bn = cand; ks = k; }
        }
    }
    if (!mask) {
        // third-generation kernel (CTA pairs, all operand planes of a chunk per stage) unless the round-1 kernel's
        // cluster split-K is modelled faster for this shape (its epilogue has no activation-backward mask)
        const int r3 = conv_fwd_tc3(in, P, B, Cin, T, w, taps, dil, Cout, bias, bias_bstride, residual, out, accumulate, gn_sums,
                                    gn_groups, (force_bn || force_ks) ? 0 : best, st);
        if (r3 <= 0) return r3;
    }
    if (force_bn) bn = force_bn;
    if (force_ks) ks = force_ks;
    if (!fits(bn, ks)) { set_error("conv_fwd_tc: BN %d with cluster %d does not fit shared memory", bn, ks); return VNB_E_ARG; }
    if (bn == 256) return launch_conv_tc<256>(in, w, Cin_p, ks, a, st);
    if (bn == 128) return launch_conv_tc<128>(in, w, Cin_p, ks, a, st);
    return launch_conv_tc<64>(in, w, Cin_p, ks, a, st);
}

}  // namespace vnb
// debug: number of bounded mbarrier waits that timed out since load (synchronises the device)
extern "C" int vnb_debug_tc_timeline(long long* host_out128) {
    return cudaMemcpyFromSymbol(host_out128, vnb::g_tc_ts, sizeof(long long) * 128) == cudaSuccess ? 0 : -2;
}
extern "C" int vnb_debug_tc_timeouts(void) {
    int v = -1;
    if (cudaMemcpyFromSymbol(&v, vnb::tc::g_tc_timeouts, sizeof(int)) != cudaSuccess) return -1;
    return v;
}
#include "tc_wgrad.cuh"
#include "tc_vq.cuh"
#include "tc_vq_fused.cuh"
