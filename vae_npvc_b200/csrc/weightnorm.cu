// Weight-norm materialisation (g, v) -> GEMM operand planes, and its backward.
// Replaces the nn.utils.weight_norm pre-hook the reference installs on all 55 conv modules
// (vae_npvc/model/vqvae.py:203-208, 329-334).  SURVEY.md section 9 G; F8 for ConvTranspose.
#include "common.cuh"

namespace vnb {

// One block per index of dim 0.  v row = d1*k contiguous floats.
__global__ void __launch_bounds__(256) wn_prepare_kernel(const float* __restrict__ v, const float* __restrict__ g,
                                                         int d0, int d1, int k, int transposed, int P,
                                                         bf16* __restrict__ wf, bf16* __restrict__ wb,
                                                         float* __restrict__ w32) {
    pdl_wait();
    __shared__ float red[32];
    const int i = blockIdx.x;
    const int n = d1 * k;
    const float* row = v + (int64_t)i * n;
    float scale = 1.f;
    if (g != nullptr) {
        float ss = 0.f;
        for (int e = threadIdx.x; e < n; e += blockDim.x) {
            float x = row[e];
            ss += x * x;
        }
        ss = block_sum(ss, red);
        scale = g[i] / sqrtf(ss);
    }
    const int co_n = transposed ? d1 : d0;
    const int ci_n = transposed ? d0 : d1;
    const int ci_p = (ci_n + 7) & ~7, co_p = (co_n + 7) & ~7;
    const int64_t wf_plane = (int64_t)k * co_n * ci_p;
    const int64_t wb_plane = (int64_t)k * ci_n * co_p;
    for (int e = threadIdx.x; e < n; e += blockDim.x) {
        const int j1 = e / k, j = e - j1 * k;
        const float w = row[e] * scale;
        if (w32) w32[(int64_t)i * n + e] = w;
        const int co = transposed ? j1 : i;
        const int ci = transposed ? i : j1;
        const int tap = transposed ? (k - 1 - j) : j;      // forward conv tap
        if (wf) store_planes(wf, ((int64_t)tap * co_n + co) * ci_p + ci, wf_plane, P, w);
        if (wb) store_planes(wb, ((int64_t)(k - 1 - tap) * ci_n + ci) * co_p + co, wb_plane, P, w);
    }
}

__global__ void __launch_bounds__(256) wn_backward_kernel(const float* __restrict__ dw, const float* __restrict__ v,
                                                          const float* __restrict__ g, int n,
                                                          float* __restrict__ dv, float* __restrict__ dg) {
    pdl_wait();
    __shared__ float red[32];
    const int i = blockIdx.x;
    const float* vr = v + (int64_t)i * n;
    const float* dr = dw + (int64_t)i * n;
    float ss = 0.f, dot = 0.f;
    for (int e = threadIdx.x; e < n; e += blockDim.x) {
        float x = vr[e];
        ss += x * x;
        dot += x * dr[e];
    }
    ss = block_sum(ss, red);
    dot = block_sum(dot, red);
    const float norm = sqrtf(ss);
    const float gi = g[i];
    if (threadIdx.x == 0) dg[i] = dot / norm;
    const float a = gi / norm, b = gi * dot / (norm * ss);
    for (int e = threadIdx.x; e < n; e += blockDim.x) dv[(int64_t)i * n + e] = a * dr[e] - b * vr[e];
}

// ---- batched variants over the trainer's flat parameter / gradient buffers -------------------
// desc: int64[n_layers][WN_DESC] = {v_off, g_off (-1: plain weight), d0, d1, k, transposed, wf_off, wb_off,
// w32_off (-1: not wanted), row0, tile0, 0};  row2layer: int32[total_rows];  tile2layer: int32[total_tiles]
// (a tile = 64 rows x 64 columns x k taps of one layer's (d0, d1, k) parameter).
constexpr int WN_DESC = 12;
constexpr int WN_TILE = 64;

// scale[row] = g / ||v_row||  (1 for plain weights): one warp per weight row of ANY layer
__global__ void __launch_bounds__(256) wn_rowscale_kernel(const float* __restrict__ flat_p, const int64_t* __restrict__ desc,
                                                          const int* __restrict__ row2layer, int total_rows,
                                                          float* __restrict__ scale) {
    pdl_wait();
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= total_rows) return;
    const int64_t* L = desc + (int64_t)row2layer[row] * WN_DESC;
    if (L[1] < 0) {
        if (lane == 0) scale[row] = 1.f;
        return;
    }
    const int i = row - (int)L[9];
    const int n = (int)L[3] * (int)L[4];
    const float* v = flat_p + L[0] + (int64_t)i * n;
    float ss = 0.f;
    if ((((uintptr_t)v) & 15) == 0 && (n & 3) == 0) {
        const float4* v4 = reinterpret_cast<const float4*>(v);
        for (int e = lane; e < n / 4; e += 32) {
            const float4 x = v4[e];
            ss += (x.x * x.x + x.y * x.y) + (x.z * x.z + x.w * x.w);
        }
    } else {
        for (int e = lane; e < n; e += 32) ss += v[e] * v[e];
    }
    ss = warp_sum(ss);
    if (lane == 0) scale[row] = flat_p[L[1] + i] / sqrtf(ss);
}

// One block per 64 x 64 x k tile: coalesced read of v, scaled by the row's g/||v||, staged in shared memory, then
// written twice with 128-byte warp stores (2 bf16 per lane) - once along d1 (the operand whose contraction index
// is d1: wf for Conv1d, wb for ConvTranspose1d) and once TRANSPOSED along d0 (the other operand, taps flipped).
// The row-per-block version scattered 2-byte stores d0 * 2 bytes apart: 364 us for 125 MB in / 125 MB out.
__device__ __forceinline__ void store_planes2(bf16* base, int64_t off, int64_t plane_stride, int P, float x, float y, bool two) {
    if (P == 1 && two && (off & 1) == 0) {
        *reinterpret_cast<__nv_bfloat162*>(base + off) = __floats2bfloat162_rn(x, y);
    } else {
        store_planes(base, off, plane_stride, P, x);
        if (two) store_planes(base, off + 1, plane_stride, P, y);
    }
}

// Fast path of wn_tiles_kernel for a FULL 64 x 64 tile, bf16 operand mode (P == 1), k known at compile time:
// no runtime divisions (the generic path divides by k for every element and by the tile width for every store
// row), every load of the tile issued before the first use, and 8-byte stores (4 bf16 per lane, two tile rows per
// warp instruction) instead of 4-byte ones.  The generic path is issue-bound (1.85 TB/s of traffic for a pass that
// moves 8 bytes per parameter); this one does a third of its instructions.
// pack 4 floats into the P planes' 8-byte words and store them plane_stride apart
template <int P>
__device__ __forceinline__ void wn_store4(bf16* __restrict__ dst, int64_t plane_stride, float a, float b, float c, float d) {
    const float x[4] = {a, b, c, d};
    bf16 h[3][4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        if constexpr (P == 1) h[0][j] = __float2bfloat16_rn(x[j]);
        else split3(x[j], h[0][j], h[1][j], h[2][j]);
    }
#pragma unroll
    for (int p = 0; p < P; p++) {
        uint2 o;
        o.x = (uint32_t)__bfloat16_as_ushort(h[p][0]) | ((uint32_t)__bfloat16_as_ushort(h[p][1]) << 16);
        o.y = (uint32_t)__bfloat16_as_ushort(h[p][2]) | ((uint32_t)__bfloat16_as_ushort(h[p][3]) << 16);
        *reinterpret_cast<uint2*>(dst + p * plane_stride) = o;
    }
}

template <int K, int P>
__device__ __forceinline__ void wn_tile_full_p1(const float* __restrict__ v, float* __restrict__ w32,
                                                const float* __restrict__ scale, int row0, int d0, int d1, int r0, int c0,
                                                bf16* __restrict__ D, bf16* __restrict__ Tt, float* __restrict__ tile) {
    constexpr int TP = WN_TILE + 1;
    constexpr int SEG4 = WN_TILE * K / 4;              // float4 per tile row
    constexpr int PER = SEG4 * WN_TILE / 256;          // float4 per thread: 4 * K
    float4 x[PER];
#pragma unroll
    for (int u = 0; u < PER; u++) {
        const int i = u * 256 + threadIdx.x;
        const int rr = i / SEG4, e4 = i - rr * SEG4;
        x[u] = *reinterpret_cast<const float4*>(v + ((int64_t)(r0 + rr) * d1 + c0) * K + 4 * e4);
    }
#pragma unroll
    for (int u = 0; u < PER; u++) {
        const int i = u * 256 + threadIdx.x;
        const int rr = i / SEG4, e4 = i - rr * SEG4;
        const float sc = scale[row0 + rr];
        const float w[4] = {x[u].x * sc, x[u].y * sc, x[u].z * sc, x[u].w * sc};
        if (w32) *reinterpret_cast<float4*>(w32 + ((int64_t)(r0 + rr) * d1 + c0) * K + 4 * e4) = make_float4(w[0], w[1], w[2], w[3]);
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const int e = 4 * e4 + q;
            const int cc = e / K, j = e - cc * K;
            tile[(j * WN_TILE + rr) * TP + cc] = w[q];
        }
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int half = lane >> 4, l4 = (lane & 15) * 4;
    const int p1 = (d1 + 7) & ~7, p0 = (d0 + 7) & ~7;
    if (D) {                                           // [j][row][col]: a half-warp writes one 64-wide row
#pragma unroll
        for (int it = 0; it < K * WN_TILE / 16; it++) {
            const int idx = it * 16 + warp * 2 + half;
            const int j = idx / WN_TILE, rr = idx - j * WN_TILE;
            const float* src = tile + (j * WN_TILE + rr) * TP + l4;
            wn_store4<P>(D + ((int64_t)j * d0 + r0 + rr) * p1 + c0 + l4, (int64_t)K * d0 * p1, src[0], src[1], src[2], src[3]);
        }
    }
    if (Tt) {                                          // [K-1-j][col][row]: a half-warp writes 64 rows of one column
#pragma unroll
        for (int it = 0; it < K * WN_TILE / 16; it++) {
            const int idx = it * 16 + warp * 2 + half;
            const int j = idx / WN_TILE, cc = idx - j * WN_TILE;
            const float* src = tile + (j * WN_TILE + l4) * TP + cc;
            wn_store4<P>(Tt + ((int64_t)(K - 1 - j) * d1 + c0 + cc) * p0 + r0 + l4, (int64_t)K * d1 * p0, src[0], src[TP], src[2 * TP],
                         src[3 * TP]);
        }
    }
}

__global__ void __launch_bounds__(256, 3) wn_tiles_kernel(const float* __restrict__ flat_p, const int64_t* __restrict__ desc,
                                                       const int* __restrict__ tile2layer, const float* __restrict__ scale,
                                                       int P, bf16* __restrict__ wbuf, float* __restrict__ w32buf, int fast) {
    pdl_wait();
    extern __shared__ float wn_tile[];                // [k][64][65]
    constexpr int TP = WN_TILE + 1;
    const int64_t* L = desc + (int64_t)tile2layer[blockIdx.x] * WN_DESC;
    const int d0 = (int)L[2], d1 = (int)L[3], k = (int)L[4], transposed = (int)L[5];
    const int tcols = (d1 + WN_TILE - 1) / WN_TILE;
    const int t = blockIdx.x - (int)L[10];
    const int r0 = (t / tcols) * WN_TILE, c0 = (t % tcols) * WN_TILE;
    const int nr = d0 - r0 < WN_TILE ? d0 - r0 : WN_TILE, nc = d1 - c0 < WN_TILE ? d1 - c0 : WN_TILE;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float* v = flat_p + L[0];
    float* w32 = L[8] >= 0 ? w32buf + L[8] : nullptr;
    // Loads in batches of 8 per thread, all issued before the first use: the tile comes from HBM, and a loop of
    // load -> scale -> shared store serialised 48 DRAM round trips per warp (225 us for the 250 MB pass).
    const int seg = nc * k;                            // contiguous floats per tile row
    const int total = nr * seg;
    const int row0 = (int)L[9] + r0;
    if (fast && P <= 2 && nr == WN_TILE && nc == WN_TILE && (k == 1 || k == 3) &&
        ((((uintptr_t)(v + ((int64_t)r0 * d1 + c0) * k)) | ((uintptr_t)((int64_t)d1 * k * 4)) |
          (w32 ? (uintptr_t)(w32 + ((int64_t)r0 * d1 + c0) * k) : 0)) & 15) == 0) {
        const int64_t o_d = transposed ? L[7] : L[6], o_t = transposed ? L[6] : L[7];
        bf16* Dp = o_d >= 0 ? wbuf + o_d : nullptr;
        bf16* Tp = o_t >= 0 ? wbuf + o_t : nullptr;
        if (P == 1) {
            if (k == 3) wn_tile_full_p1<3, 1>(v, w32, scale, row0, d0, d1, r0, c0, Dp, Tp, wn_tile);
            else wn_tile_full_p1<1, 1>(v, w32, scale, row0, d0, d1, r0, c0, Dp, Tp, wn_tile);
        } else {
            if (k == 3) wn_tile_full_p1<3, 2>(v, w32, scale, row0, d0, d1, r0, c0, Dp, Tp, wn_tile);
            else wn_tile_full_p1<1, 2>(v, w32, scale, row0, d0, d1, r0, c0, Dp, Tp, wn_tile);
        }
        return;
    }
    if ((seg & 3) == 0 && ((((uintptr_t)(v + ((int64_t)r0 * d1 + c0) * k)) | ((uintptr_t)((int64_t)d1 * k * 4))) & 15) == 0) {
        const int seg4 = seg >> 2, total4 = nr * seg4;
        for (int b0 = 0; b0 < total4; b0 += 256 * 6) {
            float4 x[6];
            int rr[6], e4[6];
#pragma unroll
            for (int u = 0; u < 6; u++) {
                const int i = b0 + u * 256 + threadIdx.x;
                rr[u] = i / seg4;
                e4[u] = i - rr[u] * seg4;
                if (i < total4) x[u] = *reinterpret_cast<const float4*>(v + ((int64_t)(r0 + rr[u]) * d1 + c0) * k + 4 * e4[u]);
            }
#pragma unroll
            for (int u = 0; u < 6; u++) {
                if (b0 + u * 256 + threadIdx.x >= total4) continue;
                const float sc = scale[row0 + rr[u]];
                const float w[4] = {x[u].x * sc, x[u].y * sc, x[u].z * sc, x[u].w * sc};
                const int64_t base = ((int64_t)(r0 + rr[u]) * d1 + c0) * k + 4 * e4[u];
                if (w32) *reinterpret_cast<float4*>(w32 + base) = make_float4(w[0], w[1], w[2], w[3]);
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const int e = 4 * e4[u] + q;
                    const int cc = e / k, j = e - cc * k;
                    wn_tile[(j * WN_TILE + rr[u]) * TP + cc] = w[q];
                }
            }
        }
    } else {
        for (int b0 = 0; b0 < total; b0 += 256 * 8) {
            float x[8];
            int rr[8], ee[8];
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const int i = b0 + u * 256 + threadIdx.x;
                rr[u] = i / seg;
                ee[u] = i - rr[u] * seg;
                if (i < total) x[u] = v[((int64_t)(r0 + rr[u]) * d1 + c0) * k + ee[u]];
            }
#pragma unroll
            for (int u = 0; u < 8; u++) {
                if (b0 + u * 256 + threadIdx.x >= total) continue;
                const float w = x[u] * scale[row0 + rr[u]];
                if (w32) w32[((int64_t)(r0 + rr[u]) * d1 + c0) * k + ee[u]] = w;
                const int cc = ee[u] / k, j = ee[u] - cc * k;
                wn_tile[(j * WN_TILE + rr[u]) * TP + cc] = w;
            }
        }
    }
    __syncthreads();
    // direct: [j][row][col], pitch round8(d1);  transposed: [k-1-j][col][row], pitch round8(d0)
    const int64_t off_d = transposed ? L[7] : L[6], off_t = transposed ? L[6] : L[7];
    const int p1 = (d1 + 7) & ~7, p0 = (d0 + 7) & ~7;
    const int64_t plane_d = (int64_t)k * d0 * p1, plane_t = (int64_t)k * d1 * p0;
    if (off_d >= 0) {
        bf16* D = wbuf + off_d;
        for (int idx = warp; idx < k * nr; idx += 8) {
            const int j = idx / nr, rr = idx - j * nr;
            const int cc = 2 * lane;
            if (cc < nc) {
                const float* src = wn_tile + (j * WN_TILE + rr) * TP + cc;
                store_planes2(D, ((int64_t)j * d0 + r0 + rr) * p1 + c0 + cc, plane_d, P, src[0], cc + 1 < nc ? src[1] : 0.f, cc + 1 < nc);
            }
        }
    }
    if (off_t >= 0) {
        bf16* Tt = wbuf + off_t;
        for (int idx = warp; idx < k * nc; idx += 8) {
            const int j = idx / nc, cc = idx - j * nc;
            const int rr = 2 * lane;
            if (rr < nr) {
                const float* src = wn_tile + (j * WN_TILE + rr) * TP + cc;
                store_planes2(Tt, ((int64_t)(k - 1 - j) * d1 + c0 + cc) * p0 + r0 + rr, plane_t, P, src[0],
                              rr + 1 < nr ? src[TP] : 0.f, rr + 1 < nr);
            }
        }
    }
}

// in place on flat_g: the v slot holds dw on entry, dv on exit; the g slot receives dg
__global__ void __launch_bounds__(256) wn_backward_batch_kernel(const float* __restrict__ flat_p,
                                                                float* __restrict__ flat_g,
                                                                const int64_t* __restrict__ desc,
                                                                const int* __restrict__ row2layer) {
    pdl_wait();
    __shared__ float red[32];
    const int64_t* L = desc + (int64_t)row2layer[blockIdx.x] * WN_DESC;
    if (L[1] < 0) return;                       // plain weight: dw already is its gradient
    const int i = blockIdx.x - (int)L[9];
    const int n = (int)L[3] * (int)L[4];
    const float* vr = flat_p + L[0] + (int64_t)i * n;
    float* dr = flat_g + L[0] + (int64_t)i * n;
    const float gi = flat_p[L[1] + i];
    if ((n & 3) == 0 && n <= 4096 && ((((uintptr_t)vr) | ((uintptr_t)dr)) & 15) == 0) {
        // one pass over global memory: the row (<= 4 float4 per thread of v and of dw) stays in registers
        const float4* v4 = reinterpret_cast<const float4*>(vr);
        float4* d4 = reinterpret_cast<float4*>(dr);
        const int n4 = n >> 2;
        float4 xv[4], xd[4];
        float ss = 0.f, dot = 0.f;
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int e = threadIdx.x + u * 256;
            if (e < n4) {
                xv[u] = v4[e];
                xd[u] = d4[e];
                ss += (xv[u].x * xv[u].x + xv[u].y * xv[u].y) + (xv[u].z * xv[u].z + xv[u].w * xv[u].w);
                dot += (xv[u].x * xd[u].x + xv[u].y * xd[u].y) + (xv[u].z * xd[u].z + xv[u].w * xd[u].w);
            }
        }
        ss = block_sum(ss, red);
        dot = block_sum(dot, red);
        const float norm = sqrtf(ss);
        if (threadIdx.x == 0) flat_g[L[1] + i] = dot / norm;
        const float a = gi / norm, b = gi * dot / (norm * ss);
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int e = threadIdx.x + u * 256;
            if (e < n4)
                d4[e] = make_float4(a * xd[u].x - b * xv[u].x, a * xd[u].y - b * xv[u].y, a * xd[u].z - b * xv[u].z,
                                    a * xd[u].w - b * xv[u].w);
        }
        return;
    }
    float ss = 0.f, dot = 0.f;
    for (int e = threadIdx.x; e < n; e += blockDim.x) {
        const float x = vr[e];
        ss += x * x;
        dot += x * dr[e];
    }
    ss = block_sum(ss, red);
    dot = block_sum(dot, red);
    const float norm = sqrtf(ss);
    if (threadIdx.x == 0) flat_g[L[1] + i] = dot / norm;
    const float a = gi / norm, b = gi * dot / (norm * ss);
    for (int e = threadIdx.x; e < n; e += blockDim.x) dr[e] = a * dr[e] - b * vr[e];
}

// Second generation of the same pass: ONE WARP per weight row, rows taken in a warp-strided loop.
// The block-per-row kernel above keeps one row (<= 24 KB of v and dw) in flight per CTA and spends two block-wide
// reductions (four __syncthreads) between its loads and its stores: at ~4 resident CTAs per SM that is ~50 KB in
// flight per SM with idle phases, 0.42 of the HBM roof.  Here a warp keeps a whole row of up to 1536 floats in
// registers (12 + 12 float4 per lane, all loads issued before the first use: 12 warps x 12 KB = 144 KB in flight
// per SM), reduces with shuffles only and writes dv from the registers; longer rows (the decoder's 3072-float
// ConvTranspose rows) take two passes, the second one re-reading the row from L2.
constexpr int WNB_MAXV = 12;
constexpr int WNB_CH = 8;

__global__ void __launch_bounds__(128, 2) wn_backward_batch_v2_kernel(const float* __restrict__ flat_p,
                                                                      float* __restrict__ flat_g,
                                                                      const int64_t* __restrict__ desc,
                                                                      const int* __restrict__ row2layer, int total_rows) {
    pdl_wait();
    const int lane = threadIdx.x & 31;
    const int nwarps = gridDim.x * 4;
    for (int row = blockIdx.x * 4 + (threadIdx.x >> 5); row < total_rows; row += nwarps) {
        const int64_t* L = desc + (int64_t)row2layer[row] * WN_DESC;
        if (L[1] < 0) continue;                     // plain weight: dw already is its gradient
        const int i = row - (int)L[9];
        const int n = (int)L[3] * (int)L[4];
        const float* vr = flat_p + L[0] + (int64_t)i * n;
        float* dr = flat_g + L[0] + (int64_t)i * n;
        const float gi = flat_p[L[1] + i];
        float ss = 0.f, dot = 0.f;
        const bool vec = (n & 3) == 0 && ((((uintptr_t)vr) | ((uintptr_t)dr)) & 15) == 0;
        if (vec && (n >> 2) <= 32 * WNB_MAXV) {
            const float4* v4 = reinterpret_cast<const float4*>(vr);
            float4* d4 = reinterpret_cast<float4*>(dr);
            const int n4 = n >> 2;
            float4 xv[WNB_MAXV], xd[WNB_MAXV];
#pragma unroll
            for (int u = 0; u < WNB_MAXV; u++) {
                const int e = lane + 32 * u;
                if (e < n4) {
                    xv[u] = v4[e];
                    xd[u] = d4[e];
                }
            }
#pragma unroll
            for (int u = 0; u < WNB_MAXV; u++) {
                if (lane + 32 * u < n4) {
                    ss += (xv[u].x * xv[u].x + xv[u].y * xv[u].y) + (xv[u].z * xv[u].z + xv[u].w * xv[u].w);
                    dot += (xv[u].x * xd[u].x + xv[u].y * xd[u].y) + (xv[u].z * xd[u].z + xv[u].w * xd[u].w);
                }
            }
            ss = warp_sum(ss);
            dot = warp_sum(dot);
            const float norm = sqrtf(ss);
            const float a = gi / norm, b = gi * dot / (norm * ss);
#pragma unroll
            for (int u = 0; u < WNB_MAXV; u++) {
                const int e = lane + 32 * u;
                if (e < n4)
                    d4[e] = make_float4(a * xd[u].x - b * xv[u].x, a * xd[u].y - b * xv[u].y, a * xd[u].z - b * xv[u].z,
                                        a * xd[u].w - b * xv[u].w);
            }
            if (lane == 0) flat_g[L[1] + i] = dot / norm;
        } else if (vec) {
            // longer rows: two passes in chunks of 8 float4 of each tensor in flight per lane; the second pass finds the row in L2
            const float4* v4 = reinterpret_cast<const float4*>(vr);
            float4* d4 = reinterpret_cast<float4*>(dr);
            const int n4 = n >> 2;
            for (int e0 = 0; e0 < n4; e0 += 32 * WNB_CH) {
                float4 xv[WNB_CH], xd[WNB_CH];
#pragma unroll
                for (int u = 0; u < WNB_CH; u++) {
                    const int e = e0 + lane + 32 * u;
                    if (e < n4) {
                        xv[u] = v4[e];
                        xd[u] = d4[e];
                    }
                }
#pragma unroll
                for (int u = 0; u < WNB_CH; u++) {
                    if (e0 + lane + 32 * u < n4) {
                        ss += (xv[u].x * xv[u].x + xv[u].y * xv[u].y) + (xv[u].z * xv[u].z + xv[u].w * xv[u].w);
                        dot += (xv[u].x * xd[u].x + xv[u].y * xd[u].y) + (xv[u].z * xd[u].z + xv[u].w * xd[u].w);
                    }
                }
            }
            ss = warp_sum(ss);
            dot = warp_sum(dot);
            const float norm = sqrtf(ss);
            const float a = gi / norm, b = gi * dot / (norm * ss);
            for (int e0 = 0; e0 < n4; e0 += 32 * WNB_CH) {
                float4 xv[WNB_CH], xd[WNB_CH];
#pragma unroll
                for (int u = 0; u < WNB_CH; u++) {
                    const int e = e0 + lane + 32 * u;
                    if (e < n4) {
                        xv[u] = v4[e];
                        xd[u] = d4[e];
                    }
                }
#pragma unroll
                for (int u = 0; u < WNB_CH; u++) {
                    const int e = e0 + lane + 32 * u;
                    if (e < n4)
                        d4[e] = make_float4(a * xd[u].x - b * xv[u].x, a * xd[u].y - b * xv[u].y,
                                            a * xd[u].z - b * xv[u].z, a * xd[u].w - b * xv[u].w);
                }
            }
            if (lane == 0) flat_g[L[1] + i] = dot / norm;
        } else {
            for (int e = lane; e < n; e += 32) {
                const float x = vr[e];
                ss += x * x;
                dot += x * dr[e];
            }
            ss = warp_sum(ss);
            dot = warp_sum(dot);
            const float norm = sqrtf(ss);
            const float a = gi / norm, b = gi * dot / (norm * ss);
            for (int e = lane; e < n; e += 32) dr[e] = a * dr[e] - b * vr[e];
            if (lane == 0) flat_g[L[1] + i] = dot / norm;
        }
    }
}

}  // namespace vnb

extern "C" int vnb_wn_prepare_batch(const float* flat_p, const int64_t* desc, const int* row2layer, int total_rows,
                                    const int* tile2layer, int total_tiles, int max_k, int P, float* row_scale,
                                    void* wbuf, float* w32buf, void* stream) {
    VNB_REQUIRE(flat_p && desc && row2layer && tile2layer && row_scale && wbuf && total_rows > 0 && total_tiles > 0 &&
                    (P >= 1 && P <= 3), "wn_prepare_batch: bad arguments");
    const size_t smem = (size_t)max_k * vnb::WN_TILE * (vnb::WN_TILE + 1) * sizeof(float);
    VNB_REQUIRE(max_k >= 1 && smem <= 200 * 1024, "wn_prepare_batch: kernel size %d too large for the tile buffer", max_k);
    static size_t smem_set = 0;
    if (smem > smem_set) {
        if (cudaFuncSetAttribute(vnb::wn_tiles_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
            vnb::set_error("wn_prepare_batch: cudaFuncSetAttribute(%zu) failed", smem);
            return VNB_E_CUDA;
        }
        smem_set = smem;
    }
    cudaStream_t st = (cudaStream_t)stream;
    vnb::launch_k(vnb::wn_rowscale_kernel, vnb::ceil_div(total_rows, 8), 256, 0, st, flat_p, desc, row2layer, total_rows, row_scale);
    VNB_LAUNCH_CHECK("wn_rowscale");
    vnb::launch_k(vnb::wn_tiles_kernel, total_tiles, 256, smem, st, flat_p, desc, tile2layer, row_scale, P, (vnb::bf16*)wbuf, w32buf,
                  vnb::option(VNB_OPT_WN_V2));
    VNB_LAUNCH_CHECK("wn_tiles");
    return VNB_OK;
}

extern "C" int vnb_wn_backward_batch(const float* flat_p, float* flat_g, const int64_t* desc, const int* row2layer,
                                     int total_rows, void* stream) {
    VNB_REQUIRE(flat_p && flat_g && desc && row2layer && total_rows > 0, "wn_backward_batch: bad arguments");
    if (vnb::option(VNB_OPT_WN_V2)) {
        int grid = vnb::ceil_div(total_rows, 4);
        if (grid > 148 * 3 * 4) grid = 148 * 3 * 4;           // three resident 4-warp CTAs per SM, several rows per warp
        vnb::launch_k(vnb::wn_backward_batch_v2_kernel, grid, 128, 0, (cudaStream_t)stream, flat_p, flat_g, desc, row2layer,
                      total_rows);
        VNB_LAUNCH_CHECK("wn_backward_batch_v2");
        return VNB_OK;
    }
    vnb::launch_k(vnb::wn_backward_batch_kernel, total_rows, 256, 0, (cudaStream_t)stream, flat_p, flat_g, desc, row2layer);
    VNB_LAUNCH_CHECK("wn_backward_batch");
    return VNB_OK;
}

extern "C" int vnb_wn_prepare(const float* v, const float* g, int d0, int d1, int k, int transposed, int P,
                              void* wf, void* wb, float* w32, void* stream) {
    VNB_REQUIRE(v && d0 > 0 && d1 > 0 && k > 0 && (P >= 1 && P <= 3), "wn_prepare: bad arguments");
    vnb::launch_k(vnb::wn_prepare_kernel, d0, 256, 0, (cudaStream_t)stream, v, g, d0, d1, k, transposed, P, (vnb::bf16*)wf,
                                                                (vnb::bf16*)wb, w32);
    VNB_LAUNCH_CHECK("wn_prepare");
    return VNB_OK;
}

extern "C" int vnb_wn_backward(const float* dw, const float* v, const float* g, int d0, int d1, int k, float* dv,
                               float* dg, void* stream) {
    VNB_REQUIRE(dw && v && g && dv && dg && d0 > 0, "wn_backward: bad arguments");
    vnb::launch_k(vnb::wn_backward_kernel, d0, 256, 0, (cudaStream_t)stream, dw, v, g, d1 * k, dv, dg);
    VNB_LAUNCH_CHECK("wn_backward");
    return VNB_OK;
}
