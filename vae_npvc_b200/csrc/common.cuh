// Shared helpers for the vae_npvc_b200 kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/vae_npvc_b200.h"

namespace vnb {

void set_error(const char* fmt, ...);
void count_launch(int n = 1);
int backend();
int option(int which);
void set_option(int which, int value);
// compiled defaults of the runtime options (environment variables VNB_EW_V2 / VNB_FAST_GATE and vnb_set_option
// override them).  Both on: the second generation passed the GPU parity suite and is 1.3-1.8x faster at large
// frame counts (profiles/r01_kernels_ew_v1_vs_v2.txt).
#ifndef VNB_EW_V2_DEFAULT
#define VNB_EW_V2_DEFAULT 1
#endif
#ifndef VNB_WN_V2_DEFAULT
#define VNB_WN_V2_DEFAULT 1
#endif
#ifndef VNB_FAST_GATE_DEFAULT
#define VNB_FAST_GATE_DEFAULT 1
#endif

#define VNB_REQUIRE(cond, ...)              \
    do {                                    \
        if (!(cond)) {                      \
            vnb::set_error(__VA_ARGS__);    \
            return VNB_E_ARG;               \
        }                                   \
    } while (0)

// Check the launch itself (not completion - never synchronise inside the library).
#define VNB_LAUNCH_CHECK(name)                                                      \
    do {                                                                            \
        cudaError_t e_ = cudaGetLastError();                                        \
        if (e_ != cudaSuccess) {                                                    \
            vnb::set_error("%s: launch failed: %s", name, cudaGetErrorString(e_));  \
            return VNB_E_CUDA;                                                      \
        }                                                                           \
        vnb::count_launch();                                                        \
    } while (0)

// ---- programmatic dependent launch (PDL) ---------------------------------------------------------------
// Every kernel of the library is launched with cudaLaunchAttributeProgrammaticStreamSerialization (captured as a
// programmatic edge inside the trainer's CUDA graph) and starts with pdl_wait(): the next kernel's launch,
// block scheduling and - for the tcgen05 kernels - barrier / TMEM / tensor-map prologue overlap the tail of the
// previous one; pdl_wait() returns once every prerequisite grid has completed and flushed its memory, so data
// dependences (RAW and WAR) are exactly those of plain stream order.  VNB_PDL=0 switches the attribute off.
bool pdl_enabled();
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() {
    asm volatile("griddepcontrol.launch_dependents;\n\tgriddepcontrol.wait;" ::: "memory");
}
#endif

template <typename... KArgs, typename... Args>
static inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                   Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

static inline int ceil_div(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }
static inline int round_up8(int x) { return (x + 7) & ~7; }

typedef __nv_bfloat16 bf16;

// hi/mid/lo split of an fp32 value into bf16 planes: x == hi + mid + lo exactly (3 x 8
// significant bits) unless the tail underflows.
__device__ __forceinline__ void split3(float x, bf16& hi, bf16& mid, bf16& lo) {
    hi = __float2bfloat16_rn(x);
    float r = x - __bfloat162float(hi);
    mid = __float2bfloat16_rn(r);
    r = r - __bfloat162float(mid);
    lo = __float2bfloat16_rn(r);
}

// Operand modes: P = 1 one bf16 plane (8 significant bits); P = 2 hi + mid (16 bits: the GEMMs contract hi*hi +
// hi*mid + mid*hi, ~2^-16 relative); P = 3 hi + mid + lo (24 bits, 6 products: fp32-faithful).  The planes of P = 2
// are the first two of P = 3.
// Write value x at element offset `off` of plane 0 (and planes 1[,2] at +stride each for P >= 2).
__device__ __forceinline__ void store_planes(bf16* base, int64_t off, int64_t plane_stride, int P, float x) {
    if (P == 1) {
        base[off] = __float2bfloat16_rn(x);
    } else {
        bf16 a, b, c;
        split3(x, a, b, c);
        base[off] = a;
        base[off + plane_stride] = b;
        if (P == 3) base[off + 2 * plane_stride] = c;
    }
}

// ---- VEC-wide (1 or 4 channels) accessors for the channels-last kernels --------------------
template <int VEC>
__device__ __forceinline__ void ld_vec(const float* __restrict__ p, float (&v)[VEC]) {
    if constexpr (VEC == 4) {
        const float4 t = *reinterpret_cast<const float4*>(p);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    } else {
#pragma unroll
        for (int j = 0; j < VEC; j++) v[j] = p[j];
    }
}
template <int VEC>
__device__ __forceinline__ void st_vec(float* __restrict__ p, const float (&v)[VEC]) {
    if constexpr (VEC == 4) {
        *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    } else {
#pragma unroll
        for (int j = 0; j < VEC; j++) p[j] = v[j];
    }
}
// VEC consecutive channels into plane 0 (and 1,2 for P==3) at element offset `off` (8-byte aligned for VEC==4)
template <int VEC>
__device__ __forceinline__ void st_planes_vec(bf16* base, int64_t off, int64_t plane_stride, int P, const float (&x)[VEC]) {
    if constexpr (VEC == 4) {
        bf16 h[3][4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            if (P == 1) h[0][j] = __float2bfloat16_rn(x[j]);
            else split3(x[j], h[0][j], h[1][j], h[2][j]);
        }
        for (int p = 0; p < P; p++) {
            uint2 u;
            u.x = (uint32_t)__bfloat16_as_ushort(h[p][0]) | ((uint32_t)__bfloat16_as_ushort(h[p][1]) << 16);
            u.y = (uint32_t)__bfloat16_as_ushort(h[p][2]) | ((uint32_t)__bfloat16_as_ushort(h[p][3]) << 16);
            *reinterpret_cast<uint2*>(base + off + p * plane_stride) = u;
        }
    } else {
#pragma unroll
        for (int j = 0; j < VEC; j++) store_planes(base, off + j, plane_stride, P, x[j]);
    }
}

__device__ __forceinline__ float load_planes(const bf16* base, int64_t off, int64_t plane_stride, int P) {
    float x = __bfloat162float(base[off]);
    if (P >= 2) x += __bfloat162float(base[off + plane_stride]);
    if (P == 3) x += __bfloat162float(base[off + 2 * plane_stride]);
    return x;
}

__device__ __forceinline__ float lrelu02(float x) { return x >= 0.f ? x : 0.2f * x; }
__device__ __forceinline__ float apply_act(float x, int act) {
    if (act == VNB_ACT_LRELU) return lrelu02(x);
    if (act == VNB_ACT_RELU) return x > 0.f ? x : 0.f;
    return x;
}
__device__ __forceinline__ float act_grad(float x, int act) {
    if (act == VNB_ACT_LRELU) return x >= 0.f ? 1.f : 0.2f;
    if (act == VNB_ACT_RELU) return x > 0.f ? 1.f : 0.f;
    return 1.f;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Block-wide sum; result valid in every thread.  `red` needs 32 floats of shared memory.
__device__ __forceinline__ float block_sum(float v, float* red) {
    v = warp_sum(v);
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) red[w] = v;
    __syncthreads();
    int nw = (blockDim.x + 31) >> 5;
    float r = (threadIdx.x < nw) ? red[threadIdx.x] : 0.f;
    if (w == 0) r = warp_sum(r);
    if (threadIdx.x == 0) red[0] = r;
    __syncthreads();
    return red[0];
}

}  // namespace vnb
