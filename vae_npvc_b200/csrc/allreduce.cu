// Gradient all-reduce over NVSwitch multicast memory (NVLS): one small kernel per bucket instead of a library
// collective.
//
// Why: the step all-reduces its flat fp32 gradient buffer (31.2 M elements) in buckets while backward is still
// running.  A library all-reduce kernel occupies 24-32 SMs with CTAs whose register / shared-memory footprint
// keeps the conv GEMM CTAs (197 KB of shared memory, one per SM, 128 per launch) off those SMs: the 128-CTA
// GEMM grids then need two waves and everything overlapped with the collective runs at half speed (measured:
// 8 GPUs 2.89 ms/step with bucketed NCCL overlap vs 2.45 ms on one GPU, although the collective alone is
// 0.37 ms).  This kernel uses no shared memory and <= 32 registers x 512 threads per CTA, so a GEMM CTA fits
// beside it on the same SM, and the reduction itself is done by the switch:
//
//   barrier (peers have produced their gradients)
//   rank r owns the r-th 1/world of the bucket:  v = multimem.ld_reduce.add(mc + i)   (switch sums all GPUs)
//                                                multimem.st(mc + i, v * scale)       (switch broadcasts)
//   barrier (everybody's part has landed everywhere)
//
// Every element is summed once, by the switch, and the same bits are broadcast to all ranks, so the replicas
// stay bit-identical.  `scale` folds the 1/world of the gradient mean into the same pass.
//
// The buffer and the flag words live in symmetric memory set up by the host (torch symmetric-memory rendezvous:
// plumbing); `mc_base` is the multicast address of the gradient buffer, `flags[p]` the unicast address of rank
// p's flag words.  Barriers are per CTA index: CTA b of rank r exchanges one flag word with CTA b of every peer
// (CAS 0->1 to signal, CAS 1->0 to consume: self-resetting, so the kernel can be replayed from a CUDA graph).
// A CTA only ever waits for the same-index CTA of its peers, never for another CTA of its own grid, and a wait
// gives up with a trap after 20 s instead of hanging the GPU.
#include "common.cuh"

namespace vnb {

__device__ __forceinline__ uint32_t cas_release_sys(uint32_t* p, uint32_t cmp, uint32_t val) {
    uint32_t old;
    asm volatile("atom.global.release.sys.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(p), "r"(cmp), "r"(val) : "memory");
    return old;
}
__device__ __forceinline__ uint32_t cas_acquire_sys(uint32_t* p, uint32_t cmp, uint32_t val) {
    uint32_t old;
    asm volatile("atom.global.acquire.sys.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(p), "r"(cmp), "r"(val) : "memory");
    return old;
}
__device__ __forceinline__ uint64_t global_ns() {
    uint64_t t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// one flag word per (CTA, peer): word [slot * world + sender] of the receiver's flag array
__device__ __forceinline__ void peer_barrier(uint32_t* const* flags, int rank, int world, int slot) {
    __syncthreads();
    if ((int)threadIdx.x < world && (int)threadIdx.x != rank) {
        const int peer = threadIdx.x;
        const uint64_t t0 = global_ns();
        uint32_t* theirs = flags[peer] + (size_t)slot * world + rank;
        while (cas_release_sys(theirs, 0u, 1u) != 0u) {                 // previous signal not consumed yet
            if (global_ns() - t0 > 20000000000ull) __trap();
        }
        uint32_t* mine = flags[rank] + (size_t)slot * world + peer;
        while (cas_acquire_sys(mine, 1u, 0u) != 1u) {
            if (global_ns() - t0 > 20000000000ull) __trap();
        }
    }
    __syncthreads();
}

__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(mc)
                 : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st_f4(float* mc, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z),
                 "f"(v.w)
                 : "memory");
}

constexpr int AR_THREADS = 512;
constexpr int AR_UNROLL = 4;

// count4 = float4 elements of the bucket; rank r reduces float4 [r * per, min((r+1) * per, count4))
__global__ void __launch_bounds__(AR_THREADS) allreduce_multimem_kernel(float* __restrict__ mc, int64_t count4,
                                                                         uint32_t* const* __restrict__ flags, int rank,
                                                                         int world, float scale) {
    pdl_wait();
    peer_barrier(flags, rank, world, blockIdx.x);
    const int64_t per = (count4 + world - 1) / world;
    const int64_t lo = per * rank, hi = lo + per < count4 ? lo + per : count4;
    const int64_t stride = (int64_t)gridDim.x * AR_THREADS;
    int64_t i = lo + (int64_t)blockIdx.x * AR_THREADS + threadIdx.x;
    for (; i + (AR_UNROLL - 1) * stride < hi; i += AR_UNROLL * stride) {
        float4 v[AR_UNROLL];
#pragma unroll
        for (int u = 0; u < AR_UNROLL; u++) v[u] = multimem_ld_reduce_f4(mc + 4 * (i + u * stride));
#pragma unroll
        for (int u = 0; u < AR_UNROLL; u++) {
            v[u].x *= scale; v[u].y *= scale; v[u].z *= scale; v[u].w *= scale;
            multimem_st_f4(mc + 4 * (i + u * stride), v[u]);
        }
    }
    for (; i < hi; i += stride) {
        float4 v = multimem_ld_reduce_f4(mc + 4 * i);
        v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale;
        multimem_st_f4(mc + 4 * i, v);
    }
    __threadfence_system();
    peer_barrier(flags, rank, world, blockIdx.x);
}

}  // namespace vnb

// mc_base: multicast address of the symmetric fp32 buffer; [offset, offset + count) must be 16-byte aligned
// (offset % 4 == 0, count % 4 == 0).  flags: DEVICE array of `world` pointers to each rank's flag words
// (>= nblocks * world zero-initialised uint32 each).  Every rank must launch the same (offset, count, nblocks)
// sequence.
extern "C" int vnb_allreduce_multimem(void* mc_base, int64_t offset, int64_t count, const void* flags, int rank, int world,
                                      float scale, int nblocks, void* stream) {
    VNB_REQUIRE(mc_base && flags && world >= 2 && world <= 32 && rank >= 0 && rank < world, "allreduce_multimem: bad arguments");
    VNB_REQUIRE(offset >= 0 && count >= 0 && (offset & 3) == 0 && (count & 3) == 0,
                "allreduce_multimem: offset %lld / count %lld must be multiples of 4 elements", (long long)offset,
                (long long)count);
    VNB_REQUIRE(nblocks >= 1 && nblocks <= 128, "allreduce_multimem: nblocks %d out of range", nblocks);
    if (count == 0) return VNB_OK;
    // an SM switches its L1 / shared-memory split only when idle: ask for the GEMMs' split (maximum shared memory) so
    // that a conv CTA (197 KB) can be placed on an SM that is already running one of these CTAs
    static bool carveout_set = false;
    if (!carveout_set) {
        cudaFuncSetAttribute(vnb::allreduce_multimem_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                             cudaSharedmemCarveoutMaxShared);
        carveout_set = true;
    }
    vnb::launch_k(vnb::allreduce_multimem_kernel, nblocks, vnb::AR_THREADS, 0, (cudaStream_t)stream,
                  (float*)mc_base + offset, count / 4, (uint32_t* const*)flags, rank, world, scale);
    VNB_LAUNCH_CHECK("allreduce_multimem");
    return VNB_OK;
}
