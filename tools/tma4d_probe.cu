// GPU probe (not product code): (1) does cuTensorMapEncodeTiled accept a 4-D map whose chunk dimension has a
// SMALLER stride than the row dimension (dims: 64 ch-in-chunk, rows, chunk, sample), and does one box {64,R,2,1}
// land as two canonical SWIZZLE_128B [R][128 B] tiles back to back?  (2) clock64 cost of issuing TMA loads.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../vae_npvc_b200/csrc/tc_common.cuh"
using namespace vnb;
using namespace vnb::tc;

__global__ void probe(const __grid_constant__ CUtensorMap m3, const __grid_constant__ CUtensorMap m4, int R, bf16* out3,
                      bf16* out4, long long* cyc) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t bar[4];
    const uint32_t tile = R * 128;
    if (threadIdx.x == 0) {
        for (int i = 0; i < 4; i++) mbar_init(&bar[i], 1);
        fence_barrier_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        // reference: two 3-D loads (chunk 0, chunk 1) of sample 1
        mbar_expect_tx(&bar[0], 2 * tile);
        tma_load_3d(smem, &m3, &bar[0], 0, 0, 1);
        tma_load_3d(smem + tile, &m3, &bar[0], 64, 0, 1);
        // candidate: one 4-D load of both chunks
        mbar_expect_tx(&bar[1], 2 * tile);
        tma_load_4d(smem + 2 * tile, &m4, &bar[1], 0, 0, 0, 1);
    }
    bool ok = mbar_wait(&bar[0], 0) && mbar_wait(&bar[1], 0);
    if (threadIdx.x == 0 && !ok) printf("probe: wait timed out\n");
    const int n = 2 * tile / 2;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        out3[i] = ((bf16*)smem)[i];
        out4[i] = ((bf16*)(smem + 2 * tile))[i];
    }
    __syncthreads();
    // ---- issue cost: 8 back-to-back loads by one thread (3-D single tile / 4-D double tile), then by two warps ----
    if (threadIdx.x == 0) {
        mbar_expect_tx(&bar[2], 8 * tile);
        long long t0 = clock64();
        for (int i = 0; i < 8; i++) tma_load_3d(smem + (i & 1) * tile, &m3, &bar[2], 0, 0, i & 1);
        long long t1 = clock64();
        cyc[0] = t1 - t0;
        mbar_wait(&bar[2], 0);
        cyc[1] = clock64() - t1;
        mbar_expect_tx(&bar[3], 16 * tile);
        t0 = clock64();
        for (int i = 0; i < 8; i++) tma_load_4d(smem, &m4, &bar[3], 0, 0, 0, i & 1);
        t1 = clock64();
        cyc[2] = t1 - t0;
        mbar_wait(&bar[3], 0);
        cyc[3] = clock64() - t1;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_init(&bar[0], 2);
        fence_barrier_init();
    }
    __syncthreads();
    if ((threadIdx.x & 31) == 0 && threadIdx.x < 64) {
        const int w = threadIdx.x >> 5;
        mbar_expect_tx(&bar[0], 8 * tile);
        long long t0 = clock64();
        for (int i = 0; i < 8; i++) tma_load_3d(smem + w * tile, &m3, &bar[0], 0, 0, i & 1);
        cyc[4 + w] = clock64() - t0;
    }
    mbar_wait(&bar[0], 0);
}

int main() {
    const int T = 128, C = 512, PB = 2;
    std::vector<bf16> h((size_t)PB * T * C);
    for (size_t i = 0; i < h.size(); i++) h[i] = __float2bfloat16((float)((i * 7) % 509));
    bf16 *d, *o3, *o4;
    long long* cyc;
    cudaMalloc(&d, h.size() * 2);
    cudaMemcpy(d, h.data(), h.size() * 2, cudaMemcpyHostToDevice);
    for (int R : {128, 64}) {
        const int n = 2 * R * 64;
        cudaMalloc(&o3, n * 2);
        cudaMalloc(&o4, n * 2);
        cudaMalloc(&cyc, 64);
        CUtensorMap m3, m4;
        if (!make_tmap_3d(&m3, d, C, T, PB, C * 2, (uint64_t)T * C * 2, 64, R, 1)) { printf("3d: %s\n", vnb_last_error()); return 1; }
        const uint64_t dims[4] = {64, (uint64_t)T, (uint64_t)C / 64, (uint64_t)PB};
        const uint64_t strides[3] = {(uint64_t)C * 2, 128, (uint64_t)T * C * 2};
        const uint32_t box[4] = {64, (uint32_t)R, 2, 1};
        if (!make_tmap(&m4, d, 4, dims, strides, box)) { printf("4d encode REJECTED: %s\n", vnb_last_error()); return 2; }
        printf("4d encode accepted (R=%d)\n", R);
        cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
        probe<<<1, 128, 100 * 1024>>>(m3, m4, R, o3, o4, cyc);
        cudaError_t e = cudaDeviceSynchronize();
        printf("kernel: %s\n", cudaGetErrorString(e));
        if (e != cudaSuccess) return 3;
        std::vector<bf16> r3(n), r4(n);
        long long c[8];
        cudaMemcpy(r3.data(), o3, n * 2, cudaMemcpyDeviceToHost);
        cudaMemcpy(r4.data(), o4, n * 2, cudaMemcpyDeviceToHost);
        cudaMemcpy(c, cyc, 64, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int i = 0; i < n; i++) bad += __bfloat162float(r3[i]) != __bfloat162float(r4[i]);
        printf("R=%d: 4d vs 2x3d mismatches %d / %d\n", R, bad, n);
        printf("issue 8x 3d(%d rows): %lld cyc (%.0f each), drain %lld | 8x 4d(2x%d rows): %lld cyc (%.0f each), drain %lld | "
               "two warps 8x each: %lld / %lld cyc\n", R, c[0], c[0] / 8.0, c[1], R, c[2], c[2] / 8.0, c[3], c[4], c[5]);
    }
    return 0;
}
