// GPU debug probe (not product code): one TMA 3-D tile load through the library's tensor-map encoder.
//   tma_probe <case>   0: weight-like map (0,0,0)  1: activation map (0,0,0)  2: activation map (-1,0,0)
//                      3: activation map, 2 chunks  4: like 1 but no swizzle check (prints first values)
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../vae_npvc_b200/csrc/tc_common.cuh"
using namespace vnb;
using namespace vnb::tc;

__global__ void probe(const __grid_constant__ CUtensorMap map, int c0, int c1, int c2, uint32_t bytes, bf16* out, int n) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        fence_barrier_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_expect_tx(&bar, bytes);
        tma_load_3d(smem, &map, &bar, c0, c1, c2);
    }
    bool ok = mbar_wait(&bar, 0);
    if (threadIdx.x == 0 && !ok) printf("probe: wait timed out\n");
    for (int i = threadIdx.x; i < n; i += blockDim.x) out[i] = ((bf16*)smem)[i];
}

int main(int argc, char** argv) {
    int cs = argc > 1 ? atoi(argv[1]) : 0;
    const int T = 128, C = 64, PB = 2, Tp = 128;
    std::vector<bf16> h((size_t)PB * C * Tp);
    for (size_t i = 0; i < h.size(); i++) h[i] = __float2bfloat16((float)(i % 512));
    bf16 *d, *out;
    cudaMalloc(&d, h.size() * 2);
    cudaMemcpy(d, h.data(), h.size() * 2, cudaMemcpyHostToDevice);
    const int n = 64 * 64;
    cudaMalloc(&out, n * 2);
    CUtensorMap m;
    bool okm;
    int c0 = 0;
    if (cs == 0) okm = make_tmap_3d(&m, d, 64, 128, 1, 64 * 2, 128 * 64 * 2, 64, 64, 1);   // weights (ci,co,pt)
    else okm = make_tmap_3d(&m, d, T, C, PB, Tp * 2, (uint64_t)C * Tp * 2, 64, 64, 1);
    if (!okm) { printf("tmap failed: %s\n", vnb_last_error()); return 1; }
    if (cs == 2) c0 = -1;
    if (cs == 5) c0 = 1;
    if (cs == 6) c0 = -8;
    if (cs == 7) c0 = 8;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    probe<<<1, 128, 32 * 1024>>>(m, c0, 0, 0, 64 * 64 * 2, out, n);
    cudaError_t e = cudaDeviceSynchronize();
    printf("case %d: %s\n", cs, cudaGetErrorString(e));
    if (e == cudaSuccess) {
        std::vector<bf16> r(n);
        cudaMemcpy(r.data(), out, n * 2, cudaMemcpyDeviceToHost);
        for (int i = 0; i < 16; i++) printf("%g ", __bfloat162float(r[i]));
        printf("\n");
        for (int i = 64; i < 80; i++) printf("%g ", __bfloat162float(r[i]));
        printf("\n");
    }
    return 0;
}
