// Cost of one tcgen05.mma (kind::f16, bf16 operands from shared memory, SS mode) as a function of M and N: issue `iters` MMAs of
// K = 16 back to back from one thread into the same accumulator (operands: whatever is in shared memory), commit, wait.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I ../../pytorch-zappa-serverless_b200/csrc -I ../../include -o mma_probe mma_probe.cu
#include <cstdio>
#include <cstdlib>
#include "gemm_tc.cuh"
namespace awd { int set_error(int code, const char*, ...) { return code; } std::atomic<int64_t> g_launches{0}; }
using namespace awd::tc;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__global__ void __launch_bounds__(128, 1) mma_kernel(int M, int N, int iters, int nacc, unsigned long long* clk) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
    const uint32_t sA = base, sB = base + 16384, bar = base + 16384 + 32768, tslot = bar + 16;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += 128) reinterpret_cast<uint32_t*>(gen)[i] = 0x3c003c00u;   // small bf16 values
    if (threadIdx.x == 0) { mbar_init(bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    if (warp == 1) tmem_alloc(tslot, 512);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *reinterpret_cast<uint32_t*>(gen + (tslot - base));
    if (warp == 1 && lane == 0) {
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
        const uint64_t adesc = umma_desc_sw128(sA), bdesc = umma_desc_sw128(sB);
        long long t0 = clock64();
        for (int i = 0; i < iters; ++i) umma_bf16(tmem_base + (i % nacc) * N, adesc + 2 * (i & 3), bdesc + 2 * (i & 3), idesc, i >= nacc);
        umma_commit(bar);
        mbar_wait(bar, 0);
        long long t1 = clock64();
        clk[blockIdx.x] = (unsigned long long)(t1 - t0);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, 512); }
}

int main() {
    CK(cudaSetDevice(0));
    unsigned long long* clk; CK(cudaMalloc(&clk, 148 * 8));
    CK(cudaFuncSetAttribute(mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
    // warm the clocks up
    for (int r = 0; r < 20; ++r) mma_kernel<<<148, 128, 64 * 1024>>>(128, 256, 20000, 1, clk);
    CK(cudaDeviceSynchronize());
    for (int M : {128, 64}) for (int N : {16, 32, 64, 128, 256}) for (int nacc : {1, 2}) {
        if (M == 64 && N % 8) continue;
        const int iters = 4000;
        mma_kernel<<<148, 128, 64 * 1024>>>(M, N, iters, nacc, clk);
        CK(cudaDeviceSynchronize());
        unsigned long long h[148]; CK(cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost));
        printf("M %3d N %3d, %d accumulator(s): %.1f clocks per tcgen05.mma (K=16), all 148 SMs busy\n", M, N, nacc, (double)h[0] / iters);
    }
    return 0;
}
