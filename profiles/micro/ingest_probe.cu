// How fast can the SMs ingest the operands of ONE LSTM phase at 128 streams (one M tile)?  D[128 x N] = A[128 x K] * W[N x K]^T with
// N split over the CTAs (BN columns each): every CTA streams the whole shared A plus its own weight rows.  Variants:
//   cluster 1: plain per-CTA TMA loads (what mega_generate_kernel does);
//   cluster c: each CTA loads 128/c rows of every A K block and TMA-multicasts them to the c CTAs of its cluster
//              (tensor map over this GPU's own memory), the weights stay per-CTA.
// Same ring / mbarrier / tcgen05 structure as the engine's kernels; the accumulators are never read.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I ../../pytorch-zappa-serverless_b200/csrc -I ../../include -o ingest_probe ingest_probe.cu -lcuda
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "gemm_tc.cuh"

namespace awd { int set_error(int code, const char*, ...) { return code; } std::atomic<int64_t> g_launches{0}; }
using namespace awd::tc;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ void tma_load_2d_mcast(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar, uint16_t mask) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4}], [%2], %5;"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "h"(mask) : "memory");
}
__device__ __forceinline__ void umma_commit_mcast(uint32_t bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar), "h"(mask) : "memory");
}

struct P { int num_kb, bn, stages, csize, iters, ncta_active, nacc; unsigned long long* clk; };

__global__ void __launch_bounds__(128, 1) ingest_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const P p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
    const uint32_t B_STAGE = (uint32_t)p.bn * BK * 2;
    const uint32_t sA = base, sB = base + p.stages * A_STAGE_BYTES;
    const uint32_t bars = sB + p.stages * ((B_STAGE + 1023) & ~1023u);
    const uint32_t full0 = bars, empty0 = bars + 128, tslot = bars + 256;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int csize = p.csize;
    const int rank = csize > 1 ? (int)cluster_ctarank() : 0;
    if (threadIdx.x == 0) {
        for (int s = 0; s < p.stages; ++s) { mbar_init(full0 + 8 * s, 1); mbar_init(empty0 + 8 * s, csize); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) tmem_alloc(tslot, 512);
    tc_fence_before();
    __syncthreads();
    if (csize > 1) cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *reinterpret_cast<uint32_t*>(gen + (tslot - base));
    const bool active = (int)blockIdx.x < p.ncta_active;
    long long t0 = clock64();
    if (warp == 0 && lane == 0) {
        int stage = 0; uint32_t phase = 0;
        const uint16_t mask_all = (uint16_t)((1u << csize) - 1);
        const int a_rows = BM / csize;
        const int n0 = blockIdx.x * p.bn;
        for (int it = 0; it < p.iters; ++it)
            for (int kb = 0; kb < p.num_kb; ++kb) {
                mbar_wait(empty0 + 8 * stage, phase ^ 1);
                const uint32_t fb = full0 + 8 * stage;
                mbar_expect_tx(fb, A_STAGE_BYTES + B_STAGE);
                if (csize == 1) tma_load_2d(sA + stage * A_STAGE_BYTES, &tmA, kb * BK, 0, fb);
                else tma_load_2d_mcast(sA + stage * A_STAGE_BYTES + rank * a_rows * 128, &tmA, kb * BK, rank * a_rows, fb, mask_all);
                tma_load_2d(sB + stage * ((B_STAGE + 1023) & ~1023u), &tmB, kb * BK, active ? n0 : 0, fb);
                if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
    } else if (warp == 1 && lane == 0) {
        int stage = 0; uint32_t phase = 0;
        const uint32_t idesc = umma_idesc_bf16(p.bn);
        const uint16_t mask_all = (uint16_t)((1u << csize) - 1);
        for (int it = 0; it < p.iters; ++it)
            for (int kb = 0; kb < p.num_kb; ++kb) {
                mbar_wait(full0 + 8 * stage, phase);
                tc_fence_after();
                const uint64_t adesc = umma_desc_sw128(sA + stage * A_STAGE_BYTES);
                const uint64_t bdesc = umma_desc_sw128(sB + stage * ((B_STAGE + 1023) & ~1023u));
#pragma unroll
                for (int k = 0; k < BK / UMMA_K; ++k) umma_bf16(tmem_base + (k % p.nacc) * p.bn, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | (k / p.nacc)) != 0);
                if (csize == 1) umma_commit(empty0 + 8 * stage); else umma_commit_mcast(empty0 + 8 * stage, mask_all);
                if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
        // drain: wait until every stage has been released once more (all MMAs retired, also the peers')
        for (int s = 0; s < p.stages; ++s) { mbar_wait(empty0 + 8 * stage, phase ^ 1); if (++stage == p.stages) { stage = 0; phase ^= 1; } }
    }
    tc_fence_before();
    __syncthreads();
    long long t1 = clock64();
    if (csize > 1) cluster_sync_all();
    if (threadIdx.x == 0) p.clk[blockIdx.x] = (unsigned long long)(t1 - t0);
    if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, 512); }
}

static void make_map(CUtensorMap* m, void* base, uint64_t rows, uint64_t cols, uint32_t box_rows) {
    typedef CUresult (*Fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                           CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static Fn fn = nullptr;
    if (!fn) { void* f = nullptr; cudaDriverEntryPointQueryResult q; CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q)); fn = (Fn)f; }
    cuuint64_t dims[2] = {cols, rows}; cuuint64_t strides[1] = {cols * 2}; cuuint32_t box[2] = {64, box_rows}; cuuint32_t es[2] = {1, 1};
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("tensor map encode failed %d\n", (int)r); exit(1); }
}

int main() {
    CK(cudaSetDevice(0));
    const int K = 2304, N = 4608;
    void *A, *W; unsigned long long* clk;
    CK(cudaMalloc(&A, 128 * K * 2)); CK(cudaMalloc(&W, (size_t)N * K * 2)); CK(cudaMalloc(&clk, 148 * 8));
    CK(cudaMemset(A, 0, 128 * K * 2)); CK(cudaMemset(W, 0, (size_t)N * K * 2));
    CK(cudaFuncSetAttribute(ingest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    CK(cudaFuncSetAttribute(ingest_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    struct Cfg { int csize, grid, bn, stages, nacc; };
    std::vector<Cfg> cfgs = {{1, 144, 32, 8, 1}, {1, 144, 32, 8, 2}, {1, 144, 32, 8, 4}, {1, 72, 64, 8, 4}, {1, 36, 128, 6, 4}, {1, 36, 128, 6, 2}, {2, 144, 32, 8, 4}, {4, 132, 32, 8, 4},
                             {4, 96, 48, 8, 4}, {8, 96, 48, 8, 4}, {8, 120, 32, 8, 4}, {4, 72, 64, 8, 4}, {4, 36, 128, 6, 4}};
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (auto c : cfgs) {
        CUtensorMap tmA, tmB;
        make_map(&tmA, A, 128, K, 128 / c.csize);
        make_map(&tmB, W, N, K, c.bn);
        P p{K / 64, c.bn, c.stages, c.csize, 50, (N + c.bn - 1) / c.bn, c.nacc, clk};
        const int smem = c.stages * (A_STAGE_BYTES + ((c.bn * 128 + 1023) & ~1023)) + 1024 + 512;
        cudaLaunchConfig_t cfg{}; cfg.gridDim = dim3(c.grid); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem;
        cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = c.csize; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = c.csize > 1 ? 1 : 0;
        float best = 1e9;
        for (int rep = 0; rep < 3; ++rep) {
            CK(cudaEventRecord(e0));
            CK(cudaLaunchKernelEx(&cfg, ingest_kernel, tmA, tmB, p));
            CK(cudaEventRecord(e1));
            CK(cudaDeviceSynchronize());
            float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
        }
        const double us = best * 1e3 / p.iters;
        const double mb_l2 = (c.grid * (128.0 / c.csize) * K * 2 + (double)c.grid * c.bn * K * 2) / 1e6;
        printf("cluster %d, %3d CTAs, BN %3d, %2d stages, %d accumulators: %.2f us per phase (L2 reads %.0f MB -> %.1f TB/s; per-SM ingest %.0f KB)\n", c.csize, c.grid, c.bn, c.stages, c.nacc, us, mb_l2,
               mb_l2 / us, (128.0 * K * 2 + c.bn * K * 2.0) / 1e3);
    }
    return 0;
}
