// Clocks per 64-wide K block of the engine's TMA ring + tcgen05 main loop as a function of how the tiles are laid over the SMs:
// grid CTAs, tile (128 x bn), tiles_m M tiles (CTA -> M tile cta % tiles_m, N tile cta / tiles_m), ring depth, and whether the
// weights come from L2 (the same 21 MB slab every iteration) or from HBM (a different slab of a 340 MB buffer every iteration).
// Options per configuration: an L2 prefetch of the weight block `pf` K blocks ahead (> 0: TMA prefetch by the producer, < 0: prefetch.global.L2 by another
// warp); eight more warps kept busy the way epilogue warps would be -- mode 1 TMEM reads, 2 FFMA, 3 MUFU, 4 global atomics + loads,
// 5 64-bit integer multiplies (IMAD.WIDE), 6 32-bit IMAD, 7 shifts / LOP3 / IADD, 8 / 9 / 10: mode 5 only in the warps of SM sub-partitions 2+3 / 1 (the MMA
// issuer's) / 0 (the TMA producer's), 11 strided fp32 row loads + stores (the cell epilogue's state traffic), 12 waiting on an mbarrier.
// usage: kb_probe  (runs a fixed list)
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I ../../pytorch-zappa-serverless_b200/csrc -I ../../include -o kb_probe kb_probe.cu -lcuda
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "gemm_tc.cuh"

namespace awd { int set_error(int code, const char*, ...) { return code; } std::atomic<int64_t> g_launches{0}; }
using namespace awd::tc;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

struct P { int conv; int num_kb, bn, stages, tiles_m, iters, slabs, slab_rows, pf, epi; unsigned long long* clk; const void* w; int num_kb_total; };

__global__ void __launch_bounds__(320, 1) kb_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const P p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
    const uint32_t B_STAGE = (uint32_t)p.bn * BK * 2;
    const uint32_t sA = base, sB = base + p.stages * A_STAGE_BYTES;
    const uint32_t bars = sB + p.stages * B_STAGE;
    const uint32_t full0 = bars, empty0 = bars + 128, tslot = bars + 256;
    volatile int* progress = reinterpret_cast<volatile int*>(gen + (tslot + 16 - base));
    if (threadIdx.x == 0) *progress = 0;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        mbar_init(tslot + 32, 1);
        for (int s = 0; s < p.stages; ++s) { mbar_init(full0 + 8 * s, 1); mbar_init(empty0 + 8 * s, 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) tmem_alloc(tslot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *reinterpret_cast<uint32_t*>(gen + (tslot - base));
    const int mt = blockIdx.x % p.tiles_m, tile = blockIdx.x / p.tiles_m;
    long long t0 = clock64();
    if (warp == 0 && lane == 0) {
        int stage = 0; uint32_t phase = 0;
        for (int it = 0; it < p.iters; ++it) {
            const int n0 = (it % p.slabs) * p.slab_rows + tile * p.bn;
            for (int kb = 0; kb < p.num_kb; ++kb) {
                *progress = it * p.num_kb + kb;
                if (p.pf > 0 && mt == (kb % p.tiles_m)) {           // one CTA of the tiles_m that share the weight tile asks it into L2
                    const int kk = kb + p.pf, it2 = it + kk / p.num_kb;
                    tma_prefetch_l2_2d(&tmB, (kk % p.num_kb) * BK, (it2 % p.slabs) * p.slab_rows + tile * p.bn);
                }
                mbar_wait(empty0 + 8 * stage, phase ^ 1);
                const uint32_t fb = full0 + 8 * stage;
                mbar_expect_tx(fb, A_STAGE_BYTES + B_STAGE);
                tma_load_2d(sA + stage * A_STAGE_BYTES, &tmA, kb * BK, mt * BM, fb);
                tma_load_2d(sB + stage * B_STAGE, &tmB, kb * BK, n0, fb);
                if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1 && p.conv) {
        int stage = 0; uint32_t phase = 0;
        const uint32_t idesc = umma_idesc_bf16(p.bn);
        for (int it = 0; it < p.iters; ++it)
            for (int kb = 0; kb < p.num_kb; ++kb) {
                mbar_wait(full0 + 8 * stage, phase);
                tc_fence_after();
                const uint64_t adesc = umma_desc_sw128(sA + stage * A_STAGE_BYTES);
                const uint64_t bdesc = umma_desc_sw128(sB + stage * B_STAGE);
#pragma unroll
                for (int k = 0; k < BK / UMMA_K; ++k) umma_bf16_elect(tmem_base, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
                umma_commit_elect(empty0 + 8 * stage);
                if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
        for (int s = 0; s < p.stages; ++s) { mbar_wait(empty0 + 8 * stage, phase ^ 1); if (++stage == p.stages) { stage = 0; phase ^= 1; } }
    } else if (warp == 1 && lane == 0) {
        int stage = 0; uint32_t phase = 0;
        const uint32_t idesc = umma_idesc_bf16(p.bn);
        for (int it = 0; it < p.iters; ++it)
            for (int kb = 0; kb < p.num_kb; ++kb) {
                mbar_wait(full0 + 8 * stage, phase);
                tc_fence_after();
                const uint64_t adesc = umma_desc_sw128(sA + stage * A_STAGE_BYTES);
                const uint64_t bdesc = umma_desc_sw128(sB + stage * B_STAGE);
#pragma unroll
                for (int k = 0; k < BK / UMMA_K; ++k) umma_bf16(tmem_base, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
                umma_commit(empty0 + 8 * stage);
                if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
        for (int s = 0; s < p.stages; ++s) { mbar_wait(empty0 + 8 * stage, phase ^ 1); if (++stage == p.stages) { stage = 0; phase ^= 1; } }
    }
    else if (warp == 2 && p.pf < 0) {
        // per-thread prefetch.global.L2 of the weight block -pf K blocks ahead of the producer (no TMA involved): the tiles_m CTAs
        // sharing a weight tile take turns by K block; 256 lines of 128 bytes per block of a 256-row tile
        const char* wbase = reinterpret_cast<const char*>(p.w);
        const int total = p.iters * p.num_kb;
        for (int idx = 0; idx < total; ++idx) {
            if (idx % p.tiles_m != mt) continue;
            while (idx > *progress - p.pf) { }
            const int it = idx / p.num_kb, kb = idx % p.num_kb;
            const int64_t row0 = (int64_t)(it % p.slabs) * p.slab_rows + tile * p.bn;
            for (int r = lane; r < p.bn; r += 32) {
                const char* a = wbase + ((row0 + r) * (int64_t)(p.num_kb_total * 64) + kb * 64) * 2;
                asm volatile("prefetch.global.L2 [%0];" ::"l"(a));
            }
        }
    }
    else if (warp >= 2 && p.epi) {
        // stand-ins for epilogue warps, busy until the main loop is through (progress reaches the end)
        const int total = p.iters * p.num_kb - 1;
        float acc = (float)lane, b = 1.0001f;
        unsigned u = threadIdx.x * 2654435761u;
        if (p.epi == 1) {
            if (warp < 4) while (*progress < total) { float v[32]; tmem_ld32(tmem_base + ((uint32_t)((warp & 3) * 32) << 16), v);
#pragma unroll
                for (int j = 0; j < 32; ++j) acc += v[j]; }
        } else if (p.epi == 2) {
            while (*progress < total) {
#pragma unroll
                for (int j = 0; j < 64; ++j) acc = fmaf(acc, b, 0.5f);
            }
        } else if (p.epi == 3) {
            while (*progress < total) {
#pragma unroll
                for (int j = 0; j < 16; ++j) acc = exp2f(acc * 0.001f) + 0.25f;
            }
        } else if (p.epi == 4) {
            int* gp = reinterpret_cast<int*>(const_cast<unsigned long long*>(p.clk)) + 512 + (blockIdx.x * 8 + (warp - 2)) * 32 + lane;
            while (*progress < total) { atomicMax(gp, (int)u); u = u * 1664525u + 1013904223u; acc += (float)__ldcg(gp); }
        } else if (p.epi == 11) {                  // strided fp32 row stores + loads, the cell epilogue's c / h traffic (32 rows x 32 B per instruction)
            float* gp = reinterpret_cast<float*>(const_cast<void*>(p.w)) + ((size_t)blockIdx.x * 8 + (warp - 2)) * 32 * 1152 + (size_t)lane * 1152;
            int col = 0;
            while (*progress < total) {
                float4 c4 = *reinterpret_cast<float4*>(gp + col);
                c4.x += acc; c4.y += 1.f;
                *reinterpret_cast<float4*>(gp + col) = c4;
                *reinterpret_cast<float2*>(gp + 1024 + ((col >> 3) << 1)) = make_float2(c4.x, c4.y);
                col = (col + 8) & 1023;
            }
        } else if (p.epi == 12) {                  // eight warps waiting on an mbarrier that never completes before the end (the epilogue's wait for the accumulator)
            while (*progress < total) {
                uint32_t done = 0;
                asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(done) : "r"(tslot + 32), "r"(0u) : "memory");
                acc += (float)done;
            }
        } else if (p.epi == 6) {
            while (*progress < total) {
#pragma unroll
                for (int j = 0; j < 64; ++j) u = u * 1664525u + 1013904223u;          // 32-bit IMAD
            }
        } else if (p.epi == 7) {
            while (*progress < total) {
#pragma unroll
                for (int j = 0; j < 64; ++j) u = ((u << 3) ^ (u >> 5)) + 0x9e3779b9u;   // shifts, LOP3, IADD
            }
        } else if (p.epi == 5 || (p.epi == 8 && (warp & 3) >= 2) || (p.epi == 9 && (warp & 3) == 1) || (p.epi == 10 && (warp & 3) == 0)) {
            while (*progress < total) {
#pragma unroll
                for (int j = 0; j < 32; ++j) { const unsigned long long m = (unsigned long long)u * 0xD2511F53u; u = (unsigned)(m >> 32) ^ (unsigned)m ^ j; }
            }
        }
        if (acc == 123.456f || u == 77u) p.clk[147] = 1;
    }
    tc_fence_before();
    __syncthreads();
    long long t1 = clock64();
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
    const int K = 2304, N = 4608, SLABS = 16;
    void *A, *W; unsigned long long* clk;
    CK(cudaMalloc(&A, 1024 * K * 2)); CK(cudaMalloc(&W, (size_t)SLABS * N * K * 2)); CK(cudaMalloc(&clk, 148 * 8 + 4096 + 148 * 8 * 32 * 4)); CK(cudaMemset(clk, 0, 148 * 8 + 4096 + 148 * 8 * 32 * 4));
    CK(cudaMemset(A, 0, 1024 * K * 2)); CK(cudaMemset(W, 0, (size_t)SLABS * N * K * 2));
    CK(cudaFuncSetAttribute(kb_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    struct Cfg { int grid, bn, tiles_m, num_kb, stages, pf, epi, dec, conv; };     // dec: the decoder's walk -- W[60000][448], 13 tiles of 256 rows per group of tiles_m CTAs
    std::vector<Cfg> cfgs = {{144, 256, 8, 36, 4, 0, 0, 0, 1}, {144, 256, 8, 36, 4, 0, 11, 0, 1}, {144, 256, 8, 36, 4, 0, 12, 0, 1}, {144, 256, 8, 36, 4, 0, 12, 0, 0}, {144, 256, 8, 36, 4, 0, 3, 0, 1}};
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (auto c : cfgs)
        for (int cold = 0; cold < 2; ++cold) {
            CUtensorMap tmA, tmB;
            make_map(&tmA, A, 1024, K, 128);
            const int groups = c.grid / c.tiles_m;
            if (c.dec) make_map(&tmB, W, 60000, 448, c.bn); else make_map(&tmB, W, (uint64_t)SLABS * N, K, c.bn);
            P p{c.conv, c.num_kb, c.bn, c.stages, c.tiles_m, c.dec ? 60000 / (groups * c.bn) : 48, cold ? SLABS : 1, N, c.pf, c.epi, clk, W, K / 64};
            if (c.dec) { p.slabs = cold ? 1 << 20 : 1; p.slab_rows = groups * c.bn; p.num_kb_total = 7; }
            const int smem = c.stages * (A_STAGE_BYTES + c.bn * 128) + 1024 + 512;
            float best = 1e9;
            for (int rep = 0; rep < 3; ++rep) {
                CK(cudaEventRecord(e0));
                kb_kernel<<<c.grid, 320, smem>>>(tmA, tmB, p);
                CK(cudaEventRecord(e1));
                CK(cudaDeviceSynchronize());
                float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
            }
            unsigned long long h[148]; CK(cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost));
            double mx = 0; for (int i = 0; i < c.grid; ++i) mx = h[i] > mx ? h[i] : mx;
            printf("%s%3d CTAs, tile 128x%3d, %d M tiles, %2d K blocks, %d stages, prefetch %2d ahead, busy warps mode %2d, %s MMA issue, weights from %s: %6.0f clocks per K block (slowest CTA), %.2f us per pass, L2->SM %.1f KB/clk\n",
                   c.dec ? "decoder walk: " : "", c.grid, c.bn, c.tiles_m, c.num_kb, c.stages, c.pf, c.epi, c.conv ? "convergent-warp" : "single-thread ", cold ? "HBM" : "L2 ", mx / p.iters / c.num_kb, best * 1e3 / p.iters,
                   c.grid * (16384.0 + c.bn * 128) / (mx / p.iters / c.num_kb) / 1024);
        }
    return 0;
}
