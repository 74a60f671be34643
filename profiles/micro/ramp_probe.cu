// What does a phase of the persistent kernel wait for right after a grid barrier?  Every CTA writes its part of an activation
// buffer, all CTAs pass a grid barrier (as in mega_generate_kernel), then one thread per CTA issues TMA loads and times their arrival:
//   A: one 16 KB K block of the activations just written by other SMs (L2-resident, dirty);
//   W: one 16 KB tile of a large weight buffer -- a new address every round (HBM) or the same one (L2) -- optionally 6 tiles at once.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I ../../pytorch-zappa-serverless_b200/csrc -I ../../include -o ramp_probe ramp_probe.cu -lcuda
#include <cstdio>
#include <cstdlib>
#include "gemm_tc.cuh"
namespace awd { int set_error(int code, const char*, ...) { return code; } std::atomic<int64_t> g_launches{0}; }
using namespace awd::tc;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ void grid_sync(unsigned* counter, unsigned& epoch) {
    asm volatile("fence.proxy.async.global;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {
        epoch += gridDim.x;
        __threadfence();
        atomicAdd(counter, 1u);
        unsigned v = 0;
        for (unsigned spin = 0;; ++spin) {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
            if ((int)(v - epoch) >= 0) break;
            if (spin > (1u << 26)) __trap();
        }
        __threadfence();
    }
    __syncthreads();
    asm volatile("fence.proxy.async.global;" ::: "memory");
}

struct P { __nv_bfloat16* hb; int iters, cold, nw, write_first; unsigned* counter; long long* out; };

__global__ void __launch_bounds__(128, 1) ramp_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, const P p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t sA = base, sW = base + 16384, barA = base + 16384 + 8 * 16384, barW = barA + 8;
    unsigned epoch = 0;
    if (threadIdx.x == 0) { mbar_init(barA, 1); mbar_init(barW, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); prefetch_tmap(&tmA); prefetch_tmap(&tmW); }
    __syncthreads();
    long long accA = 0, accW = 0, accB = 0;
    for (int it = 0; it < p.iters; ++it) {
        if (p.write_first) {    // every CTA rewrites its 1/gridDim of the 128 x 2304 activation buffer
            const int total = 128 * 2304 / 8, per = (total + gridDim.x - 1) / gridDim.x;
            for (int i = blockIdx.x * per + threadIdx.x; i < min(total, (int)(blockIdx.x + 1) * per); i += 128)
                reinterpret_cast<uint4*>(p.hb)[i] = make_uint4(it, it, it, it);
        }
        long long tb = clock64();
        grid_sync(p.counter, epoch);
        if (threadIdx.x == 0) {
            const long long t0 = clock64();
            accB += t0 - tb;
            mbar_expect_tx(barA, 16384);
            tma_load_2d(sA, &tmA, ((it + blockIdx.x) % 36) * 64, 0, barA);
            mbar_expect_tx(barW, 16384 * p.nw);
            for (int w = 0; w < p.nw; ++w) {
                const int row = p.cold ? ((it * 148 + blockIdx.x) * 128) % (1 << 20) : blockIdx.x * 128;
                tma_load_2d(sW + w * 16384, &tmW, w * 64, row, barW);
            }
            mbar_wait(barA, it & 1);
            const long long t1 = clock64();
            mbar_wait(barW, it & 1);
            const long long t2 = clock64();
            accA += t1 - t0; accW += t2 - t0;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) { p.out[blockIdx.x * 3] = accA; p.out[blockIdx.x * 3 + 1] = accW; p.out[blockIdx.x * 3 + 2] = accB; }
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
    void *hb, *W; unsigned* counter; long long* out;
    const size_t wrows = 1 << 20;    // 1 Mi rows x 512 cols bf16 = 1 GiB
    CK(cudaMalloc(&hb, 128 * 2304 * 2)); CK(cudaMalloc(&W, wrows * 512 * 2)); CK(cudaMalloc(&counter, 64)); CK(cudaMalloc(&out, 148 * 3 * 8));
    CK(cudaMemset(hb, 0, 128 * 2304 * 2)); CK(cudaMemset(W, 0, wrows * 512 * 2));
    CUtensorMap tmA, tmW;
    make_map(&tmA, hb, 128, 2304, 128);
    make_map(&tmW, W, wrows, 512, 128);
    CK(cudaFuncSetAttribute(ramp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    for (int write_first : {1, 0}) for (int cold : {1, 0}) for (int nw : {1, 6}) {
        CK(cudaMemset(counter, 0, 64));
        P p{(__nv_bfloat16*)hb, 200, cold, nw, write_first, counter, out};
        void* args[] = {(void*)&tmA, (void*)&tmW, (void*)&p};
        CK(cudaLaunchCooperativeKernel((void*)ramp_kernel, dim3(148), dim3(128), args, 200 * 1024, 0));
        CK(cudaDeviceSynchronize());
        long long h[148 * 3]; CK(cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost));
        double a = 0, w = 0, b = 0; for (int i = 0; i < 148; ++i) { a += h[3 * i]; w += h[3 * i + 1]; b += h[3 * i + 2]; }
        printf("activations %s, weights %s, %d x 16 KB weight tile(s): A lands after %.0f clocks, W after %.0f clocks (mean over CTAs and rounds); write+barrier %.0f clocks\n",
               write_first ? "rewritten by all SMs before the barrier" : "untouched", cold ? "from HBM" : "L2-resident", nw, a / 148 / p.iters, w / 148 / p.iters, b / 148 / p.iters);
    }
    return 0;
}
