// Micro-measurements behind the small-batch generation kernel's design (B200):
//   1. how many CTAs of a cluster-launched, cooperative, 1-CTA-per-SM kernel fit (cluster size 1/2/4/8)
//   2. cost of one grid-wide barrier (atomic counter + acquire spin) with idle CTAs
//   3. point-to-point flag latency between two SMs (st.release -> ld.acquire ping-pong)
//   4. DSMEM exchange: every CTA of a cluster of 4 writes 16 KB to each peer, then cluster barrier
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o sync_probe sync_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }

__global__ void __launch_bounds__(320, 1) occ_kernel(int* out) { extern __shared__ uint8_t sm[]; if (threadIdx.x == 0 && out) out[blockIdx.x] = sm[0]; }

__device__ __forceinline__ void grid_sync(unsigned* counter, unsigned& epoch, int mode) {
    if (mode == 0) asm volatile("fence.proxy.async.global;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {
        epoch += gridDim.x;
        if (mode == 0) { __threadfence(); atomicAdd(counter, 1u); }
        else asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(counter) : "memory");
        unsigned v = 0;
        for (unsigned spin = 0;; ++spin) {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
            if ((int)(v - epoch) >= 0) break;
            if (spin > (1u << 24)) __trap();
        }
        if (mode == 0) __threadfence();
    }
    __syncthreads();
    if (mode == 0) asm volatile("fence.proxy.async.global;" ::: "memory");
}

__global__ void __launch_bounds__(320, 1) barrier_kernel(unsigned* counter, int iters, int mode, float* sink, unsigned long long* ns) {
    unsigned epoch = 0;
    unsigned long long t0 = 0;
    grid_sync(counter, epoch, mode);
    if (blockIdx.x == 0 && threadIdx.x == 0) t0 = gtime();
    for (int i = 0; i < iters; ++i) {
        sink[blockIdx.x * 320 + threadIdx.x] = (float)i;      // one store per thread per phase (so the fences have work)
        grid_sync(counter, epoch, mode);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) ns[0] = gtime() - t0;
}

// ping-pong between CTA 0 and CTA `peer`: flag latency one way = rtt / 2
__global__ void __launch_bounds__(32, 1) pingpong_kernel(unsigned* flags, int iters, int peer, unsigned long long* ns) {
    if (threadIdx.x != 0) return;
    unsigned* a = flags, *b = flags + 64;
    if (blockIdx.x == 0) {
        unsigned long long t0 = gtime();
        for (int i = 1; i <= iters; ++i) {
            asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(a), "r"((unsigned)i) : "memory");
            unsigned v = 0;
            for (unsigned spin = 0; v != (unsigned)i; ++spin) { asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(b) : "memory"); if (spin > (1u << 24)) __trap(); }
        }
        ns[0] = gtime() - t0;
    } else if ((int)blockIdx.x == peer) {
        for (int i = 1; i <= iters; ++i) {
            unsigned v = 0;
            for (unsigned spin = 0; v != (unsigned)i; ++spin) { asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(a) : "memory"); if (spin > (1u << 24)) __trap(); }
            asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(b), "r"((unsigned)i) : "memory");
        }
    }
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t mapa(uint32_t addr, uint32_t rank) { uint32_t r; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank)); return r; }
__device__ __forceinline__ void cluster_sync() { asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory"); }

// mode 0: st.shared::cluster.v4.f32 from 256 threads, layout [src][j][row] (lanes -> consecutive 16 B)
// mode 1: stage in local smem, then cp.async.bulk.shared::cluster (one 16 KB bulk copy per peer) with mbarrier complete_tx
template <int CS>
__global__ void __launch_bounds__(256, 1) dsmem_kernel(int iters, int mode, unsigned long long* clk, float* sink) {
    extern __shared__ __align__(1024) uint8_t sm[];
    float* recv = reinterpret_cast<float*>(sm);                    // [CS][8][128] float4 = CS * 16 KB
    float* stage = reinterpret_cast<float*>(sm + CS * 16384);      // [CS][8][128] float4
    uint64_t* bar = reinterpret_cast<uint64_t*>(sm + 2 * CS * 16384);
    uint32_t rank; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    const int row = threadIdx.x & 127, half = threadIdx.x >> 7;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    cluster_sync();
    long long t0 = clock64();
    float acc = 0.f;
    for (int it = 0; it < iters; ++it) {
        if (mode == 0) {
            for (int p = 0; p < CS; ++p) {
                if (p == (int)rank) continue;
                const uint32_t dst = mapa(smem_u32(recv + ((rank * 8) * 128) * 4), p);
#pragma unroll
                for (int j = half * 4; j < half * 4 + 4; ++j) {
                    const float v = (float)(it + j + row);
                    asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(dst + (j * 128 + row) * 16), "f"(v), "f"(v), "f"(v), "f"(v) : "memory");
                }
            }
            cluster_sync();
        } else {
            if (threadIdx.x == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"((CS - 1) * 16384) : "memory");
            for (int p = 0; p < CS; ++p) {
#pragma unroll
                for (int j = half * 4; j < half * 4 + 4; ++j) {
                    const float v = (float)(it + j + row);
                    *reinterpret_cast<float4*>(stage + ((p * 8 + j) * 128 + row) * 4) = make_float4(v, v, v, v);
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();
            cluster_sync();       // peers have armed their barriers and finished reading the previous round
            if (threadIdx.x < CS && threadIdx.x != rank) {
                const int p = threadIdx.x;
                const uint32_t dst = mapa(smem_u32(recv + ((rank * 8) * 128) * 4), p);
                const uint32_t rbar = mapa(smem_u32(bar), p);
                asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(dst), "r"(smem_u32(stage + ((p * 8) * 128) * 4)), "r"(16384), "r"(rbar) : "memory");
            }
            uint32_t done = 0;
            for (unsigned spin = 0; !done; ++spin) {
                asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(done) : "r"(smem_u32(bar)), "r"((uint32_t)(it & 1)) : "memory");
                if (spin > (1u << 24)) __trap();
            }
        }
        // consume: thread = row reads its 3 x 32 floats
        for (int p = 0; p < CS; ++p) {
            if (p == (int)rank) continue;
#pragma unroll
            for (int j = half * 4; j < half * 4 + 4; ++j) {
                const float4 v = *reinterpret_cast<const float4*>(recv + ((p * 8 + j) * 128 + row) * 4);
                acc += v.x + v.y + v.z + v.w;
            }
        }
        if (mode == 0) cluster_sync();   // everyone has read before the next round overwrites
    }
    long long t1 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) clk[0] = (unsigned long long)(t1 - t0);
    sink[blockIdx.x * 256 + threadIdx.x] = acc;
    cluster_sync();
}

int main() {
    int dev = 0; CK(cudaSetDevice(dev));
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, dev));
    printf("device %s, %d SMs, clock %d kHz\n", prop.name, prop.multiProcessorCount, prop.clockRate);
    // 1. occupancy of clusters
    CK(cudaFuncSetAttribute(occ_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(occ_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    for (int cs : {1, 2, 4, 8, 16}) {
        cudaLaunchConfig_t cfg{}; cfg.gridDim = dim3(prop.multiProcessorCount / cs * cs); cfg.blockDim = dim3(320); cfg.dynamicSmemBytes = 200 * 1024;
        cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        int n = 0; cudaError_t e = cudaOccupancyMaxActiveClusters(&n, occ_kernel, &cfg);
        printf("cluster size %2d: max active clusters %d (%d CTAs) %s\n", cs, n, n * cs, e == cudaSuccess ? "" : cudaGetErrorString(e));
        cudaGetLastError();
    }
    unsigned* counter; float* sink; unsigned long long* ns; unsigned long long hns;
    CK(cudaMalloc(&counter, 1024)); CK(cudaMalloc(&sink, 148 * 320 * 4)); CK(cudaMalloc(&ns, 64));
    // 2. grid barrier
    for (int mode = 0; mode < 2; ++mode) for (int grid : {148, 132}) {
        CK(cudaMemset(counter, 0, 1024));
        int iters = 2000;
        void* args[] = {&counter, &iters, &mode, &sink, &ns};
        CK(cudaLaunchCooperativeKernel((void*)barrier_kernel, dim3(grid), dim3(320), args, 0, 0));
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(&hns, ns, 8, cudaMemcpyDeviceToHost));
        printf("grid barrier mode %d (%s), %d CTAs: %.3f us per barrier\n", mode, mode ? "red.release + ld.acquire only" : "as mega_generate_kernel", grid, hns * 1e-3 / iters);
    }
    // 3. flag ping-pong
    for (int peer : {1, 2, 37, 74, 147}) {
        CK(cudaMemset(counter, 0, 1024));
        int iters = 2000;
        void* args[] = {&counter, &iters, &peer, &ns};
        CK(cudaLaunchCooperativeKernel((void*)pingpong_kernel, dim3(148), dim3(32), args, 0, 0));
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(&hns, ns, 8, cudaMemcpyDeviceToHost));
        printf("flag ping-pong CTA 0 <-> CTA %d: one-way %.3f us\n", peer, hns * 1e-3 / iters / 2);
    }
    // 4. DSMEM exchange
    {
        auto k4 = dsmem_kernel<4>; auto k2 = dsmem_kernel<2>;
        CK(cudaFuncSetAttribute(k4, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 4 * 16384 + 64));
        CK(cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 2 * 16384 + 64));
        for (int cs : {4, 2}) for (int mode = 0; mode < 2; ++mode) {
            cudaLaunchConfig_t cfg{}; cfg.gridDim = dim3(cs == 4 ? 132 : 148); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = 2 * cs * 16384 + 64;
            cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            int iters = 200;
            if (cs == 4) CK(cudaLaunchKernelEx(&cfg, k4, iters, mode, ns, sink)); else CK(cudaLaunchKernelEx(&cfg, k2, iters, mode, ns, sink));
            CK(cudaDeviceSynchronize());
            CK(cudaMemcpy(&hns, ns, 8, cudaMemcpyDeviceToHost));
            printf("DSMEM exchange cluster %d mode %d (%s): %.0f clocks per round (%d KB sent and received per CTA, all clusters at once)\n", cs, mode,
                   mode ? "cp.async.bulk smem->dsmem" : "st.shared::cluster.v4", (double)hns / iters, (cs - 1) * 16);
        }
    }
    return 0;
}
