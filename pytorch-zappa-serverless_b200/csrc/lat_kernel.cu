// Low-latency generation for a handful of streams (bs <= 8): the persistent recurrent kernel of
// BASELINE.json's north star.
//
// With one stream a token is ~88 MB of bf16 weights against ~88 MFLOP: tensor cores have nothing to
// chew on, latency is weight delivery plus synchronisation.  So: ONE cooperative launch for the whole
// request (prompt + generated tokens), one CTA per SM, and
//   * every CTA owns a fixed slice of hidden units of every layer and keeps that slice of W_hh
//     RESIDENT IN SHARED MEMORY for the life of the kernel (shipped size: 158 KB per SM, 22.4 MB over
//     the chip); W_ih and the decoder rows stream from L2 with coalesced 128-bit loads;
//   * a warp computes the four gate rows of one hidden unit for all streams, reduces with warp
//     shuffles and applies sigmoid/tanh + the cell update in registers; the cell state never leaves
//     the SM;
//   * one grid-wide barrier per layer and per decoder pass; the sampled token is chosen on the device
//     (exponential race with the reference's 10-draw / first-allowed rule, main.py:61-68) and fed back
//     without a host round trip.
#include <cstdio>
#include <cstdlib>

#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_api.h"

namespace awd {
namespace tc {

constexpr int LAT_MAX_WARPS = 16;       // thread count is a template parameter: 512 for 1-2 streams, 256 above

struct LatParams {
    int n_layers, bs, V, P, n_new, Epad, par0;
    int H[4], Ipad[4], Hpad[4], upc[4];
    int w_off[4];                        // byte offset of the resident W_hh slice in shared memory, or -1 (streamed)
    int wi_off[4];                       // same for the W_ih slice (what is left of shared memory, first layers first)
    int c_off[4];                        // float offset of the layer's cell-state slice in the smem state array
    int vec_off, vh_off, state_off, red_off;   // byte offsets
    const __nv_bfloat16* wcat[4];
    const float* bias[4];
    float* c[4];
    __nv_bfloat16* hb[4][2];
    const __nv_bfloat16* table;
    const __nv_bfloat16* dec_w;
    const float* dec_bias;
    const int64_t* prompt;
    int64_t* gen;
    SampleParams sp;
    float4* part;                        // [grid][8]
    float* banned_keys;                  // [8][64]
    unsigned* sync_counter;
    int* err_flag;
    const int32_t* start;                // ragged prompts: stream b is fed from step start[b] on (nullptr: 0)
    int debug;                           // AWD_LAT_DEBUG=1: CTA 0 prints where a token's time goes
};

// Grid-wide barrier in two halves.  arrive: CTA barrier (orders every thread's writes before thread 0), thread 0 bumps the
// counter with release semantics.  wait: thread 0 spins with acquire loads, CTA barrier.  Everything another CTA wrote is
// afterwards read with ld.global.cg (L2), so the other threads need no fence of their own.  Between the two halves a CTA
// does the work of the NEXT phase that does not depend on the other CTAs (the recurrent half of the next layer).
__device__ __forceinline__ void lat_arrive(unsigned* counter, unsigned& epoch) {
    __syncthreads();
    if (threadIdx.x == 0) {
        epoch += gridDim.x;
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(counter) : "memory");
    }
}
__device__ __forceinline__ void lat_wait(unsigned* counter, unsigned epoch) {
    if (threadIdx.x == 0) {
        unsigned v = 0;
        for (unsigned spin = 0;; ++spin) {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
            if ((int)(v - epoch) >= 0) break;
            if (spin > (1u << 26)) __trap();
        }
    }
    __syncthreads();
}

__device__ __forceinline__ uint4 lat_philox(uint4 c, uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k0, lo1, hi0 ^ c.w ^ k1, lo0);
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return c;
}
__device__ __forceinline__ bool lat_better(float ka, int ia, float kb, int ib) { return (ka > kb) || (ka == kb && ia < ib); }

// acc[r][b] += dot(w_r[8 bf16], v_b[8 bf16]) for one 16-byte chunk
template <int LB>
__device__ __forceinline__ void fma_chunk(float (&acc)[4][LB], const uint4 (&w)[4], const uint4* __restrict__ vec, int ld_vec16, int kc) {
#pragma unroll
    for (int b = 0; b < LB; ++b) {
        const uint4 v = vec[b * ld_vec16 + kc];
        const __nv_bfloat162* vp = reinterpret_cast<const __nv_bfloat162*>(&v);
        float2 vf[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) vf[q] = __bfloat1622float2(vp[q]);
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const __nv_bfloat162* wp = reinterpret_cast<const __nv_bfloat162*>(&w[r]);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float2 wf = __bfloat1622float2(wp[q]);
                acc[r][b] = fmaf(wf.x, vf[q].x, acc[r][b]);
                acc[r][b] = fmaf(wf.y, vf[q].y, acc[r][b]);
            }
        }
    }
}

template <int LB>
__device__ __forceinline__ void warp_reduce_acc(float (&acc)[4][LB]) {
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int b = 0; b < LB; ++b) {
            float v = acc[r][b];
#pragma unroll
            for (int s = 16; s > 0; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
            acc[r][b] = v;
        }
}

// Sum acc[r][b] over the 32 lanes with a transposing butterfly: after it, lane (r, *) = lanes 8r .. 8r+7 hold row r's
// total.  6 shuffles per stream instead of 20: the first two stages halve the number of live values instead of
// reducing each of the four separately.
template <int LB>
__device__ __forceinline__ void warp_reduce_rows(float (&acc)[4][LB], float (&out)[LB], int lane) {
    const bool hi16 = lane & 16, hi8 = lane & 8;
#pragma unroll
    for (int b = 0; b < LB; ++b) {
        float k0 = hi16 ? acc[2][b] : acc[0][b], k1 = hi16 ? acc[3][b] : acc[1][b];
        const float s0 = hi16 ? acc[0][b] : acc[2][b], s1 = hi16 ? acc[1][b] : acc[3][b];
        k0 += __shfl_xor_sync(0xffffffffu, s0, 16);
        k1 += __shfl_xor_sync(0xffffffffu, s1, 16);
        float k = hi8 ? k1 : k0;
        const float sv = hi8 ? k0 : k1;
        k += __shfl_xor_sync(0xffffffffu, sv, 8);
        k += __shfl_xor_sync(0xffffffffu, k, 4);
        k += __shfl_xor_sync(0xffffffffu, k, 2);
        k += __shfl_xor_sync(0xffffffffu, k, 1);
        out[b] = k;
    }
}

__device__ __forceinline__ float sig_acc(float x) { return 1.0f / (1.0f + __expf(-x)); }

template <int LB, int LAT_THREADS>
__global__ void __launch_bounds__(LAT_THREADS, 1) lat_generate_kernel(const __grid_constant__ LatParams p) {
    constexpr int LAT_WARPS = LAT_THREADS / 32;
    extern __shared__ __align__(16) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int cta = blockIdx.x, ncta = gridDim.x;
    const int L = p.n_layers, bs = p.bs;
    uint4* vin = reinterpret_cast<uint4*>(smem + p.vec_off);       // [LB][maxIpad/8] layer input (or decoder input)
    uint4* vh = reinterpret_cast<uint4*>(smem + p.vh_off);         // [LB][maxHpad/8] previous hidden state
    float* cstate = reinterpret_cast<float*>(smem + p.state_off);  // per layer: [upc][LB]
    float* red = reinterpret_cast<float*>(smem + p.red_off);       // [LAT_WARPS][LB][4] + tokens
    int* tok_s = reinterpret_cast<int*>(red + LAT_WARPS * LB * 4);
    unsigned epoch = 0;
    unsigned long long tacc[6] = {0, 0, 0, 0, 0, 0}, tprev = 0;
    const bool dbg = p.debug && blockIdx.x == 0 && threadIdx.x == 0;
    auto stamp = [&](int k) { if (dbg) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); tacc[k] += t - tprev; tprev = t; } };

    // ---- one-time set-up: resident W_hh slices and this CTA's cell state
    for (int l = 0; l < L; ++l) {
        const int u0 = cta * p.upc[l], u1 = min(p.H[l], u0 + p.upc[l]);
        const int ldw16 = (p.Ipad[l] + p.Hpad[l]) / 8, h16 = p.Hpad[l] / 8, i16 = p.Ipad[l] / 8;
        if (p.wi_off[l] >= 0 && u1 > u0) {
            uint4* dst = reinterpret_cast<uint4*>(smem + p.wi_off[l]);
            const uint4* src = reinterpret_cast<const uint4*>(p.wcat[l]);
            const int rows = 4 * (u1 - u0);
            for (int i = tid; i < rows * i16; i += LAT_THREADS) {
                const int r = i / i16, k = i - r * i16;
                dst[i] = __ldg(src + (int64_t)(4 * u0 + r) * ldw16 + k);
            }
        }
        if (p.w_off[l] >= 0 && u1 > u0) {
            uint4* dst = reinterpret_cast<uint4*>(smem + p.w_off[l]);
            const uint4* src = reinterpret_cast<const uint4*>(p.wcat[l]);
            const int rows = 4 * (u1 - u0);
            for (int i = tid; i < rows * h16; i += LAT_THREADS) {
                const int r = i / h16, k = i - r * h16;
                dst[i] = __ldg(src + (int64_t)(4 * u0 + r) * ldw16 + i16 + k);
            }
        }
        for (int i = tid; i < p.upc[l] * LB; i += LAT_THREADS) {
            const int u = u0 + i / LB, b = i % LB;
            cstate[p.c_off[l] + i] = (u < u1 && b < bs) ? p.c[l][(int64_t)b * p.Hpad[l] + u] : 0.f;
        }
    }
    __syncthreads();

    const int T = p.P + p.n_new;
    int par = p.par0;
    // recurrent half computed ahead of a barrier wait: vh <- h of `layer` (buffer `hpar`), acc_pre <- W_hh slice x h for the warp's
    // first unit of that layer
    float acc_pre[4][LB];
    bool have_pre = false;
    auto hh_ahead = [&](int layer, int hpar) {
        const int i16 = p.Ipad[layer] / 8, h16 = p.Hpad[layer] / 8, ldw16 = i16 + h16;
        const int u0 = cta * p.upc[layer], u1 = min(p.H[layer], u0 + p.upc[layer]);
        const uint4* wg = reinterpret_cast<const uint4*>(p.wcat[layer]);
        const uint4* ws = p.w_off[layer] >= 0 ? reinterpret_cast<const uint4*>(smem + p.w_off[layer]) : nullptr;
        const uint4* src = reinterpret_cast<const uint4*>(p.hb[layer][hpar]);
        for (int i = tid; i < LB * h16; i += LAT_THREADS) {
            const int b = i / h16;
            vh[i] = (b < bs) ? __ldcg(src + i) : make_uint4(0, 0, 0, 0);
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int b = 0; b < LB; ++b) acc_pre[r][b] = 0.f;
        const int u = u0 + warp;
        if (u < u1) {
            for (int kc = lane; kc < h16; kc += 32) {
                uint4 w[4];
#pragma unroll
                for (int r = 0; r < 4; ++r)
                    w[r] = ws ? ws[(4 * (u - u0) + r) * h16 + kc] : __ldg(wg + (int64_t)(4 * u + r) * ldw16 + i16 + kc);
                fma_chunk<LB>(acc_pre, w, vh, h16, kc);
            }
        }
        have_pre = true;
    };
    if (dbg) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tprev));
    for (int s = 0; s < T; ++s) {
        // ---- token of every stream for this step, and its embedding row (K1: 128-bit gather)
        if (s < p.P && tid < bs) {
            int64_t t = p.prompt[(int64_t)s * bs + tid];
            if (t < 0 || t >= p.V) { atomicExch(p.err_flag, 1); t = 0; }
            tok_s[tid] = (int)t;
        }
        __syncthreads();
        const int e16 = p.Epad / 8;
        for (int i = tid; i < LB * e16; i += LAT_THREADS) {
            const int b = i / e16, k = i - b * e16;
            vin[b * e16 + k] = (b < bs) ? __ldg(reinterpret_cast<const uint4*>(p.table) + (int64_t)tok_s[b] * e16 + k)
                                        : make_uint4(0, 0, 0, 0);
        }
        stamp(0);
        // ---- the layers
        for (int l = 0; l < L; ++l) {
            const int i16 = p.Ipad[l] / 8, h16 = p.Hpad[l] / 8, ldw16 = i16 + h16;
            const int u0 = cta * p.upc[l], u1 = min(p.H[l], u0 + p.upc[l]);
            const uint4* wg = reinterpret_cast<const uint4*>(p.wcat[l]);
            const uint4* ws = p.w_off[l] >= 0 ? reinterpret_cast<const uint4*>(smem + p.w_off[l]) : nullptr;
            const uint4* wis = p.wi_off[l] >= 0 ? reinterpret_cast<const uint4*>(smem + p.wi_off[l]) : nullptr;
            // the first W_ih chunk of this warp's first unit does not depend on the inputs: fetch it from L2 while
            // the input vectors are still being staged
            uint4 wc[4];
            {
                const int u = u0 + warp;
#pragma unroll
                for (int r = 0; r < 4; ++r)
                    wc[r] = (wis == nullptr && u < u1 && lane < i16) ? __ldg(wg + (int64_t)(4 * u + r) * ldw16 + lane) : make_uint4(0, 0, 0, 0);
            }
            if (l > 0) {    // input = the layer below at this step
                const uint4* src = reinterpret_cast<const uint4*>(p.hb[l - 1][par ^ 1]);
                for (int i = tid; i < LB * i16; i += LAT_THREADS) {
                    const int b = i / i16;
                    vin[i] = (b < bs) ? __ldcg(src + i) : make_uint4(0, 0, 0, 0);
                }
            }
            if (!have_pre) {        // (otherwise vh already holds h(t-1) of this layer, and acc_pre the recurrent half of the warp's first unit)
                const uint4* src = reinterpret_cast<const uint4*>(p.hb[l][par]);
                for (int i = tid; i < LB * h16; i += LAT_THREADS) {
                    const int b = i / h16;
                    vh[i] = (b < bs) ? __ldcg(src + i) : make_uint4(0, 0, 0, 0);
                }
            }
            __syncthreads();
            for (int u = u0 + warp; u < u1; u += LAT_WARPS) {
                const bool pre = have_pre && u == u0 + warp;
                float acc[4][LB];
#pragma unroll
                for (int r = 0; r < 4; ++r)
#pragma unroll
                    for (int b = 0; b < LB; ++b) acc[r][b] = pre ? acc_pre[r][b] : 0.f;
                if (!pre) {
                    for (int kc = lane; kc < h16; kc += 32) {           // h * W_hh^T : weights resident in smem
                        uint4 w[4];
#pragma unroll
                        for (int r = 0; r < 4; ++r)
                            w[r] = ws ? ws[(4 * (u - u0) + r) * h16 + kc] : __ldg(wg + (int64_t)(4 * u + r) * ldw16 + i16 + kc);
                        fma_chunk<LB>(acc, w, vh, h16, kc);
                    }
                }
                if (wis != nullptr) {
                    for (int kc = lane; kc < i16; kc += 32) {           // x * W_ih^T : this layer's slice is resident too
                        uint4 w[4];
#pragma unroll
                        for (int r = 0; r < 4; ++r) w[r] = wis[(4 * (u - u0) + r) * i16 + kc];
                        fma_chunk<LB>(acc, w, vin, i16, kc);
                    }
                } else {
                if (u != u0 + warp) {                                   // (only the first unit was prefetched)
#pragma unroll
                    for (int r = 0; r < 4; ++r) wc[r] = (lane < i16) ? __ldg(wg + (int64_t)(4 * u + r) * ldw16 + lane) : make_uint4(0, 0, 0, 0);
                }
                for (int kc = lane; kc < i16; kc += 32) {               // x * W_ih^T : weights from L2, next chunk in flight
                    uint4 wn[4];
                    const bool more = kc + 32 < i16;
#pragma unroll
                    for (int r = 0; r < 4; ++r) wn[r] = more ? __ldg(wg + (int64_t)(4 * u + r) * ldw16 + kc + 32) : make_uint4(0, 0, 0, 0);
                    fma_chunk<LB>(acc, wc, vin, i16, kc);
#pragma unroll
                    for (int r = 0; r < 4; ++r) wc[r] = wn[r];
                }
                }
                warp_reduce_acc<LB>(acc);
                // lane b finishes stream b: gates (i,f,g,o) -> cell -> h
                float g4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int b = 0; b < LB; ++b)
                    if (lane == b) { g4[0] = acc[0][b]; g4[1] = acc[1][b]; g4[2] = acc[2][b]; g4[3] = acc[3][b]; }
                if (lane < bs && p.start != nullptr && s < __ldg(p.start + lane)) {
                    p.hb[l][par ^ 1][(int64_t)lane * p.Hpad[l] + u] = __float2bfloat16_rn(0.f);   // not begun: stays at reset
                } else if (lane < bs) {
                    const float4 bz = __ldg(reinterpret_cast<const float4*>(p.bias[l] + 4 * u));
                    float* cp = cstate + p.c_off[l] + (u - u0) * LB + lane;
                    const float ig = sig_acc(g4[0] + bz.x), fg = sig_acc(g4[1] + bz.y);
                    const float gg = tanhf(g4[2] + bz.z), og = sig_acc(g4[3] + bz.w);
                    const float cv = fg * (*cp) + ig * gg;
                    *cp = cv;
                    p.hb[l][par ^ 1][(int64_t)lane * p.Hpad[l] + u] = __float2bfloat16_rn(og * tanhf(cv));
                }
            }
            stamp(1);
            lat_arrive(p.sync_counter, epoch);
            // While the other CTAs finish this layer: the recurrent half of what this CTA does next.  Its input is a step old --
            // h(t-1) of layer l+1, or, behind the last layer of a step that has no decoder pass, h(t) of layer 0 (complete since
            // this step's first barrier) -- so it needs nobody.  (Behind the last layer of a decoding step the decoder runs next;
            // its recurrent hand-over happens at the decoder's own barrier.)
            have_pre = false;
            {
                const bool decode_next = (l == L - 1) && s >= p.P - 1 && s < T - 1;
                const int nl = l + 1 < L ? l + 1 : 0;
                const bool usable = (l + 1 < L) || (L > 1 && !decode_next && s + 1 < T);
                if (usable) hh_ahead(nl, l + 1 < L ? par : par ^ 1);
            }
            lat_wait(p.sync_counter, epoch);
            stamp(2);
        }
        par ^= 1;
        // ---- decoder + sampling race over this CTA's vocabulary rows (after the last prompt token and
        //      after every generated token but the last: main.py:56-71)
        if (s >= p.P - 1 && s < T - 1) {
            const int step = s - (p.P - 1);
            const uint4* src = reinterpret_cast<const uint4*>(p.hb[L - 1][par]);
            for (int i = tid; i < LB * e16; i += LAT_THREADS) {
                const int b = i / e16;
                vh[i] = (b < bs) ? __ldcg(src + i) : make_uint4(0, 0, 0, 0);
            }
            __syncthreads();
            const int rpc = (p.V + ncta - 1) / ncta;
            const int v0 = cta * rpc, v1 = min(p.V, v0 + rpc);
            const uint4* dw = reinterpret_cast<const uint4*>(p.dec_w);
            // A warp takes 32 rows at a time: eight groups of 4 rows go through the dot products and the transposing
            // reduction, each group hands its four totals to lanes 4j .. 4j+3, and then lane i owns row base + i for
            // every stream: one Philox call per (row, stream) per 32 rows instead of one (mostly masked) per 4 rows.
            float ka[LB], ko[LB];
            int ia[LB], io[LB];
#pragma unroll
            for (int b = 0; b < LB; ++b) { ka[b] = ko[b] = -INFINITY; ia[b] = io[b] = 0x7fffffff; }
            for (int base = v0 + 32 * warp; base < v1; base += 32 * LAT_WARPS) {
                float zrow[LB];
#pragma unroll
                for (int b = 0; b < LB; ++b) zrow[b] = 0.f;
                // software pipeline over the groups: the first 16-byte chunk of the NEXT group's four rows and the second
                // chunk of this group are in flight while this group's first chunk is multiplied
                uint4 wa[4];
#pragma unroll
                for (int r = 0; r < 4; ++r) wa[r] = (lane < e16) ? __ldg(dw + (int64_t)min(base + r, v1 - 1) * e16 + lane) : make_uint4(0, 0, 0, 0);
#pragma unroll 1
                for (int j = 0; j < 8; ++j) {
                    const int g0 = base + 4 * j;
                    if (g0 >= v1) break;                                  // warp-uniform
                    float acc[4][LB];
#pragma unroll
                    for (int r = 0; r < 4; ++r)
#pragma unroll
                        for (int b = 0; b < LB; ++b) acc[r][b] = 0.f;
                    const bool second = lane + 32 < e16;
                    uint4 wb[4], wn[4];
#pragma unroll
                    for (int r = 0; r < 4; ++r) wb[r] = second ? __ldg(dw + (int64_t)min(g0 + r, v1 - 1) * e16 + lane + 32) : make_uint4(0, 0, 0, 0);
                    const bool more = j + 1 < 8 && g0 + 4 < v1 && lane < e16;
#pragma unroll
                    for (int r = 0; r < 4; ++r) wn[r] = more ? __ldg(dw + (int64_t)min(g0 + 4 + r, v1 - 1) * e16 + lane) : make_uint4(0, 0, 0, 0);
                    if (lane < e16) fma_chunk<LB>(acc, wa, vh, e16, lane);
                    if (second) fma_chunk<LB>(acc, wb, vh, e16, lane + 32);
                    for (int kc = lane + 64; kc < e16; kc += 32) {       // wider embeddings than 512
                        uint4 w[4];
#pragma unroll
                        for (int r = 0; r < 4; ++r) w[r] = __ldg(dw + (int64_t)min(g0 + r, v1 - 1) * e16 + kc);
                        fma_chunk<LB>(acc, w, vh, e16, kc);
                    }
#pragma unroll
                    for (int r = 0; r < 4; ++r) wa[r] = wn[r];
                    float tot[LB];
                    warp_reduce_rows<LB>(acc, tot, lane);                 // lanes 8r .. 8r+7: row g0 + r
#pragma unroll
                    for (int b = 0; b < LB; ++b) {
                        const float v = __shfl_sync(0xffffffffu, tot[b], (lane & 3) * 8);
                        if ((lane >> 2) == j) zrow[b] = v;
                    }
                }
                const int id = base + lane;
                if (id < v1) {
                    const float bias = p.dec_bias ? __ldg(p.dec_bias + id) : 0.f;
                    int tb = -1;
                    for (int t = 0; t < p.sp.n_banned; ++t)
                        if (p.sp.banned[t] == id) tb = t;
#pragma unroll
                    for (int b = 0; b < LB; ++b) {
                        if (b >= bs) break;
                        float k = (zrow[b] + bias) * p.sp.inv_temperature;
                        if (!p.sp.greedy) {
                            const uint4 rn = lat_philox(make_uint4((uint32_t)id, (uint32_t)b, (uint32_t)step, 4u), p.sp.seed_lo, p.sp.seed_hi);
                            k += gumbel_from_bits(rn.x);
                        }
                        if (lat_better(k, id, ka[b], ia[b])) { ka[b] = k; ia[b] = id; }
                        if (tb >= 0) p.banned_keys[b * 64 + tb] = k;
                        else if (lat_better(k, id, ko[b], io[b])) { ko[b] = k; io[b] = id; }
                    }
                }
            }
            // combine the lanes of the warp (per stream), then the warps
#pragma unroll
            for (int b = 0; b < LB; ++b) {
#pragma unroll
                for (int sft = 16; sft > 0; sft >>= 1) {
                    float k = __shfl_xor_sync(0xffffffffu, ka[b], sft);
                    int i = __shfl_xor_sync(0xffffffffu, ia[b], sft);
                    if (lat_better(k, i, ka[b], ia[b])) { ka[b] = k; ia[b] = i; }
                    k = __shfl_xor_sync(0xffffffffu, ko[b], sft);
                    i = __shfl_xor_sync(0xffffffffu, io[b], sft);
                    if (lat_better(k, i, ko[b], io[b])) { ko[b] = k; io[b] = i; }
                }
                if (lane == 0) {
                    float* r4 = red + (warp * LB + b) * 4;
                    r4[0] = ka[b]; r4[1] = __int_as_float(ia[b]); r4[2] = ko[b]; r4[3] = __int_as_float(io[b]);
                }
            }
            __syncthreads();
            if (tid < bs) {
                float fa = -INFINITY, fo = -INFINITY;
                int ja = 0x7fffffff, jo = 0x7fffffff;
                for (int w = 0; w < LAT_WARPS; ++w) {
                    const float* r4 = red + (w * LB + tid) * 4;
                    if (lat_better(r4[0], __float_as_int(r4[1]), fa, ja)) { fa = r4[0]; ja = __float_as_int(r4[1]); }
                    if (lat_better(r4[2], __float_as_int(r4[3]), fo, jo)) { fo = r4[2]; jo = __float_as_int(r4[3]); }
                }
                p.part[cta * 8 + tid] = make_float4(fa, __int_as_float(ja), fo, __int_as_float(jo));
            }
            stamp(3);
            lat_arrive(p.sync_counter, epoch);
            if (L > 1) hh_ahead(0, par);          // next step's layer 0: its h(t) has been complete since this step's first barrier
            lat_wait(p.sync_counter, epoch);
            stamp(4);
            // every CTA resolves the same winner: warp b merges the per-CTA partials of stream b
            if (warp < bs) {
                float ka = -INFINITY, ko = -INFINITY;
                int ia = 0x7fffffff, io = 0x7fffffff;
                for (int c = lane; c < ncta; c += 32) {
                    const float4 q = __ldcg(p.part + c * 8 + warp);
                    if (lat_better(q.x, __float_as_int(q.y), ka, ia)) { ka = q.x; ia = __float_as_int(q.y); }
                    if (lat_better(q.z, __float_as_int(q.w), ko, io)) { ko = q.z; io = __float_as_int(q.w); }
                }
                int outrank = 0;
                float kb[2] = {-INFINITY, -INFINITY};
#pragma unroll
                for (int u = 0; u < 2; ++u)
                    if (lane + 32 * u < p.sp.n_banned && p.sp.banned[lane + 32 * u] < p.V)
                        kb[u] = __ldcg(p.banned_keys + warp * 64 + lane + 32 * u);
#pragma unroll
                for (int sft = 16; sft > 0; sft >>= 1) {
                    float k = __shfl_xor_sync(0xffffffffu, ka, sft);
                    int i = __shfl_xor_sync(0xffffffffu, ia, sft);
                    if (lat_better(k, i, ka, ia)) { ka = k; ia = i; }
                    k = __shfl_xor_sync(0xffffffffu, ko, sft);
                    i = __shfl_xor_sync(0xffffffffu, io, sft);
                    if (lat_better(k, i, ko, io)) { ko = k; io = i; }
                }
#pragma unroll
                for (int u = 0; u < 2; ++u)
                    if (lane + 32 * u < p.sp.n_banned && p.sp.banned[lane + 32 * u] < p.V)
                        outrank += lat_better(kb[u], p.sp.banned[lane + 32 * u], ko, io) ? 1 : 0;
#pragma unroll
                for (int sft = 16; sft > 0; sft >>= 1) outrank += __shfl_xor_sync(0xffffffffu, outrank, sft);
                // main.py:63-68: first allowed of the n_draws candidates, else the first candidate
                const int pick = (io == 0x7fffffff || outrank >= p.sp.n_draws) ? ia : io;
                if (lane == 0) {
                    tok_s[warp] = pick;
                    if (cta == 0) p.gen[(int64_t)step * bs + warp] = pick;
                }
            }
            __syncthreads();
            stamp(5);
        }
    }
    if (dbg) printf("lat kernel (CTA 0) us/step: gather %.2f | layers compute %.2f sync %.2f | decoder %.2f sync %.2f merge %.2f\n", tacc[0] * 1e-3 / T, tacc[1] * 1e-3 / T, tacc[2] * 1e-3 / T, tacc[3] * 1e-3 / T, tacc[4] * 1e-3 / T, tacc[5] * 1e-3 / T);
    // ---- hand the cell state back
    for (int l = 0; l < L; ++l) {
        const int u0 = cta * p.upc[l], u1 = min(p.H[l], u0 + p.upc[l]);
        for (int i = tid; i < p.upc[l] * LB; i += LAT_THREADS) {
            const int u = u0 + i / LB, b = i % LB;
            if (u < u1 && b < bs) p.c[l][(int64_t)b * p.Hpad[l] + u] = cstate[p.c_off[l] + i];
        }
    }
}

int lat_max_streams() { return 8; }

int launch_lat_generate(const LatArgs& a, cudaStream_t st) {
    if (a.bs < 1 || a.bs > 8 || a.n_layers < 1 || a.n_layers > 4) return set_error(AWD_ELIMIT, "latency kernel: 1..8 streams, 1..4 layers");
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int LB = a.bs == 1 ? 1 : a.bs == 2 ? 2 : a.bs <= 4 ? 4 : 8;
    LatParams p{};
    p.n_layers = a.n_layers; p.bs = a.bs; p.V = a.V; p.P = a.P; p.n_new = a.n_new; p.Epad = a.Epad; p.par0 = a.par0;
    int maxI = a.Epad, maxH = a.Epad, states = 0;
    for (int l = 0; l < a.n_layers; ++l) {
        p.H[l] = a.H[l]; p.Ipad[l] = a.Ipad[l]; p.Hpad[l] = a.Hpad[l];
        p.upc[l] = (a.H[l] + sms - 1) / sms;
        p.wcat[l] = (const __nv_bfloat16*)a.wcat[l]; p.bias[l] = a.bias[l]; p.c[l] = a.c[l];
        p.hb[l][0] = (__nv_bfloat16*)a.hb[l][0]; p.hb[l][1] = (__nv_bfloat16*)a.hb[l][1];
        maxI = std::max(maxI, a.Ipad[l]); maxH = std::max(maxH, a.Hpad[l]);
        p.c_off[l] = states;
        states += p.upc[l] * LB;
    }
    // shared-memory plan: vectors, state and reduction scratch first; W_hh slices while they fit
    int off = 0;
    p.vec_off = off; off += LB * maxI * 2;
    p.vh_off = off; off += LB * maxH * 2;
    p.state_off = off; off += ((states * 4 + 15) / 16) * 16;
    p.red_off = off; off += LAT_MAX_WARPS * LB * 16 + 64;
    off = (off + 15) / 16 * 16;
    const int budget = 227 * 1024;
    for (int l = 0; l < a.n_layers; ++l) {
        const int bytes = 4 * p.upc[l] * a.Hpad[l] * 2;
        if (off + bytes <= budget) { p.w_off[l] = off; off += bytes; }
        else p.w_off[l] = -1;            // does not fit: this layer's W_hh streams from L2 like W_ih
    }
    for (int l = 0; l < a.n_layers; ++l) {   // what is left goes to W_ih slices (shipped size: layer 0's 28 KB fits)
        const int bytes = 4 * p.upc[l] * a.Ipad[l] * 2;
        if (p.w_off[l] >= 0 && off + bytes <= budget) { p.wi_off[l] = off; off += bytes; }
        else p.wi_off[l] = -1;
    }
    p.table = (const __nv_bfloat16*)a.table; p.dec_w = (const __nv_bfloat16*)a.dec_w; p.dec_bias = a.dec_bias;
    p.prompt = a.prompt; p.gen = a.gen; p.sp = a.sp; p.part = (float4*)a.part; p.banned_keys = a.banned_keys;
    p.sync_counter = a.sync_counter; p.err_flag = a.err_flag; p.start = a.start;
    { static int dbg = -1; if (dbg < 0) { const char* s = getenv("AWD_LAT_DEBUG"); dbg = s ? atoi(s) : 0; } p.debug = dbg; }
    AWD_CUDA(cudaMemsetAsync(a.sync_counter, 0, sizeof(unsigned), st));
    // more warps = more independent weight loads in flight per SM (the decoder pass is load-latency bound);
    // the 4- and 8-stream variants need too many registers for 512 threads
    void (*kern)(LatParams) = LB == 1 ? lat_generate_kernel<1, 512> : LB == 2 ? lat_generate_kernel<2, 512>
                              : LB == 4 ? lat_generate_kernel<4, 256> : lat_generate_kernel<8, 256>;
    const int nthreads = LB <= 2 ? 512 : 256;
    AWD_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, budget));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(sms);
    cfg.blockDim = dim3(nthreads);
    cfg.dynamicSmemBytes = off;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    AWD_CUDA(cudaLaunchKernelEx(&cfg, kern, p));
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

}  // namespace tc
}  // namespace awd
