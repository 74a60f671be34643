// Blackwell (sm_100a) tensor-core GEMM of the AWD-LSTM engine: hand-written tcgen05.mma with TMEM
// accumulators, operands staged by TMA into 128B-swizzled shared memory, mbarrier pipelines.
//
//   D[M][N] = A[M][K] * W[N][K]^T        bf16 operands (both K-major), fp32 accumulate in TMEM
//
// One persistent CTA per SM, 320 threads:
//   warp 0      TMA producer   (one elected lane issues cp.async.bulk.tensor into a STAGES-deep ring)
//   warp 1      MMA issuer     (one elected lane issues tcgen05.mma 128 x BN x 16; owns the TMEM allocation)
//   warps 2..9  epilogue       (tcgen05.ld 32 lanes x 32 columns -> registers; thread = output row,
//                               EPI_PARTS warps per TMEM lane quarter, each taking a part of the tile's columns)
// The accumulator is double-buffered in TMEM (2 x BN columns), so the epilogue of tile i overlaps the
// MMAs of tile i+1.  The epilogue is a functor; three are used by the engine:
//   TcEpiStore   C = acc + bias                        (input projection of a whole prompt, decoder logits)
//   TcEpiCell    gates -> sigmoid/tanh -> c,h update    (one LSTM step; weight rows are gate-interleaved so a
//                                                       thread's 4 adjacent columns are i,f,g,o of one unit)
//   TcEpiSample  exponential-race keys + running best   (decoder fused with the sampling rule: logits are
//                                                       consumed from TMEM and never written to HBM)
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"

namespace awd {
namespace tc {

constexpr int BM = 128;       // rows per tile = TMEM lanes = UMMA M
constexpr int BK = 64;        // bf16 elements per 128-byte swizzle row
constexpr int UMMA_K = 16;
// 8 epilogue warps, 2 per TMEM lane quarter.  12 warps (3 column parts) were measured: the 128-register cap
// they force spills in the sampling epilogue and the decoder phase got 25 % slower.
constexpr int NUM_EPI_THREADS = 256;
constexpr int EPI_PARTS = NUM_EPI_THREADS / 128;     // column parts a tile is split into
// Store epilogue: every epilogue warp stages 32 rows x 32 fp32 columns (4 KB, 128-byte swizzle) in one of two buffers
// and hands it to a TMA tensor store; other epilogues need no scratch.
constexpr int STORE_WARP_SCRATCH = 2 * 4096;
constexpr int NUM_THREADS = 64 + NUM_EPI_THREADS;    // + TMA producer warp + MMA issuer warp
constexpr int A_STAGE_BYTES = BM * BK * 2;

template <int BN> struct TileCfg {
    static constexpr int B_STAGE_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
    static constexpr int STAGES = (BN >= 256) ? 4 : 6;
    static constexpr int TMEM_COLS = 2 * BN;                      // power of two >= 32 for BN in {64,128,256}
    static constexpr int TAIL_BYTES = 1024 /*alignment slack*/ + 256 /*barriers*/ + 2 * BN * 4 /*bias*/;   // + Epi::kScratchBytes
    static constexpr int SMEM_MAX = 227 * 1024;
};

struct GemmShape {
    int M, N;
    int kb0, kb1;          // 64-wide K blocks taken from operand A0, then A1
    int a0_row, a1_row;    // row offsets into the A tensor maps
    int b_k0;              // first K element of the weight map
    int tiles_m, tiles_n;
    int a_res;             // 1: A-resident schedule (whole K of one M tile kept in shared memory)
    int row_stat;          // 1: row-stationary schedule (CTA keeps one M tile, walks N tiles); implied by a_res
    int stages;            // depth of the TMA ring (<= 8)
    int col_w;             // persistent kernel, row-stationary phase: columns per CTA group (multiple of 32; 0 = whole tiles dealt round-robin)
    int x_n0, x_w;         // ... and the column range [x_n0, x_n0 + x_w) multiplied by the CTAs left over after the last whole group (PhaseTiles)
};

// ---------------------------------------------------------------------------------------- PTX
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Bounded spin: a protocol bug traps (CUDA error) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    for (uint32_t spin = 0; !done; ++spin) {
        asm volatile(
            "{\n.reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (spin > (1u << 26)) __trap();
    }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
// Pull a tensor map (kernel parameter / constant space) into the TMA unit's descriptor cache ahead of its first use
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* m) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
// Ask for a tensor box to be brought into L2 (no shared-memory destination, nothing to wait for)
__device__ __forceinline__ void tma_prefetch_l2_2d(const CUtensorMap* map, int c0, int c1) {
    asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];" ::"l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, uint32_t src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(map)), "r"(src), "r"(c0), "r"(c1) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template <int N> __device__ __forceinline__ void tma_store_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
// L2 eviction-priority hints (the encodings CUTLASS uses for createpolicy results)
constexpr uint64_t L2_EVICT_FIRST = 0x12F0000000000000ull, L2_EVICT_LAST = 0x14F0000000000000ull;
__device__ __forceinline__ void tma_load_2d_hint(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;"
        ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "l"(policy)
        : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum)
        : "memory");
}
// The MMA-issuing warp runs its loop with all 32 lanes (uniform values, no divergence) and ONE ELECTED lane issues each tcgen05
// instruction.  Issued from inside `if (lane == 0)`, every tcgen05.mma was wrapped by the compiler in an ELECT / 7 x R2UR.BROADCAST /
// BRA.U.ANY loop (its operands live in uniform registers): ~17 instructions per MMA, which made a 128 x N x 16 MMA cost 119 clocks
// for every N <= 128 and let busy epilogue warps on the same SM sub-partition slow the main loop by 20 % (profiles/micro/kb_probe.cu:
// 474 -> 395 clocks per K block at N <= 128; 620 -> 512 at N = 256 next to integer-heavy warps).
__device__ __forceinline__ void umma_bf16_elect(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n.reg .pred p, q;\nsetp.ne.b32 p, %4, 0;\nelect.sync _|q, 0xffffffff;\n"
        "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum)
        : "memory");
}
__device__ __forceinline__ void umma_commit_elect(uint32_t bar) {
    asm volatile("{\n.reg .pred q;\nelect.sync _|q, 0xffffffff;\n@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n}"
                 ::"r"(bar) : "memory");
}
// ... and the producer's instructions the same way (UTMALDG takes its operands from uniform registers too)
__device__ __forceinline__ void mbar_expect_tx_elect(uint32_t bar, uint32_t bytes) {
    asm volatile("{\n.reg .pred q;\nelect.sync _|q, 0xffffffff;\n@q mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n}" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_2d_elect(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
    asm volatile(
        "{\n.reg .pred q;\nelect.sync _|q, 0xffffffff;\n"
        "@q cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n}"
        ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d_hint_elect(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar, uint64_t policy) {
    asm volatile(
        "{\n.reg .pred q;\nelect.sync _|q, 0xffffffff;\n"
        "@q cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;\n}"
        ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "l"(policy)
        : "memory");
}
// K-major operand tile in shared memory, 128-byte swizzle: rows are 128 B, 8-row groups 1024 B apart.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int n) {
    // c=f32 (bit 4), a=b=bf16 (bits 7, 10), K-major A and B, N>>3 at bit 17, M>>4 at bit 24
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}
// ------------------------------------------------------------------------------------- kernel
// Epi interface (const functor + per-thread State):
//   const float* bias;                       // [N] or nullptr -- staged in shared memory per tile, times bias_scale()
//   bool stage_bias() const;                 // false: skip the staging and its epilogue-wide barrier (CTA-uniform)
//   static constexpr int kScratchBytes;      // shared memory the kernel sets aside for the epilogue warps (split evenly)
//   void drain() const;                      // called by every epilogue thread after its last tile
//   void tile_begin(State&, int row, bool row_ok, int n_first /* first column this thread will see in the tile */,
//                   int nchunks /* 32-column chunks it will see */) const;   // called BEFORE the accumulator is ready
//   void chunk(State&, int row, bool row_ok, int n /*first of 32 columns*/, int N, const float (&v)[32],
//              const float* sbias /*shared memory, 32 floats for this chunk*/, float* warp_scratch) const;
//   void tile_end(State&, int row, bool row_ok, int slot /* = tile_n * EPI_PARTS + column part */) const;
//   void init(State&) const;                 // once per thread, before the first tile
//   void finish(State&, int row, bool row_ok, int slot) const;   // row-stationary schedules only, after the last tile
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(NUM_EPI_THREADS) : "memory"); }

// Static persistent schedule shared by the three warp roles.  Streaming mode: tiles in M-fastest order.
// A-resident mode (g.a_res): the CTA keeps ONE M tile's A operand (all of K) in shared memory and walks
// N tiles, so only the weights stream -- used when K is small enough (the decoder: K = emb_sz), where it
// cuts the bytes each SM must ingest per tile by a third.
// `pair`: the two CTAs of a tc_gemm_pair_kernel cluster take adjacent M tiles of the same N tile.
struct TileIter {
    int tm, tn;
    int cur, step, end, tiles_mc, cm_, cmi;
    bool res;
    __device__ __forceinline__ TileIter(const GemmShape& g, int rank, bool pair) {
        res = g.a_res != 0 || g.row_stat != 0;
        cm_ = pair ? 2 : 1; cmi = rank;
        tiles_mc = (g.tiles_m + cm_ - 1) / cm_;
        if (res) {
            tm = blockIdx.x % g.tiles_m;
            cur = blockIdx.x / g.tiles_m; step = gridDim.x / g.tiles_m; end = g.tiles_n;
        } else {
            cur = blockIdx.x / cm_; step = gridDim.x / cm_;
            end = tiles_mc * g.tiles_n;
        }
    }
    // explicit CTA index / count (persistent multi-phase kernels)
    __device__ __forceinline__ TileIter(const GemmShape& g, int cta, int ncta, bool /*tag*/) {
        res = g.row_stat != 0;
        cm_ = 1; cmi = 0; tiles_mc = g.tiles_m;
        if (res) { tm = cta % g.tiles_m; cur = cta / g.tiles_m; step = ncta / g.tiles_m; end = g.tiles_n; }
        else { cur = cta; step = ncta; end = g.tiles_m * g.tiles_n; }
    }
    __device__ __forceinline__ bool valid() {
        if (cur >= end) return false;
        if (res) tn = cur;
        else { tm = (cur % tiles_mc) * cm_ + cmi; tn = cur / tiles_mc; }
        return true;
    }
    __device__ __forceinline__ void next() { cur += step; }
};

template <int BN, class Epi>
__global__ void __launch_bounds__(NUM_THREADS, 1)
tc_gemm_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
               const __grid_constant__ CUtensorMap tmB, const GemmShape g, const __grid_constant__ Epi epi) {
    using Cfg = TileCfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    const int num_kb = g.kb0 + g.kb1;
    const int STAGES = g.stages;
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t a_bytes = (g.a_res ? num_kb : STAGES) * A_STAGE_BYTES;
    const uint32_t sA = base, sB = base + a_bytes;
    const uint32_t scr0 = sB + STAGES * Cfg::B_STAGE_BYTES;          // epilogue scratch (1024-aligned), then the barriers
    const uint32_t bars = scr0 + Epi::kScratchBytes;
    const uint32_t full0 = bars, empty0 = bars + 64, tfull0 = bars + 128, tempty0 = bars + 144, afull = bars + 160,
                   tmem_slot = bars + 168;
    uint8_t* gen_base = smem_raw + (base - smem_u32(smem_raw));
    uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(gen_base + (tmem_slot - base));
    float* sbias = reinterpret_cast<float*>(gen_base + (bars - base) + 256);   // [2][BN]
    float* wscr = (Epi::kScratchBytes && (threadIdx.x >> 5) >= 2)               // per epilogue warp
                      ? reinterpret_cast<float*>(gen_base + (scr0 - base) + ((threadIdx.x >> 5) - 2) * (Epi::kScratchBytes / (NUM_EPI_THREADS / 32)))
                      : nullptr;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(full0 + 8 * s, 1);
            mbar_init(empty0 + 8 * s, 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(tfull0 + 8 * a, 1);
            mbar_init(tempty0 + 8 * a, NUM_EPI_THREADS / 32);
        }
        mbar_init(afull, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            TileIter it(g, 0, false);
            if (g.a_res && it.valid()) {       // the whole A operand of this CTA's M tile, once
                mbar_expect_tx(afull, num_kb * A_STAGE_BYTES);
                for (int kb = 0; kb < num_kb; ++kb) {
                    if (kb < g.kb0) tma_load_2d(sA + kb * A_STAGE_BYTES, &tmA0, kb * BK, g.a0_row + it.tm * BM, afull);
                    else tma_load_2d(sA + kb * A_STAGE_BYTES, &tmA1, (kb - g.kb0) * BK, g.a1_row + it.tm * BM, afull);
                }
            }
            for (; it.valid(); it.next()) {
                const int m0 = it.tm * BM, n0 = it.tn * BN;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(empty0 + 8 * stage, phase ^ 1);
                    const uint32_t fb = full0 + 8 * stage;
                    const uint32_t da = sA + stage * A_STAGE_BYTES, db = sB + stage * Cfg::B_STAGE_BYTES;
                    if (g.a_res) {
                        mbar_expect_tx(fb, Cfg::B_STAGE_BYTES);
                    } else {
                        mbar_expect_tx(fb, Cfg::STAGE_BYTES);
                        if (kb < g.kb0) tma_load_2d(da, &tmA0, kb * BK, g.a0_row + m0, fb);
                        else tma_load_2d(da, &tmA1, (kb - g.kb0) * BK, g.a1_row + m0, fb);
                    }
                    tma_load_2d(db, &tmB, g.b_k0 + kb * BK, n0, fb);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        {   // all 32 lanes, one elected issuer per instruction (umma_bf16_elect)
            constexpr uint32_t idesc = umma_idesc_bf16(BN);
            int stage = 0, acc = 0;
            uint32_t phase = 0, acc_phase = 0;
            TileIter it(g, 0, false);
            if (g.a_res && it.valid()) mbar_wait(afull, 0);
            for (; it.valid(); it.next()) {
                mbar_wait(tempty0 + 8 * acc, acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(full0 + 8 * stage, phase);
                    tc_fence_after();
                    const uint64_t adesc = umma_desc_sw128(sA + (g.a_res ? kb : stage) * A_STAGE_BYTES);
                    const uint64_t bdesc = umma_desc_sw128(sB + stage * Cfg::B_STAGE_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k)   // +32 bytes per K step inside the swizzle row
                        umma_bf16_elect(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
                    umma_commit_elect(empty0 + 8 * stage);       // frees the smem slot when the MMAs retire
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                umma_commit_elect(tfull0 + 8 * acc);         // accumulator complete -> epilogue
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else {
        // epilogue warps: EPI_PARTS per TMEM lane quarter (hardware: a warp may read lanes 32*(warp%4)..+31),
        // each taking one part of the tile's columns
        const int ew = warp - 2;
        const int q = warp & 3, part = ew >> 2;                 // TMEM lane quarter, column part
        const int c0 = ((BN / 32) * part / EPI_PARTS) * 32, c1 = ((BN / 32) * (part + 1) / EPI_PARTS) * 32;
        const int et = ew * 32 + lane;                       // 0..255
        int acc = 0;
        uint32_t acc_phase = 0;
        typename Epi::State es;
        epi.init(es);
        for (TileIter it(g, 0, false); it.valid(); it.next()) {
            const int tm = it.tm, tn = it.tn;
            const int row = tm * BM + q * 32 + lane;
            const bool row_ok = row < g.M;
            // stage this tile's bias while the MMAs are still running
            float* sb = sbias + acc * BN;
            if (epi.stage_bias()) {                       // uniform over the CTA
                if (et < BN) {
                    const int n = tn * BN + et;
                    sb[et] = (epi.bias != nullptr && n < g.N) ? __ldg(epi.bias + n) * epi.bias_scale() : 0.f;
                }
                epi_bar_sync();
            }
            epi.tile_begin(es, row, row_ok, tn * BN + c0, (c1 - c0) / 32);
            mbar_wait(tfull0 + 8 * acc, acc_phase);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN;
#pragma unroll 1
            for (int c = c0; c < c1; c += 32) {
                float v[32];
                tmem_ld32(taddr + c, v);
                if (tn * BN + c < g.N) epi.chunk(es, row, row_ok, tn * BN + c, g.N, v, sb + c, wscr);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty0 + 8 * acc);
            epi.tile_end(es, row, row_ok, tn * EPI_PARTS + part);
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if (g.a_res || g.row_stat) {        // row-stationary: one result per (row, CTA, column part)
            const int row = (blockIdx.x % g.tiles_m) * BM + q * 32 + lane;
            epi.finish(es, row, row < g.M, (blockIdx.x / g.tiles_m) * EPI_PARTS + part);
        }
        epi.drain();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
    }
}

// ------------------------------------------------------------------------ CTA-pair variant
// cta_group::2: two CTAs of a cluster (two SMs) run ONE 256 x 256 x 16 MMA per instruction.  Each CTA
// stages its own 128 A rows and only HALF of the B tile (128 of the 256 weight rows); the tensor cores
// of both SMs read both halves.  Per SM this halves the weight bytes that must come out of L2 and cuts
// the shared-memory read traffic per MMA by a third -- the two things the single-CTA kernel is bound by.
//   * the leader CTA (cluster rank 0) issues every MMA; TMA loads of BOTH CTAs signal the leader's `full`
//     barrier (cp.async.bulk.tensor .cta_group::2), tcgen05.commit multicasts to both CTAs' `empty`
//     (smem slot free) and `tfull` (accumulator ready) barriers, and the epilogue warps of both CTAs
//     arrive on the leader's `tempty` barrier.
//   * each CTA's epilogue drains its own 128 TMEM lanes, exactly as in the single-CTA kernel.
__device__ __forceinline__ uint32_t mapa_rank(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar_cluster) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar_cluster), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t dst_smem, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum)
        : "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)3) : "memory");
}

constexpr int PAIR_BN = 256;
constexpr int PAIR_B_BYTES = (PAIR_BN / 2) * BK * 2;          // half of the weight tile per CTA
constexpr int PAIR_STAGE_BYTES = A_STAGE_BYTES + PAIR_B_BYTES;
constexpr int PAIR_TAIL_BYTES = 1024 + 256 + 2 * PAIR_BN * 4;
constexpr int PAIR_SMEM_MAX = 227 * 1024;

template <class Epi>
__global__ void __launch_bounds__(NUM_THREADS, 1)
tc_gemm_pair_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                    const __grid_constant__ CUtensorMap tmB, const GemmShape g, const Epi epi) {
    constexpr int BN = PAIR_BN;
    extern __shared__ uint8_t smem_raw[];
    const int num_kb = g.kb0 + g.kb1;
    const int STAGES = g.stages;
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t a_bytes = (g.a_res ? num_kb : STAGES) * A_STAGE_BYTES;
    const uint32_t sA = base, sB = base + a_bytes;
    const uint32_t bars = sB + STAGES * PAIR_B_BYTES;
    const uint32_t full0 = bars, empty0 = bars + 64, tfull0 = bars + 128, tempty0 = bars + 144, afull = bars + 160,
                   tmem_slot = bars + 168;
    uint8_t* gen_base = smem_raw + (base - smem_u32(smem_raw));
    uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(gen_base + (tmem_slot - base));
    float* sbias = reinterpret_cast<float*>(gen_base + (bars - base) + 256);   // [2][BN]
    float* wscr = nullptr;                 // (the pair kernel gives the store epilogue no staging: it falls back to direct stores)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();
    const bool leader = rank == 0;

    if (warp == 0 && lane == 0) {
        for (int s = 0; s < STAGES; ++s) {
            // the leader's copy is used: ONE arrival (the leader's expect_tx, which announces the bytes of BOTH
            // CTAs); the peer's TMA only completes transactions on it.  A per-stage remote arrive from the peer
            // cost ~1300 clocks per K block (measured) and serialised its producer.
            mbar_init(full0 + 8 * s, 1);
            mbar_init(empty0 + 8 * s, 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(tfull0 + 8 * a, 1);
            mbar_init(tempty0 + 8 * a, 2 * (NUM_EPI_THREADS / 32));   // epilogue warps of both CTAs
        }
        mbar_init(afull, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) tmem_alloc_pair(tmem_slot, 2 * BN);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            TileIter it(g, rank, true);
            if (g.a_res && it.valid()) {       // the whole A operand of this CTA's M tile, once
                const uint32_t af = mapa_rank(afull, 0);
                if (leader) mbar_expect_tx(afull, 2 * num_kb * A_STAGE_BYTES);
                for (int kb = 0; kb < num_kb; ++kb) {
                    if (kb < g.kb0) tma_load_2d_pair(sA + kb * A_STAGE_BYTES, &tmA0, kb * BK, g.a0_row + it.tm * BM, af);
                    else tma_load_2d_pair(sA + kb * A_STAGE_BYTES, &tmA1, (kb - g.kb0) * BK, g.a1_row + it.tm * BM, af);
                }
            }
            const uint32_t stage_tx = 2 * ((g.a_res ? 0 : A_STAGE_BYTES) + PAIR_B_BYTES);
            for (; it.valid(); it.next()) {
                const int m0 = it.tm * BM, n0 = it.tn * BN + rank * (BN / 2);
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(empty0 + 8 * stage, phase ^ 1);
                    const uint32_t fb = mapa_rank(full0 + 8 * stage, 0);
                    if (leader) mbar_expect_tx(full0 + 8 * stage, stage_tx);
                    if (!g.a_res) {
                        const uint32_t da = sA + stage * A_STAGE_BYTES;
                        if (kb < g.kb0) tma_load_2d_pair(da, &tmA0, kb * BK, g.a0_row + m0, fb);
                        else tma_load_2d_pair(da, &tmA1, (kb - g.kb0) * BK, g.a1_row + m0, fb);
                    }
                    tma_load_2d_pair(sB + stage * PAIR_B_BYTES, &tmB, g.b_k0 + kb * BK, n0, fb);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && leader) {
            // M = 256 (two CTAs x 128 lanes), N = 256
            constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
            int stage = 0, acc = 0;
            uint32_t phase = 0, acc_phase = 0;
            TileIter it(g, rank, true);
            if (g.a_res && it.valid()) mbar_wait(afull, 0);
            for (; it.valid(); it.next()) {
                mbar_wait(tempty0 + 8 * acc, acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(full0 + 8 * stage, phase);
                    tc_fence_after();
                    const uint64_t adesc = umma_desc_sw128(sA + (g.a_res ? kb : stage) * A_STAGE_BYTES);
                    const uint64_t bdesc = umma_desc_sw128(sB + stage * PAIR_B_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k)
                        umma_bf16_pair(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
                    umma_commit_pair(empty0 + 8 * stage);    // slot free in both CTAs once the MMAs retire
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                umma_commit_pair(tfull0 + 8 * acc);          // accumulator halves complete in both CTAs
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else {
        const int ew = warp - 2;
        const int q = warp & 3, part = ew >> 2;                 // TMEM lane quarter, column part
        const int c0 = ((BN / 32) * part / EPI_PARTS) * 32, c1 = ((BN / 32) * (part + 1) / EPI_PARTS) * 32;
        const int et = ew * 32 + lane;
        int acc = 0;
        uint32_t acc_phase = 0;
        typename Epi::State es;
        epi.init(es);
        const uint32_t te0 = mapa_rank(tempty0, 0);
        for (TileIter it(g, rank, true); it.valid(); it.next()) {
            const int tm = it.tm, tn = it.tn;
            const int row = tm * BM + q * 32 + lane;
            const bool row_ok = row < g.M && tn < g.tiles_n;
            float* sb = sbias + acc * BN;
            if (epi.stage_bias()) {                       // uniform over the CTA
                if (et < BN) {
                    const int n = tn * BN + et;
                    sb[et] = (epi.bias != nullptr && n < g.N) ? __ldg(epi.bias + n) * epi.bias_scale() : 0.f;
                }
                epi_bar_sync();
            }
            epi.tile_begin(es, row, row_ok, tn * BN + c0, (c1 - c0) / 32);
            mbar_wait(tfull0 + 8 * acc, acc_phase);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN;
#pragma unroll 1
            for (int c = c0; c < c1; c += 32) {
                float v[32];
                tmem_ld32(taddr + c, v);
                if (tn * BN + c < g.N) epi.chunk(es, row, row_ok, tn * BN + c, g.N, v, sb + c, wscr);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(te0 + 8 * acc);
            epi.tile_end(es, row, row_ok, tn * EPI_PARTS + part);
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if (g.a_res || g.row_stat) {
            const int row = (blockIdx.x % g.tiles_m) * BM + q * 32 + lane;
            epi.finish(es, row, row < g.M, (blockIdx.x / g.tiles_m) * EPI_PARTS + part);
        }
        epi.drain();
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_pair(tmem_base, 2 * BN);
    }
}

// ---------------------------------------------------------------------------------- epilogues
struct TcEpiStore {
    float* C;              // [M][ldc] fp32
    int64_t ldc;
    const float* bias;     // [N] or nullptr
    int M;
    // A thread owns one output row in TMEM, so a direct store would put 32 lanes on 32 different rows (16 of every
    // 32-byte sector per instruction).  Instead the warp writes its 32 x 32 block into shared memory in the 128-byte
    // swizzle pattern (conflict-free 16-byte stores) and one lane hands it to a TMA tensor store: full 128-byte
    // lines, edges clipped by the tensor map, the next block staged while the previous one is still being read.
    CUtensorMap tmC;       // fp32 [M][N] pitch ldc, box 32 x 32, 128-byte swizzle; valid iff use_tma
    int use_tma;
    static constexpr int kScratchBytes = (NUM_EPI_THREADS / 32) * STORE_WARP_SCRATCH;
    struct State { int pending; int n[2]; float* scr; };   // blocks staged but not yet handed to TMA, their first columns, this warp's staging
    __device__ __forceinline__ float bias_scale() const { return 1.f; }
    __device__ __forceinline__ bool stage_bias() const { return true; }   // chunk() always reads the staged values (zeros without a bias)
    __device__ __forceinline__ void init(State& s) const { s.pending = 0; s.scr = nullptr; }
    __device__ __forceinline__ void finish(State&, int, bool, int) const {}
    __device__ __forceinline__ void tile_begin(State&, int, bool, int, int) const {}
    // hand the staged blocks (one or two) to TMA: one proxy fence and one bulk group for both
    __device__ __forceinline__ void flush(State& s, int row0, float* wscr) const {
        if (s.pending == 0) return;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if ((threadIdx.x & 31) == 0) {
            for (int k = 0; k < s.pending; ++k)
                asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                             ::"l"(reinterpret_cast<uint64_t>(&tmC)), "r"(smem_u32(reinterpret_cast<uint8_t*>(wscr) + k * 4096)), "r"(s.n[k]), "r"(row0) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        s.pending = 0;
    }
    __device__ __forceinline__ void tile_end(State& s, int row, bool, int) const {
        if (use_tma && s.pending) flush(s, row - (int)(threadIdx.x & 31), s.scr);
    }
    // every staged block has been read out of shared memory (call before the CTA may exit)
    __device__ __forceinline__ void drain() const {
        if (use_tma && (threadIdx.x & 31) == 0) tma_store_wait_read<0>();
    }
    __device__ __forceinline__ void chunk(State& s, int row, bool row_ok, int n, int N, const float (&v)[32],
                                          const float* sb, float* wscr) const {
        const int lane = threadIdx.x & 31;
        if (use_tma && wscr != nullptr) {             // warp-uniform
            if (s.pending == 0) {                     // both buffers were handed over: wait until TMA has read them
                if (lane == 0) tma_store_wait_read<0>();
                __syncwarp();
            }
            uint8_t* buf = reinterpret_cast<uint8_t*>(wscr) + s.pending * 4096;
#pragma unroll
            for (int j = 0; j < 8; ++j)
                *reinterpret_cast<float4*>(buf + lane * 128 + ((j ^ (lane & 7)) << 4)) =
                    make_float4(v[4 * j] + sb[4 * j], v[4 * j + 1] + sb[4 * j + 1], v[4 * j + 2] + sb[4 * j + 2], v[4 * j + 3] + sb[4 * j + 3]);
            s.n[s.pending] = n;
            s.scr = wscr;
            if (++s.pending == 2) flush(s, row - lane, wscr);
        } else if (row_ok) {
            float* p = C + (int64_t)row * ldc + n;
#pragma unroll
            for (int i = 0; i < 32; ++i)
                if (n + i < N) p[i] = v[i] + sb[i];
        }
    }
};

// One MUFU each (tanh.approx: max relative error ~2^-11, far inside the bf16-mode tolerance of 2e-2).
__device__ __forceinline__ float fast_tanh(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float fast_sigmoid(float x) { return fmaf(fast_tanh(0.5f * x), 0.5f, 0.5f); }

// 256-bit global accesses (sm_100): the 8 cell states of a chunk are 32 contiguous bytes per row, one request per lane instead of two.
// The cell epilogue's rows are Hpad floats apart, so every lane of a warp instruction touches its own line: these scattered requests
// are what slows the TMA main loop running next to the epilogue (profiles/micro/kb_probe.cu, mode 11: 512 -> 917 clocks per K block).
__device__ __forceinline__ void ldg_f32x8(const float* p, float4& a, float4& b) {
    asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p));
}
__device__ __forceinline__ void stg_f32x8(float* p, const float (&v)[8]) {
    asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 ::"l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]) : "memory");
}
__device__ __forceinline__ void stg_b32x8(void* p, const uint4& a, const uint4& b) {
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 ::"l"(p), "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b.x), "r"(b.y), "r"(b.z), "r"(b.w) : "memory");
}
struct TcEpiCell {
    const float* addend;   // [M][ld_add] fp32 x*W_ih^T + b for this timestep (column order 4j+g) or nullptr
    int64_t ld_add;
    const float* bias;     // [4H] used when addend == nullptr
    float* c;              // [M][ld_state] fp32 in/out; or, with c_rows != 0, blocked [ld_state / 8][c_rows][8] (cell_at)
    int c_rows;            // 0: row-major.  Blocked: the 8 states of a chunk of 32 consecutive rows are 1 KB contiguous, so a warp's
                           // access is 8 full lines instead of 32 lines with one sector each (persistent generation kernel)
    float* h;              // [M][ld_state] fp32 out (RNNCore.hidden[l][0]); nullptr: derived later from hb
    int ld_state;
    __nv_bfloat16* hb;     // [M][ld_hb] bf16 out: next step's / next layer's GEMM operand
    int ld_hb;
    float* raw_out;        // optional caller tensor [M][H]
    int64_t ld_raw;
    const int32_t* start;  // ragged prompts: stream `row` is fed from timestep start[row] on (nullptr: all from 0)
    int t;                 // timestep of this launch
    // The old cell state of every unit this thread will update in the tile is fetched in tile_begin,
    // i.e. while the tile's MMAs are still running, so the epilogue never waits on global memory.
    struct State {
        float4 c[8];       // up to 4 chunks x 8 units
        int n_first;
        uint4 h_even;      // bf16 outputs of the last even chunk, waiting for the odd one: the two go out as ONE 32-byte store
        int h_pending;     // unit offset of that chunk, -1: none
    };
    __device__ __forceinline__ float bias_scale() const { return 1.f; }
    __device__ __forceinline__ bool stage_bias() const { return true; }   // chunk() always reads the staged values (zeros without a bias)
    static constexpr int kScratchBytes = 0;
    __device__ __forceinline__ float* cell_at(int row, int unit0) const {       // unit0: multiple of 8
        return c_rows ? c + ((int64_t)(unit0 >> 3) * c_rows + row) * 8 : c + (int64_t)row * ld_state + unit0;
    }
    __device__ __forceinline__ void drain() const {}
    __device__ __forceinline__ void init(State&) const {}
    __device__ __forceinline__ void finish(State&, int, bool, int) const {}
    __device__ __forceinline__ void tile_begin(State& s, int row, bool row_ok, int n_first, int nchunks) const {
        s.n_first = n_first;
        s.h_pending = -1;
        if (!row_ok) return;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (k < nchunks && (n_first >> 2) + 8 * k + 8 <= ld_state) {
                ldg_f32x8(cell_at(row, (n_first >> 2) + 8 * k), s.c[2 * k], s.c[2 * k + 1]);
            }
        }
    }
    __device__ __forceinline__ void tile_end(State& s, int row, bool row_ok, int) const {
        if (row_ok && s.h_pending >= 0) *reinterpret_cast<uint4*>(hb + (int64_t)row * ld_hb + s.h_pending) = s.h_even;   // odd number of chunks
        s.h_pending = -1;
    }
    __device__ __forceinline__ void chunk(State& s, int row, bool row_ok, int n, int N, const float (&v)[32],
                                          const float* sb, float*) const {
        if (!row_ok) return;
        const int j0 = n >> 2;
        if (start != nullptr && t < __ldg(start + row)) {      // prompt not begun: state stays at reset, outputs are zero
            *reinterpret_cast<uint4*>(hb + (int64_t)row * ld_hb + j0) = make_uint4(0, 0, 0, 0);
            if (raw_out) {
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    if (n + 4 * u < N) raw_out[(int64_t)row * ld_raw + j0 + u] = 0.f;
            }
            return;
        }
        // all loads first (8 units = 32 contiguous bytes of c, 8 float4 of addend/bias), then the
        // arithmetic, then vector stores: no load waits behind a possibly-aliasing store
        const int k = (n - s.n_first) >> 5;                 // which prefetched chunk (warp-uniform)
        float4 c0, c1;
        if (k == 0) { c0 = s.c[0]; c1 = s.c[1]; }
        else if (k == 1) { c0 = s.c[2]; c1 = s.c[3]; }
        else if (k == 2) { c0 = s.c[4]; c1 = s.c[5]; }
        else { c0 = s.c[6]; c1 = s.c[7]; }
        float cv[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
        float4 a[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int nn = n + 4 * u;
            if (nn >= N) a[u] = make_float4(0.f, 0.f, 0.f, 0.f);   // zero K-padding units: stay 0
            else if (addend) a[u] = *reinterpret_cast<const float4*>(addend + (int64_t)row * ld_add + nn);
            else a[u] = *reinterpret_cast<const float4*>(sb + 4 * u);
        }
        float hv[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            float ig = fast_sigmoid(v[4 * u] + a[u].x), fg = fast_sigmoid(v[4 * u + 1] + a[u].y);
            float gg = fast_tanh(v[4 * u + 2] + a[u].z), og = fast_sigmoid(v[4 * u + 3] + a[u].w);
            cv[u] = fg * cv[u] + ig * gg;
            hv[u] = og * fast_tanh(cv[u]);
        }
        stg_f32x8(cell_at(row, j0), cv);
        if (h != nullptr) stg_f32x8(h + (int64_t)row * ld_state + j0, hv);
        if (raw_out) {
#pragma unroll
            for (int u = 0; u < 8; ++u)
                if (n + 4 * u < N) raw_out[(int64_t)row * ld_raw + j0 + u] = hv[u];
        }
        // 8 bf16 = 16 bytes; columns past H inside the last chunk are the zero K-padding
        __nv_bfloat162 p0 = __floats2bfloat162_rn(hv[0], hv[1]), p1 = __floats2bfloat162_rn(hv[2], hv[3]);
        __nv_bfloat162 p2 = __floats2bfloat162_rn(hv[4], hv[5]), p3 = __floats2bfloat162_rn(hv[6], hv[7]);
        uint4 pk = make_uint4(*reinterpret_cast<uint32_t*>(&p0), *reinterpret_cast<uint32_t*>(&p1),
                              *reinterpret_cast<uint32_t*>(&p2), *reinterpret_cast<uint32_t*>(&p3));
        if ((k & 1) == 0 && (j0 & 15) == 0) { s.h_even = pk; s.h_pending = j0; }     // (k and j0 & 15 are warp-uniform)
        else if (s.h_pending == j0 - 8) { stg_b32x8(hb + (int64_t)row * ld_hb + j0 - 8, s.h_even, pk); s.h_pending = -1; }
        else *reinterpret_cast<uint4*>(hb + (int64_t)row * ld_hb + j0) = pk;
    }
};

}  // namespace tc
}  // namespace awd
