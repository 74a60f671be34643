// Host side of the tcgen05 GEMM family (gemm_tc.cuh), the fused sampling epilogue and the small bf16
// helper kernels (weight packing, embedding gather, sampler finalisation).
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <mutex>

#include "gemm_tc.cuh"
#include "tc_api.h"

namespace awd {
namespace tc {

// ------------------------------------------------------------------------------- tensor maps
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;
static std::once_flag g_encode_once;

static EncodeTiledFn encode_fn() {
    std::call_once(g_encode_once, [] {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            g_encode = reinterpret_cast<EncodeTiledFn>(fn);
    });
    return g_encode;
}

// bf16 row-major [rows][cols] with row pitch ld elements; box = box_rows x 64 columns, 128B swizzle.
int make_tmap_bf16(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return set_error(AWD_ECUDA, "cuTensorMapEncodeTiled is not available from the driver");
    if ((reinterpret_cast<uintptr_t>(base) & 15) || (ld * 2) % 16)
        return set_error(AWD_EINVAL, "TMA operand must be 16-byte aligned with a 16-byte row pitch");
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {ld * 2};
    cuuint32_t box[2] = {(cuuint32_t)BK, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(AWD_ECUDA, "cuTensorMapEncodeTiled failed with %d", (int)r);
    return AWD_OK;
}

// fp32 row-major [rows][cols] output, row pitch ld elements; box = 32 rows x 32 columns (128 bytes), 128B swizzle.
int make_tmap_f32_store(CUtensorMap* map, void* base, uint64_t rows, uint64_t cols, uint64_t ld) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return set_error(AWD_ECUDA, "cuTensorMapEncodeTiled is not available from the driver");
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {ld * 4};
    cuuint32_t box[2] = {32, 32};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(AWD_ECUDA, "cuTensorMapEncodeTiled (fp32 store) failed with %d", (int)r);
    return AWD_OK;
}

// Launch-time caches are kept PER DEVICE: one process may drive several GPUs (BatchScheduler replicas),
// and function attributes / occupancy answers belong to the device that was current when they were set.
constexpr int kMaxDev = 64;
static int cur_dev() {
    int dev = 0;
    cudaGetDevice(&dev);
    return (dev >= 0 && dev < kMaxDev) ? dev : 0;
}
static int num_sms() {
    static std::atomic<int> sms[kMaxDev];
    const int dev = cur_dev();
    int n = sms[dev].load(std::memory_order_relaxed);
    if (!n) {
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
        sms[dev].store(n, std::memory_order_relaxed);
    }
    return n;
}

// ------------------------------------------------------------------- fused sampling epilogue
// The reference draws multinomial(exp(z/T), n_draws) without replacement and keeps the first allowed
// id (main.py:61-68): an exponential race whose keys are z/T + Gumbel noise.  The race is run
// hierarchically, which is exact (max-stability of the Gumbel law): over a chunk of 32 columns
//     max_i (z_i + G_i)  ~  logsumexp(z_chunk) + G_chunk,      argmax | chunk  ~  softmax(z_chunk),
// so the GEMM epilogue needs one exp per logit and ONE Philox call per 32 logits; the id inside a chunk
// is drawn (inverse CDF, second random word) only when that chunk takes the lead of its row.
// Disallowed ids (id 0, exclude_tokens) are singletons with their own key, parked in banned_keys
// for the "else the first candidate" fallback.  Greedy mode: keys are the logits, argmax is exact.
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint32_t k0, uint32_t k1) {
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
__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float u01(uint32_t bits) { return uniform_open01(bits); }
__device__ __forceinline__ float gumbel(uint32_t bits) { return gumbel_from_bits(bits); }
__device__ __forceinline__ uint32_t pick_word(const uint4& r, int q) { return q == 0 ? r.x : q == 1 ? r.y : q == 2 ? r.z : r.w; }
__device__ __forceinline__ bool better(float ka, int ia, float kb, int ib) { return (ka > kb) || (ka == kb && ia < ib); }

// float4 index of columns 4j..4j+3 of the parked chunk of (row, slot); M = rows of the launch
__device__ __forceinline__ int64_t zwin_at(int row, int slot, int j, int M) { return ((int64_t)slot * 8 + j) * M + row; }

// Philox counter domains (4th word): 1 = chunk noise, 2 = in-chunk uniform, 3 = banned-id noise.
constexpr float kLog2e = 1.4426950408889634f, kLn2 = 0.6931471805599453f;

// kListPtr: the list of disallowed ids is read through `banned_list` instead of sp.banned.  The persistent kernel
// builds this functor in registers; indexing sp.banned[t] there would force the whole struct into local memory
// (a dozen local loads per 32-column chunk).  Launched as a kernel parameter (kListPtr = false) the list sits in
// constant memory and is indexed in place.
template <bool kListPtr>
struct TcEpiSampleT {
    SampleParams sp;
    const int* banned_list;  // kListPtr only: sp.banned of a __grid_constant__ parameter
    int ban_lo, ban_hi;      // smallest / largest disallowed id (INT_MAX / -1 without any): the per-chunk range test stays in registers
    __device__ __forceinline__ int ban(int t) const {
        if constexpr (kListPtr) return banned_list[t];
        else return sp.banned[t];
    }
    const float* bias;       // decoder bias [V] or nullptr
    float2* part;            // [M][slots]: (best allowed key, chunk index | greedy: id), one slot per (CTA group, column part)
    float* zwin;             // [slots][8][M] float4: scaled logits of each (row, slot)'s leading chunk (sampling mode); rows are the
                             // fastest index so that the 32 lanes of a warp (32 rows) park with full-line stores (zwin_at)
    float* banned_keys;      // [M][64], followed by the M row bounds, one per 128-byte line (see row_bound())
    int M;                   // streams in this launch
    int slots;
    int tiles_m;
    int n_off;               // vocabulary shard: global id of this GEMM's column 0 (multiple of 32), else 0
    struct State {
        float ko;
        int io;
        float g[4];          // Gumbel noise of the (up to) 4 chunks a thread visits per tile: one Philox call, drawn in
                             // tile_begin, i.e. while the tile's MMAs are still running
        int first;           // id of the first of those chunks
        float bound;         // best lower bound of the row's winning key any CTA has published (read once per tile; in the
                             // CTA's first tile of a phase once per chunk, one chunk ahead)
        float pub;           // what this thread last published
        int tiles;           // tiles begun in this phase
        int next_raw;        // first tile: the row bound as it was one chunk ago (load issued early, consumed here)
    };
    // Row bounds: one int per stream behind banned_keys, holding the best allowed race key seen so far by ANY
    // CTA working on the row (order-preserving int encoding, red.max).  A chunk whose key cannot exceed it
    // (key <= max logit + ln 32 + Gumbel) cannot be the row's winner, so its 32 exponentials, its log and its
    // parking are skipped when no lane of the warp needs them.  Which chunks get skipped depends on timing; the
    // winner does not: the winning chunk's bound is >= its key > every other key, so it is always evaluated.
    // (Packed, kRowBoundStride = 1: a warp's 32 rows share one 128-byte line.  One 32-byte sector or one whole line per row were
    //  measured too: no difference at 128 or 1024 streams.)
    __device__ __forceinline__ int* row_bound(int row) const { return reinterpret_cast<int*>(banned_keys + (int64_t)M * 64) + (int64_t)row * kRowBoundStride; }
    int* row_bound_host() const { return reinterpret_cast<int*>(banned_keys + (int64_t)M * 64); }
    static __device__ __forceinline__ int enc_key(float k) { const int b = __float_as_int(k); return b >= 0 ? b : b ^ 0x7fffffff; }
    static __device__ __forceinline__ float dec_key(int e) { return __int_as_float(e >= 0 ? e : e ^ 0x7fffffff); }
    // slot of the calling epilogue thread in a row-stationary launch (same formula as the kernel's finish())
    __device__ __forceinline__ int cur_slot() const { return (blockIdx.x / tiles_m) * EPI_PARTS + (((threadIdx.x >> 5) - 2) >> 2); }
    __device__ __forceinline__ float bias_scale() const { return sp.inv_temperature; }   // staged bias is pre-divided by T
    __device__ __forceinline__ bool stage_bias() const { return bias != nullptr; }       // nullptr: folded into the GEMM
    static constexpr int kScratchBytes = 0;
    __device__ __forceinline__ void drain() const {}
    __device__ __forceinline__ void init(State& s) const {
        s.ko = -INFINITY;
        s.io = 0x7fffffff;
        s.bound = s.pub = -INFINITY;
        s.next_raw = (int)0x80808080;
        s.tiles = 0;
    }
    __device__ __forceinline__ void tile_begin(State& s, int row, bool row_ok, int n_first, int nchunks) const {
        s.first = (n_first + n_off) >> 5;
        if (!sp.greedy) {
            // The noise of chunk X is word X & 3 of the Philox block numbered X & ~3, whatever tile or column part X falls into:
            // a thread whose (up to four) chunks start at a multiple of four needs one block, otherwise two (uniform over the warp).
            const int lead = s.first & 3;
            const uint4 rnd = philox4x32_10(make_uint4((uint32_t)(s.first - lead), (uint32_t)row, sp.step_lo, 1u), sp.seed_lo, sp.seed_hi);
            uint32_t b0 = rnd.x, b1 = rnd.y, b2 = rnd.z, b3 = rnd.w;
            if (lead != 0) {                               // (explicit cases: an indexed word array went to local memory)
                uint4 nxt = make_uint4(0u, 0u, 0u, 0u);
                if (lead + nchunks > 4)
                    nxt = philox4x32_10(make_uint4((uint32_t)(s.first - lead + 4), (uint32_t)row, sp.step_lo, 1u), sp.seed_lo, sp.seed_hi);
                if (lead == 1) { b0 = rnd.y; b1 = rnd.z; b2 = rnd.w; b3 = nxt.x; }
                else if (lead == 2) { b0 = rnd.z; b1 = rnd.w; b2 = nxt.x; b3 = nxt.y; }
                else { b0 = rnd.w; b1 = nxt.x; b2 = nxt.y; b3 = nxt.z; }
            }
            s.g[0] = gumbel(b0); s.g[1] = gumbel(b1); s.g[2] = gumbel(b2); s.g[3] = gumbel(b3);
            // The row bound any CTA has published is loaded at the end of the previous tile (tile_end) and consumed by the first chunk.
            // Decoding a load issued here made every epilogue warp wait for an L2 round trip on a contended line once per tile, on the
            // critical path of a phase whose epilogue is slower than its MMAs.
            // (First tile of a phase: no load here.  Without a grid barrier in front of the decoder phase the epilogue warps get here
            //  while other CTAs may still be resetting the bounds of the previous token (finalize_row); everything read from the first
            //  chunk on is ordered behind that reset through the accumulator -- the MMAs needed the last layer's h, which needed x.)
            if (s.tiles == 0) s.next_raw = (int)0x80808080;
        }
        ++s.tiles;
    }
    __device__ __forceinline__ void tile_end(State& s, int row, bool row_ok, int) const {
        if (!sp.greedy && row_ok) s.next_raw = __ldcg(row_bound(row));      // for the next tile's first chunk (issued before the red.max below)
        if (!sp.greedy && row_ok && s.ko > s.pub && s.ko > s.bound) {
            atomicMax(row_bound(row), enc_key(s.ko));       // result unused: compiles to RED.MAX
            s.pub = s.ko;
        }
    }
    __device__ __forceinline__ void finish(State& s, int row, bool row_ok, int slot) const {
        if (row_ok) part[(int64_t)row * slots + slot] = make_float2(s.ko, __int_as_float(s.io));
    }
    // Two forms of the same arithmetic.  bias != nullptr: z = acc/T + b/T with b/T staged in shared memory.
    // bias == nullptr (the engine folded the decoder bias into two spare K columns of the weight matrix, see
    // engine_bf16.cu ensure_fold): the accumulator already holds h.w + b, the 1/T factor rides in the FFMA that
    // feeds ex2 and in the few values that leave the thread -- no shared-memory reads, no per-element scaling.
    __device__ __forceinline__ void chunk(State& s, int row, bool row_ok, int n, int N, const float (&v)[32],
                                          const float* sb, float*) const {
        float z[32];
        float sc = 1.f;                               // logit = z * sc
        if (bias != nullptr) {
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
                const float4 b = *reinterpret_cast<const float4*>(sb + i);
                z[i] = fmaf(v[i], sp.inv_temperature, b.x);
                z[i + 1] = fmaf(v[i + 1], sp.inv_temperature, b.y);
                z[i + 2] = fmaf(v[i + 2], sp.inv_temperature, b.z);
                z[i + 3] = fmaf(v[i + 3], sp.inv_temperature, b.w);
            }
        } else {
            sc = sp.inv_temperature;
#pragma unroll
            for (int i = 0; i < 32; ++i) z[i] = v[i];
        }
        if (n + 32 > N) {
#pragma unroll
            for (int i = 0; i < 32; ++i)
                if (n + i >= N) z[i] = -INFINITY;
        }
        const int ng = n + n_off;                     // global id of this chunk's first column
        const int chunk_id = ng >> 5;
        if (ban_lo < ng + 32 && ban_hi >= ng) {
            uint32_t mask = 0;
            for (int t = 0; t < sp.n_banned; ++t) {
                const int id = ban(t);
                if (id >= ng && id < ng + 32 && id - n_off < N) mask |= 1u << (id - ng);
            }
            if (mask) banned_columns(row, row_ok, ng, mask, z, sc);
        }
        float m4[4] = {z[0], z[1], z[2], z[3]};
#pragma unroll
        for (int i = 4; i < 32; ++i) m4[i & 3] = fmaxf(m4[i & 3], z[i]);
        const float m = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
        if (sp.greedy) {
            // a thread visits its chunks in ascending id order, so a strict '>' keeps the lowest id on ties
            if (row_ok && m != -INFINITY && m * sc > s.ko) {
                int idx = 0;
#pragma unroll
                for (int i = 31; i >= 0; --i)
                    if (z[i] == m) idx = i;
                s.ko = m * sc;
                s.io = ng + idx;
            }
            return;
        }
        // key = logsumexp(chunk) + Gumbel <= max + ln 32 + Gumbel: evaluate the sum only if that upper end beats the best lower
        // bound known for the row's winning key (this thread's or the published one), for at least one lane of the warp.
        const int gq = chunk_id - s.first;
        const float g = gq == 0 ? s.g[0] : gq == 1 ? s.g[1] : gq == 2 ? s.g[2] : s.g[3];
        constexpr float kLn32Hi = 3.4667f, kLn32Lo = 3.4647f;   // ln 32 = 3.46574 with margins far above the ex2/lg2/rounding error
        const float ub = fmaf(m, sc, kLn32Hi) + g;
        s.bound = fmaxf(s.bound, dec_key(s.next_raw));         // (next_raw: see tile_begin; unchanged between loads)
        if (s.tiles == 1) {
            // First tile of the phase: no exact key exists yet anywhere, so a bound made of exact keys alone lets every thread
            // evaluate its first chunks in full -- at 128 streams, where a CTA owns one or two tiles per token, that was the whole
            // epilogue.  Here the chunk's bracket, which costs no exponential,
            //     max(m, mean + ln 32)  <=  logsumexp(chunk)  <=  m + ln 32          (m = max logit; Jensen for the mean)
            // gives a lower end that goes to the row bound at once (red.max): it is a valid lower bound of the ROW's winning key,
            // so every chunk of the row whose upper end stays below it is out of the race.  The winner is always evaluated: its
            // upper end >= its key >= any lower bound of the winning key.  The bound is re-read every chunk, one chunk ahead.
            float a4[4] = {z[0], z[1], z[2], z[3]};
#pragma unroll
            for (int i = 4; i < 32; ++i) a4[i & 3] += z[i];
            const float zsum = (a4[0] + a4[1]) + (a4[2] + a4[3]);   // -inf as soon as one column is masked: then only m counts
            float lb = fmaf(m, sc, g);
            if (zsum != -INFINITY) lb = fmaxf(lb, fmaf(zsum * (1.0f / 32.0f), sc, kLn32Lo) + g);
            // Publish only promising bounds (noise in the top ~8 %: the winner's is far higher): any subset of valid bounds is
            // valid, and the L2 applies the red.max of a line one lane at a time.
            if (row_ok && g > 2.5f && lb > fmaxf(fmaxf(s.ko, s.bound), s.pub)) {
                atomicMax(row_bound(row), enc_key(lb));     // result unused: compiles to RED.MAX
                s.pub = lb;
            }
            if (row_ok) s.next_raw = __ldcg(row_bound(row));    // consumed by the next chunk
        }
        const float known = fmaxf(fmaxf(s.ko, s.bound), s.pub);
        const bool need = row_ok && ub > known;                // false for an all -inf chunk
        if (!__any_sync(0xffffffffu, need)) return;
        const float cl = sc * kLog2e, ml = (m == -INFINITY ? 0.f : m) * cl;
        float s4[4] = {0.f, 0.f, 0.f, 0.f};          // four independent chains
#pragma unroll
        for (int i = 0; i < 32; ++i) s4[i & 3] += ex2_approx(fmaf(z[i], cl, -ml));
        const float sum = (s4[0] + s4[1]) + (s4[2] + s4[3]);
        const float key = fmaf(m, sc, __logf(sum)) + g;
        if (need && key > s.ko) {
            // New leader of this (row, slot).  Park the chunk's scaled logits (one 128-byte line); the finalize
            // kernel draws the id inside the overall winner from them (~ softmax(z_chunk)).
            s.ko = key;
            s.io = chunk_id;
            // (row-major [row][slot][32] cost a 128-byte line per lane and instruction: ~3000 clocks for the first chunk of a phase,
            //  when every thread becomes a leader at once)
            float4* zw = reinterpret_cast<float4*>(zwin) + zwin_at(row, cur_slot(), 0, M);
#pragma unroll
            for (int i = 0; i < 8; ++i) zw[(int64_t)i * M] = make_float4(z[4 * i] * sc, z[4 * i + 1] * sc, z[4 * i + 2] * sc, z[4 * i + 3] * sc);
        }
    }
    // Disallowed ids inside this chunk (a handful of chunks per row): their race keys go to banned_keys[row][t]
    // and their columns leave the race.  One pass per disallowed id (a rolled loop: the Philox call appears once in the code, not
    // once per column -- the tile that holds the low ids made its CTA the straggler of every decoder phase otherwise); the column
    // is picked out of the register array with a select chain, so z[] stays in registers.
    __device__ __forceinline__ void banned_columns(int row, bool row_ok, int ng, uint32_t mask, float (&z)[32], float sc) const {
#pragma unroll 1
        for (int t = 0; t < sp.n_banned; ++t) {
            const int id = ban(t);
            const int col = id - ng;
            if (col < 0 || col >= 32 || !((mask >> col) & 1u)) continue;
            float zb = 0.f;
#pragma unroll
            for (int i = 0; i < 32; ++i) zb = (i == col) ? z[i] : zb;
            zb *= sc;
            if (!sp.greedy)
                zb += gumbel(philox4x32_10(make_uint4((uint32_t)id, (uint32_t)row, sp.step_lo, 3u), sp.seed_lo, sp.seed_hi).x);
            if (row_ok) banned_keys[(int64_t)row * 64 + t] = zb;
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) z[i] = ((mask >> i) & 1u) ? -INFINITY : z[i];
    }
};
using TcEpiSample = TcEpiSampleT<false>;

// One warp per stream: merge the per-tile partials, resolve the reference's fallback rule
// (main.py:63-68), emit the id and gather its embedding row (bf16, 128-bit loads) as the next step's
// layer-0 input.
struct FinalizeArgs {
    const float2* part;
    const float* zwin;
    const float* banned_keys;   // [bs][64] keys of the disallowed ids, then bs row bounds, kRowBoundStride ints apart (TcEpiSample::row_bound)
    int tiles_n, bs, V;      // tiles_n = partial slots per row
    SampleParams sp;
    int64_t* out;
    // continuous batching (awd_generate_resume): a stream still inside its prompt is fed its own next prompt token
    // instead of the id it would have drawn: forced[step][row] while step < forced_len[row]
    const int64_t* forced;
    const int32_t* forced_len;
    int step;
    const uint4* table;       // bf16 embedding rows (gather source)
    int row_vec4;             // uint4 per row (Epad / 8)
    uint4* xnext;
};

// (sp, step: the sampling parameters and the token index.  The persistent kernel passes its own grid-constant copy -- filling
//  a.sp there copied the 64-entry list of disallowed ids to local memory once per token and thread)
__device__ __forceinline__ void finalize_row(const FinalizeArgs& a, const SampleParams& sp, uint32_t step_lo, int row, int lane) {
    float ko = -INFINITY;
    int io = 0x7fffffff, slot = 0;
    const float2* prow = a.part + (int64_t)row * a.tiles_n;
    // This step is a chain of dependent L2 / HBM round trips on the critical path of every token (partials -> parked logits ->
    // embedding row), so everything that does not depend on the chain is requested first: the keys of the disallowed ids, and
    // (up to 320 partial slots = one M tile of streams, the latency-critical case) ALL partials of the row in one round.
    float bk[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int u = 0; u < 2; ++u) {
        const int t = lane + 32 * u;
        if (t < sp.n_banned && sp.banned[t] < a.V) bk[u] = __ldcg(a.banned_keys + (int64_t)row * 64 + t);
    }
    if (a.tiles_n <= 320) {
        float2 p[10];
#pragma unroll
        for (int u = 0; u < 10; ++u) {
            const int t = u * 32 + lane;
            p[u] = t < a.tiles_n ? __ldcg(prow + t) : make_float2(-INFINITY, __int_as_float(0x7fffffff));
        }
#pragma unroll
        for (int u = 0; u < 10; ++u) {
            const int pio = __float_as_int(p[u].y);
            if (better(p[u].x, pio, ko, io)) { ko = p[u].x; io = pio; slot = u * 32 + lane; }
        }
    } else {
        for (int t0 = 0; t0 < a.tiles_n; t0 += 128) {          // 4 independent loads in flight per lane
            float2 p[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int t = t0 + u * 32 + lane;
                p[u] = t < a.tiles_n ? __ldcg(prow + t) : make_float2(-INFINITY, __int_as_float(0x7fffffff));
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int pio = __float_as_int(p[u].y);
                if (better(p[u].x, pio, ko, io)) { ko = p[u].x; io = pio; slot = t0 + u * 32 + lane; }
            }
        }
    }
    // disallowed ids: how many outrank the best allowed key, and which one leads
    int outrank = 0;
    float kb = -INFINITY;
    int ib = 0x7fffffff;
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        float k = __shfl_xor_sync(0xffffffffu, ko, s);
        int i = __shfl_xor_sync(0xffffffffu, io, s);
        int sl = __shfl_xor_sync(0xffffffffu, slot, s);
        if (better(k, i, ko, io)) { ko = k; io = i; slot = sl; }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
        const int t = lane + 32 * u;
        if (t < sp.n_banned && sp.banned[t] < a.V) {
            // in sampling mode `io` is a chunk index, not an id: an exact key tie has probability zero
            outrank += (sp.greedy ? better(bk[u], sp.banned[t], ko, io) : (bk[u] > ko)) ? 1 : 0;
            if (better(bk[u], sp.banned[t], kb, ib)) { kb = bk[u]; ib = sp.banned[t]; }
        }
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        outrank += __shfl_xor_sync(0xffffffffu, outrank, s);
        float k = __shfl_xor_sync(0xffffffffu, kb, s);
        int i = __shfl_xor_sync(0xffffffffu, ib, s);
        if (better(k, i, kb, ib)) { kb = k; ib = i; }
    }
    int pick = io;
    if (io == 0x7fffffff || outrank >= sp.n_draws) {
        pick = (ib != 0x7fffffff) ? ib : 0;                  // first candidate overall is a disallowed id
    } else if (!sp.greedy) {
        // `io` is the winning chunk: draw the id inside it from its parked logits (lane = column)
        const float z = __ldcg(a.zwin + zwin_at(row, slot, lane >> 2, a.bs) * 4 + (lane & 3));
        float m = z;
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, s));
        const float ex = (z == -INFINITY) ? 0.f : ex2_approx((z - m) * kLog2e);
        float pre = ex;                                      // inclusive prefix sum over lanes
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
            float o = __shfl_up_sync(0xffffffffu, pre, s);
            if (lane >= s) pre += o;
        }
        const float total = __shfl_sync(0xffffffffu, pre, 31);
        const uint4 r = philox4x32_10(make_uint4((uint32_t)row, (uint32_t)io, step_lo, 2u), sp.seed_lo, sp.seed_hi);
        const float target = u01(r.x) * total;
        const unsigned hit = __ballot_sync(0xffffffffu, ex > 0.f && pre >= target);
        const unsigned valid = __ballot_sync(0xffffffffu, ex > 0.f);
        const int l = hit ? (__ffs(hit) - 1) : (31 - __clz(valid));
        pick = io * 32 + l;
    }
    if (a.forced != nullptr && a.step < __ldg(a.forced_len + row)) pick = (int)__ldg(a.forced + (int64_t)a.step * a.bs + row);
    if (lane == 0) {
        a.out[row] = pick;
        // re-arm the row bound of the sampling epilogue for the next token ("no key yet")
        reinterpret_cast<int*>(const_cast<float*>(a.banned_keys) + (int64_t)a.bs * 64)[(int64_t)row * kRowBoundStride] = (int)0x80808080;
    }
    if (a.xnext != nullptr) {
        if (a.row_vec4 <= 128) {                                   // all loads of the row in flight together
            uint4 r[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (lane + 32 * u < a.row_vec4) r[u] = __ldg(a.table + (int64_t)pick * a.row_vec4 + lane + 32 * u);
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (lane + 32 * u < a.row_vec4) a.xnext[(int64_t)row * a.row_vec4 + lane + 32 * u] = r[u];
        } else {
            for (int i = lane; i < a.row_vec4; i += 32)
                a.xnext[(int64_t)row * a.row_vec4 + i] = __ldg(a.table + (int64_t)pick * a.row_vec4 + i);
        }
    }
}

// ---- vocabulary-sharded decoder: per-shard partial and cross-shard merge ------------
// Each shard (a GPU, or a logical slice on one GPU) runs the fused decoder + race over its rows and
// reduces to ONE candidate per stream: (best allowed key, id) plus the keys of the disallowed ids it owns.
// Race keys of different shards are directly comparable, so the global result is a max over shards --
// the exchange is 8 + 4*n_banned bytes per stream and rank (one small all-gather), no normaliser needed.
struct ShardPartialArgs {
    const float2* part;
    const float* zwin;
    const float* banned_keys;
    int slots, bs, v0, v1;
    SampleParams sp;
    float* out_key;            // [bs]
    int64_t* out_id;           // [bs]
    float* out_banned;         // [bs][64]
};

__global__ void __launch_bounds__(128) shard_partial_kernel(ShardPartialArgs a) {
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (row >= a.bs) return;
    const SampleParams& sp = a.sp;
    float ko = -INFINITY;
    int io = 0x7fffffff, slot = 0;
    for (int t = lane; t < a.slots; t += 32) {
        const float2 p = __ldcg(a.part + (int64_t)row * a.slots + t);
        const int pio = __float_as_int(p.y);
        if (better(p.x, pio, ko, io)) { ko = p.x; io = pio; slot = t; }
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        float k = __shfl_xor_sync(0xffffffffu, ko, s);
        int i = __shfl_xor_sync(0xffffffffu, io, s);
        int sl = __shfl_xor_sync(0xffffffffu, slot, s);
        if (better(k, i, ko, io)) { ko = k; io = i; slot = sl; }
    }
    int id = io;
    if (io != 0x7fffffff && !sp.greedy) {                  // resolve the id inside the shard's winning chunk
        const float z = __ldcg(a.zwin + zwin_at(row, slot, lane >> 2, a.bs) * 4 + (lane & 3));
        float m = z;
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, s));
        const float ex = (z == -INFINITY) ? 0.f : ex2_approx((z - m) * kLog2e);
        float pre = ex;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
            float o = __shfl_up_sync(0xffffffffu, pre, s);
            if (lane >= s) pre += o;
        }
        const float total = __shfl_sync(0xffffffffu, pre, 31);
        // the in-chunk uniform is keyed by (row, chunk) so that every sharding draws the same id
        const uint4 r = philox4x32_10(make_uint4((uint32_t)row, (uint32_t)io, sp.step_lo, 2u), sp.seed_lo, sp.seed_hi);
        const float target = u01(r.x) * total;
        const unsigned hit = __ballot_sync(0xffffffffu, ex > 0.f && pre >= target);
        const unsigned valid = __ballot_sync(0xffffffffu, ex > 0.f);
        id = io * 32 + (hit ? (__ffs(hit) - 1) : (31 - __clz(valid)));
    }
    if (lane == 0) {
        a.out_key[row] = ko;
        a.out_id[row] = (io == 0x7fffffff) ? -1 : id;
    }
    for (int t = lane; t < 64; t += 32) {
        const bool mine = t < sp.n_banned && sp.banned[t] >= a.v0 && sp.banned[t] < a.v1;
        a.out_banned[(int64_t)row * 64 + t] = mine ? __ldcg(a.banned_keys + (int64_t)row * 64 + t) : -INFINITY;
    }
}

// keys/ids/banned are the all-gathered per-shard partials: [R][bs], [R][bs], [R][bs][64]
// (shard r's arrays start r * stride_k floats / r * stride_i int64 / r * stride_b floats after shard 0's: bs, bs, bs*64 for three
//  separate [R][...] arrays, or the size of one rank's packed block when the three sit in one all-gathered buffer)
__global__ void __launch_bounds__(128) shard_merge_kernel(const float* __restrict__ keys, const int64_t* __restrict__ ids,
                                                           const float* __restrict__ banned, int R, int bs, int64_t stride_k,
                                                           int64_t stride_i, int64_t stride_b, SampleParams sp, int64_t* __restrict__ out) {
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (row >= bs) return;
    float ko = -INFINITY;
    int io = 0x7fffffff;
    for (int r = lane; r < R; r += 32) {
        const float k = keys[(int64_t)r * stride_k + row];
        const int64_t i = ids[(int64_t)r * stride_i + row];
        if (i >= 0 && better(k, (int)i, ko, io)) { ko = k; io = (int)i; }
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        float k = __shfl_xor_sync(0xffffffffu, ko, s);
        int i = __shfl_xor_sync(0xffffffffu, io, s);
        if (better(k, i, ko, io)) { ko = k; io = i; }
    }
    int outrank = 0;
    float kb = -INFINITY;
    int ib = 0x7fffffff;
    for (int t = lane; t < sp.n_banned; t += 32) {
        float k = -INFINITY;
        for (int r = 0; r < R; ++r) k = fmaxf(k, banned[(int64_t)r * stride_b + (int64_t)row * 64 + t]);   // one shard owns it
        if (k == -INFINITY) continue;
        outrank += (sp.greedy ? better(k, sp.banned[t], ko, io) : (k > ko)) ? 1 : 0;
        if (better(k, sp.banned[t], kb, ib)) { kb = k; ib = sp.banned[t]; }
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        outrank += __shfl_xor_sync(0xffffffffu, outrank, s);
        float k = __shfl_xor_sync(0xffffffffu, kb, s);
        int i = __shfl_xor_sync(0xffffffffu, ib, s);
        if (better(k, i, kb, ib)) { kb = k; ib = i; }
    }
    if (lane == 0) out[row] = (io == 0x7fffffff || outrank >= sp.n_draws) ? ((ib != 0x7fffffff) ? ib : 0) : io;
}

__global__ void __launch_bounds__(128) sample_finalize_kernel(FinalizeArgs a) {
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (row < a.bs) finalize_row(a, a.sp, a.sp.step_lo, row, threadIdx.x & 31);
}

// ----------------------------------------------------------------------------- small kernels
// K1 in bf16 mode: out[n][:] = table[tok[n]][:], rows are Epad64 bf16 = 16-byte multiples.
__global__ void gather_bf16_kernel(const uint4* __restrict__ table, int row_vec4, int V, const int64_t* __restrict__ tok,
                                   int64_t n, uint4* __restrict__ out, int* err_flag) {
    int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t row = gid / row_vec4;
    int col = (int)(gid - row * row_vec4);
    if (row >= n) return;
    int64_t t = tok[row];
    if (t < 0 || t >= V) {
        if (col == 0) atomicExch(err_flag, 1);
        t = 0;
    }
    out[row * row_vec4 + col] = __ldg(table + t * row_vec4 + col);
}

// dst[r'][k0 + k] = bf16(src[r][k]) for k < cols; columns [k0+cols, k0+kpad) are zeroed.  H > 0 reorders
// rows from torch's gate-blocked (g*H + j) to gate-interleaved (4*j + g).
__global__ void pack_rows_bf16_kernel(const float* __restrict__ src, int rows, int cols, __nv_bfloat16* __restrict__ dst,
                                      int ld_dst, int k0, int kpad, int H) {
    int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t r = gid / kpad;
    int k = (int)(gid - r * kpad);
    if (r >= rows) return;
    int64_t rd = r;
    if (H > 0) {
        int g = (int)(r / H), j = (int)(r - (int64_t)g * H);
        rd = 4 * (int64_t)j + g;
    }
    dst[rd * ld_dst + k0 + k] = __float2bfloat16_rn(k < cols ? src[r * cols + k] : 0.f);
}

__global__ void f32_to_bf16_rows_kernel(const float* __restrict__ src, int64_t ld_src, int rows, int cols,
                                        __nv_bfloat16* __restrict__ dst, int ld_dst) {
    int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t r = gid / ld_dst;
    int k = (int)(gid - r * ld_dst);
    if (r >= rows) return;
    dst[r * ld_dst + k] = __float2bfloat16_rn(k < cols ? src[r * ld_src + k] : 0.f);
}

int launch_gather_bf16(const void* table, int ld, int V, const int64_t* tok, int64_t n, void* out, int* err_flag,
                       cudaStream_t st) {
    if (n == 0) return AWD_OK;
    int v4 = ld / 8;
    int64_t total = n * v4;
    gather_bf16_kernel<<<(unsigned)cdiv(total, 256), 256, 0, st>>>((const uint4*)table, v4, V, tok, n, (uint4*)out, err_flag);
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

int launch_pack_rows_bf16(const float* src, int rows, int cols, void* dst, int ld_dst, int k0, int kpad, int H,
                          cudaStream_t st) {
    int64_t total = (int64_t)rows * kpad;
    if (total == 0) return AWD_OK;
    pack_rows_bf16_kernel<<<(unsigned)cdiv(total, 256), 256, 0, st>>>(src, rows, cols, (__nv_bfloat16*)dst, ld_dst, k0, kpad, H);
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

int launch_f32_to_bf16_rows(const float* src, int64_t ld_src, int rows, int cols, void* dst, int ld_dst, cudaStream_t st) {
    int64_t total = (int64_t)rows * ld_dst;
    if (total == 0) return AWD_OK;
    f32_to_bf16_rows_kernel<<<(unsigned)cdiv(total, 256), 256, 0, st>>>(src, ld_src, rows, cols, (__nv_bfloat16*)dst, ld_dst);
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

// --------------------------------------------------------------------------------- launchers
// Every load of these kernels is a plain (non-multicast) TMA load: operands may be caller memory
// (awd_tc_gemm_bf16), and TMA multicast loads here read only this GPU's memory, since on B200 a TMA
// multicast load from another GPU's memory hung the GPU.  (Clusters with multicast operands were also
// measured 2-20 % slower on this workload in round 1: the step is bound by the bytes each SM ingests.)
template <int BN, class Epi>
static int launch_tc(const TcOperands& o, int M, int N, const Epi& epi, cudaStream_t st, bool row_stationary = false) {
    using Cfg = TileCfg<BN>;
    static std::atomic<bool> attr_set_dev[kMaxDev];
    std::atomic<bool>& attr_set = attr_set_dev[cur_dev()];
    auto kern = tc_gemm_kernel<BN, Epi>;
    if (!attr_set) {
        AWD_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_MAX));
        attr_set = true;
    }
    GemmShape g{};
    g.M = M; g.N = N;
    g.kb0 = o.k0 / BK; g.kb1 = o.k1 / BK;
    g.a0_row = o.a0_row; g.a1_row = o.a1_row; g.b_k0 = o.b_k0;
    g.tiles_m = (int)cdiv(M, BM); g.tiles_n = (int)cdiv(N, BN);
    constexpr int kTail = Cfg::TAIL_BYTES + Epi::kScratchBytes;         // what sits behind the operand ring
    g.stages = std::min(Cfg::STAGES, (Cfg::SMEM_MAX - kTail) / Cfg::STAGE_BYTES);
    int smem_bytes = g.stages * Cfg::STAGE_BYTES + kTail;
    // A-resident schedule: whole-K A tile + a 3-deep ring of weight tiles must fit in shared memory
    {
        const int num_kb = g.kb0 + g.kb1;
        const int need = num_kb * A_STAGE_BYTES + 3 * Cfg::B_STAGE_BYTES + kTail;
        if (need <= Cfg::SMEM_MAX && (g.tiles_n >= 4 || row_stationary) && g.tiles_m <= num_sms()) {
            g.a_res = 1;
            g.stages = std::min(8, (Cfg::SMEM_MAX - kTail - num_kb * A_STAGE_BYTES) / Cfg::B_STAGE_BYTES);
            smem_bytes = num_kb * A_STAGE_BYTES + g.stages * Cfg::B_STAGE_BYTES + kTail;
        }
    }
    if (row_stationary && !g.a_res) g.row_stat = 1;
    CUtensorMap tmA0, tmA1, tmB;
    int rc = make_tmap_bf16(&tmA0, o.a0, o.a0_rows, o.k0, o.lda0, BM);
    if (rc) return rc;
    if (o.k1 > 0) rc = make_tmap_bf16(&tmA1, o.a1, o.a1_rows, o.k1, o.lda1, BM);
    else tmA1 = tmA0;
    if (rc) return rc;
    rc = make_tmap_bf16(&tmB, o.w, o.w_rows, o.w_cols, o.ldw, BN);
    if (rc) return rc;
    const int num_tiles = g.tiles_m * g.tiles_n;
    cudaLaunchConfig_t cfg{};
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = smem_bytes;
    cfg.stream = st;
    if (g.a_res || g.row_stat) cfg.gridDim = dim3(std::min(g.tiles_n, std::max(1, num_sms() / g.tiles_m)) * g.tiles_m);
    else cfg.gridDim = dim3(std::min(num_tiles, num_sms()));
    static const bool dbg = getenv("AWD_TC_DEBUG") != nullptr;
    if (dbg) fprintf(stderr, "[awd] tc launch BN=%d: grid %u, stages %d, a_res %d, row_stat %d\n", BN, cfg.gridDim.x, g.stages, g.a_res, g.row_stat);
    AWD_CUDA(cudaLaunchKernelEx(&cfg, kern, tmA0, tmA1, tmB, g, epi));
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

// CTA-pair launch (tc_gemm_pair_kernel): BN = 256, pairs of adjacent M tiles.
template <class Epi>
static int launch_tc_pair(const TcOperands& o, int M, int N, const Epi& epi, cudaStream_t st, bool row_stationary) {
    static std::atomic<bool> attr_set_dev[kMaxDev];
    static std::atomic<int> max_clusters_dev[kMaxDev];
    std::atomic<bool>& attr_set = attr_set_dev[cur_dev()];
    std::atomic<int>& max_clusters = max_clusters_dev[cur_dev()];
    auto kern = tc_gemm_pair_kernel<Epi>;
    if (!attr_set) {
        AWD_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, PAIR_SMEM_MAX));
        attr_set = true;
    }
    GemmShape g{};
    g.M = M; g.N = N;
    g.kb0 = o.k0 / BK; g.kb1 = o.k1 / BK;
    g.a0_row = o.a0_row; g.a1_row = o.a1_row; g.b_k0 = o.b_k0;
    g.tiles_m = (int)cdiv(M, BM); g.tiles_n = (int)cdiv(N, PAIR_BN);
    const int num_kb = g.kb0 + g.kb1;
    int room = PAIR_SMEM_MAX - PAIR_TAIL_BYTES;
    if (num_kb * A_STAGE_BYTES + 3 * PAIR_B_BYTES <= room && (g.tiles_n >= 4 || row_stationary) && g.tiles_m <= num_sms()) {
        g.a_res = 1;
        g.stages = std::min(8, (room - num_kb * A_STAGE_BYTES) / PAIR_B_BYTES);
    } else {
        g.row_stat = row_stationary ? 1 : 0;
        g.stages = std::min(8, room / PAIR_STAGE_BYTES);
    }
    const int smem_bytes = (g.a_res ? num_kb * A_STAGE_BYTES + g.stages * PAIR_B_BYTES : g.stages * PAIR_STAGE_BYTES) + PAIR_TAIL_BYTES;
    CUtensorMap tmA0, tmA1, tmB;
    int rc = make_tmap_bf16(&tmA0, o.a0, o.a0_rows, o.k0, o.lda0, BM);
    if (rc) return rc;
    if (o.k1 > 0) rc = make_tmap_bf16(&tmA1, o.a1, o.a1_rows, o.k1, o.lda1, BM);
    else tmA1 = tmA0;
    if (rc) return rc;
    if ((rc = make_tmap_bf16(&tmB, o.w, o.w_rows, o.w_cols, o.ldw, PAIR_BN / 2))) return rc;
    cudaLaunchConfig_t cfg{};
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = smem_bytes;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (!max_clusters) {
        cfg.gridDim = dim3(num_sms() / 2 * 2);
        int n = 0;
        if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess || n <= 0) {
            cudaGetLastError();
            n = num_sms() / 2;
        }
        max_clusters = n;
    }
    if (g.a_res || g.row_stat) {
        const int groups = std::min(g.tiles_n, std::max(1, 2 * max_clusters.load() / g.tiles_m));
        cfg.gridDim = dim3(groups * g.tiles_m);
    } else {
        const int num_ctiles = (int)(cdiv(g.tiles_m, 2) * g.tiles_n);
        cfg.gridDim = dim3(std::min(num_ctiles, max_clusters.load()) * 2);
    }
    static const bool dbg = getenv("AWD_TC_DEBUG") != nullptr;
    if (dbg) fprintf(stderr, "[awd] pair launch: grid %u, max_clusters %d, stages %d, a_res %d, row_stat %d, smem %d\n", cfg.gridDim.x,
                     max_clusters.load(), g.stages, g.a_res, g.row_stat, smem_bytes);
    AWD_CUDA(cudaLaunchKernelEx(&cfg, kern, tmA0, tmA1, tmB, g, epi));
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

// CTA pairs (tc_gemm_pair_kernel).  AWD_TC_PAIR=1/0 forces them on/off; unset: on for the LSTM-step and the
// sampled-decoder launches (measured +6 % on the per-kernel path: 5.94 -> 6.29 M tokens/s at 1024 streams), off for the
// store GEMMs (the pair kernel has no staging for TMA stores).
static bool use_pair(int M, bool default_on) {
    static int env = -2;
    if (env == -2) { const char* s = getenv("AWD_TC_PAIR"); env = s ? atoi(s) : -1; }
    const bool on = env < 0 ? default_on : env != 0;
    return on && (cdiv(M, BM) % 2 == 0);
}

static int check_operands(const TcOperands& o) {
    if (o.k0 <= 0 || o.k0 % BK || o.k1 % BK || o.k1 < 0) return set_error(AWD_EINVAL, "tc gemm: K extents must be multiples of 64");
    return AWD_OK;
}

int launch_tc_store(const TcOperands& o, int M, int N, float* C, int64_t ldc, const float* bias, cudaStream_t st) {
    int rc = check_operands(o);
    if (rc) return rc;
    TcEpiStore epi{};
    epi.C = C; epi.ldc = ldc; epi.bias = bias; epi.M = M;
    // TMA tensor stores need a 16-byte aligned base and row pitch; anything else takes the direct-store path
    static const bool no_tma = getenv("AWD_TC_STORE_TMA") != nullptr && atoi(getenv("AWD_TC_STORE_TMA")) == 0;
    if (!no_tma && C != nullptr && (reinterpret_cast<uintptr_t>(C) & 15) == 0 && (ldc % 4) == 0) {
        if ((rc = make_tmap_f32_store(&epi.tmC, C, (uint64_t)M, (uint64_t)N, (uint64_t)ldc))) return rc;
        epi.use_tma = 1;
    }
    // wide tiles when there are enough of them to fill the SMs
    if (cdiv(M, BM) * cdiv(N, 256) >= num_sms()) {
        if (use_pair(M, false)) return launch_tc_pair<TcEpiStore>(o, M, N, epi, st, false);
        return launch_tc<256, TcEpiStore>(o, M, N, epi, st);
    }
    return launch_tc<128, TcEpiStore>(o, M, N, epi, st);
}

int launch_tc_cell(const TcOperands& o, int M, int N, const TcCellArgs& a, cudaStream_t st) {
    int rc = check_operands(o);
    if (rc) return rc;
    TcEpiCell epi{a.addend, a.ld_add, a.bias, a.c, 0, a.h, a.ld_state, (__nv_bfloat16*)a.hb, a.ld_hb, a.raw_out, a.ld_raw, a.start, a.t};
    if (cdiv(M, BM) * cdiv(N, 256) >= num_sms() - 8) {
        if (use_pair(M, true)) return launch_tc_pair<TcEpiCell>(o, M, N, epi, st, false);
        return launch_tc<256, TcEpiCell>(o, M, N, epi, st);
    }
    return launch_tc<128, TcEpiCell>(o, M, N, epi, st);
}

// Partial slots per stream for the fused sampler: one per (CTA group of the row-stationary schedule, column part).
int tc_sample_slots(int M, int N) {
    const int tiles_m = (int)cdiv(M, BM);
    return EPI_PARTS * std::min((int)cdiv(N, 256), std::max(1, num_sms() / tiles_m));
}

// ... and for the persistent generation kernel, whose left-over CTAs may use one more pair (launch_mega_generate)
int tc_mega_slots(int M, int N) { return tc_sample_slots(M, N) + EPI_PARTS; }

int launch_tc_sample(const TcOperands& o, int M, int N, const SampleParams& sp, const float* bias, float* part,
                     float* zwin, float* banned_keys, cudaStream_t st) {
    int rc = check_operands(o);
    if (rc) return rc;
    if (cdiv(M, BM) > num_sms()) return set_error(AWD_ELIMIT, "fused sampler: at most %d streams per call", num_sms() * BM);
    TcEpiSample epi{};
    epi.sp = sp; epi.bias = bias; epi.part = (float2*)part; epi.zwin = zwin; epi.banned_keys = banned_keys;
    epi.slots = tc_sample_slots(M, N);
    epi.tiles_m = (int)cdiv(M, BM);
    epi.M = M;
    epi.ban_lo = sp.n_banned ? sp.banned[0] : 0x7fffffff; epi.ban_hi = sp.n_banned ? sp.banned[sp.n_banned - 1] : -1;
    AWD_CUDA(cudaMemsetAsync(epi.row_bound_host(), 0x80, (size_t)M * kRowBoundStride * sizeof(int), st));    // "no key yet" (a huge negative key)
    if (use_pair(M, true)) return launch_tc_pair<TcEpiSample>(o, M, N, epi, st, /*row_stationary=*/true);
    return launch_tc<256, TcEpiSample>(o, M, N, epi, st, /*row_stationary=*/true);
}

int launch_tc_sample_shard(const TcOperands& o, int M, int n_local, int v0, const SampleParams& sp, const float* bias,
                           float* part, float* zwin, float* banned_keys, cudaStream_t st) {
    int rc = check_operands(o);
    if (rc) return rc;
    if (v0 % 256) return set_error(AWD_EINVAL, "vocabulary shards must start at a multiple of 256 (whole decoder tiles)");
    if (cdiv(M, BM) > num_sms()) return set_error(AWD_ELIMIT, "fused sampler: at most %d streams per call", num_sms() * BM);
    TcEpiSample epi{};
    epi.sp = sp; epi.bias = bias; epi.part = (float2*)part; epi.zwin = zwin; epi.banned_keys = banned_keys;
    epi.slots = tc_sample_slots(M, n_local);
    epi.tiles_m = (int)cdiv(M, BM);
    epi.M = M;
    epi.ban_lo = sp.n_banned ? sp.banned[0] : 0x7fffffff; epi.ban_hi = sp.n_banned ? sp.banned[sp.n_banned - 1] : -1;
    AWD_CUDA(cudaMemsetAsync(epi.row_bound_host(), 0x80, (size_t)M * kRowBoundStride * sizeof(int), st));
    epi.n_off = v0;
    return launch_tc<256, TcEpiSample>(o, M, n_local, epi, st, /*row_stationary=*/true);
}

int launch_shard_partial(const float* part, const float* zwin, const float* banned_keys, int slots, int bs, int v0, int v1,
                         const SampleParams& sp, float* out_key, int64_t* out_id, float* out_banned, cudaStream_t st) {
    ShardPartialArgs a{};
    a.part = (const float2*)part; a.zwin = zwin; a.banned_keys = banned_keys; a.slots = slots; a.bs = bs; a.v0 = v0; a.v1 = v1;
    a.sp = sp; a.out_key = out_key; a.out_id = out_id; a.out_banned = out_banned;
    shard_partial_kernel<<<(unsigned)cdiv(bs, 4), 128, 0, st>>>(a);
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

int launch_shard_merge(const float* keys, const int64_t* ids, const float* banned, int R, int bs, const SampleParams& sp,
                       int64_t* out, cudaStream_t st, int64_t stride_k, int64_t stride_i, int64_t stride_b) {
    if (stride_k == 0) { stride_k = bs; stride_i = bs; stride_b = (int64_t)bs * 64; }
    shard_merge_kernel<<<(unsigned)cdiv(bs, 4), 128, 0, st>>>(keys, ids, banned, R, bs, stride_k, stride_i, stride_b, sp, out);
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

int launch_sample_finalize(const float* part, const float* zwin, const float* banned_keys, int slots, int bs, int V,
                           const SampleParams& sp, int64_t* out, const void* table, int ld_table, void* xnext,
                           cudaStream_t st, const int64_t* forced, const int32_t* forced_len, int step) {
    if (bs <= 0) return AWD_OK;
    FinalizeArgs a{};
    a.part = (const float2*)part; a.zwin = zwin; a.banned_keys = banned_keys; a.tiles_n = slots; a.bs = bs; a.V = V; a.sp = sp;
    a.out = out; a.table = (const uint4*)table; a.row_vec4 = ld_table / 8; a.xnext = (uint4*)xnext;
    a.forced = forced; a.forced_len = forced_len; a.step = step;
    sample_finalize_kernel<<<(unsigned)cdiv(bs, 4), 128, 0, st>>>(a);
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

// =============================================================================================
// Persistent generation kernel: ALL tokens of awd_generate's loop in ONE launch.
//
// The per-token chain  decoder+race -> finalize+gather -> layer 0 -> layer 1 -> ... is strictly serial,
// and as separate launches each link pays launch latency, barrier/TMEM set-up, a cold TMA pipeline and an
// exposed epilogue (measured ~8-10 us per kernel, 5 kernels per token).  Here one CTA per SM stays
// resident, keeps its TMEM allocation and warp roles, and walks the phases of every token with a
// grid-wide barrier between phases (an atomic counter in global memory; the launch is cooperative so
// that all CTAs are co-resident).  Phase bodies are the same TMA -> tcgen05 -> TMEM pipelines as
// tc_gemm_kernel (single CTA, streaming A and B).
// =============================================================================================
constexpr int MEGA_MAXL = 4;

constexpr int kTraceTiles = 128 + 148 * 8;     // AWD_MEGA_TRACE: per-tile stamps of one CTA's decoder phase start here ([16][4])
struct MegaParams {
    CUtensorMap tm_x;                       // gathered layer-0 input  [bs][Epad]
    CUtensorMap tm_hb[MEGA_MAXL][2];        // hidden operands, ping-pong
    CUtensorMap tm_w[MEGA_MAXL];            // [W_ih | W_hh], box rows = bn[l]
    CUtensorMap tm_dec;                     // decoder rows, box rows = 256
    CUtensorMap tm_dec_r;                   // same matrix, box rows = dec_col_w % 256 (the narrower last tile of a group's column range)
    CUtensorMap tm_dec_x;                   // same matrix, box rows = dec_x_w % 256 (... of the left-over CTAs' range)
    int n_layers, bs, V, n_steps, p0, Epad;
    int H[MEGA_MAXL], Ipad[MEGA_MAXL], Hpad[MEGA_MAXL], bn[MEGA_MAXL];
    const float* bias[MEGA_MAXL];
    float* c[MEGA_MAXL];                    // cell state, BLOCKED [Hpad / 8][bs][8] (TcEpiCell::cell_at): launch_mega_generate converts from / to the engine's rows
    float* h[MEGA_MAXL];
    __nv_bfloat16* hb[MEGA_MAXL][2];
    const float* dec_bias;
    SampleParams sp;
    float2* part;
    float* zwin;
    float* banned_keys;
    int ban_lo, ban_hi;                     // range of the disallowed ids (see TcEpiSampleT)
    int slots, dec_ctas;                    // decoder phase: row-stationary over dec_ctas = groups * tiles_m CTAs
    int dec_col_w;                          // vocabulary columns per group (PhaseTiles); 0: whole 256-column tiles round-robin
    int dec_x_n0, dec_x_w;                  // column range of the CTAs left over after the last whole group (0: none)
    int64_t* gen;                           // [n_steps][bs]
    const int64_t* forced;                  // awd_generate_resume: [forced_F][bs] / [bs] (see FinalizeArgs), or nullptr
    const int32_t* forced_len;
    int forced_F;
    const uint4* table;
    uint4* xb;
    unsigned* sync_counter;                 // zeroed before launch
    unsigned long long* phase_ns;           // optional [5]: time spent in decoder, finalize, layers 0..2 (CTA 0)
    // L2 eviction priority of the weight streams (chosen in launch_mega_generate)
    uint64_t dec_policy, lstm_policy;
    // <= 128 streams (one M tile): LSTM layers run as a 2-D (N tile x K slice) split, see sb_lstm_phase
    int sb;                                 // 1: use it
    int sb_prefetch;                        // decoder K blocks per CTA asked into L2 at the end of the LSTM phase
    int sb_stages;                          // operand ring of that phase: 6 x (16 + 16) KB, or 4 x (16 + 32) KB when a layer uses 256-wide tiles
    int sb_tiles[MEGA_MAXL], sb_gk[MEGA_MAXL], sb_cta0[MEGA_MAXL];
    float4* sb_scratch;                     // partial gate tiles exchanged between the K slices of an N tile, one region per layer
    int64_t sb_scratch_off[MEGA_MAXL];      // (in float4)
    unsigned* sb_counter;                   // [2][MEGA_MAXL][kSbCnt]: partials stored / pieces finished, per (layer, N tile, M tile), then x_cnt; monotonic over a launch
    long long* trace;                       // debugging (AWD_MEGA_TRACE): clock64 stamps of CTA 0 in one phase, else nullptr
};

// Grid-wide barrier (all CTAs co-resident: cooperative launch).  Same shape as cooperative groups'
// grid.sync(): the CTA barrier orders every thread's writes before thread 0's gpu-scope fence + arrive,
// and thread 0's fence after the spin orders the peers' writes before everything that follows the second
// CTA barrier.  Data written by OTHER CTAs is then read either by TMA (L2) or with ld.global.cg, never
// through a possibly stale L1 line, so no per-thread fence is needed.
__device__ __forceinline__ void grid_sync(unsigned* counter, unsigned& epoch) {
    asm volatile("fence.proxy.async.global;" ::: "memory");     // this thread's generic-proxy stores -> later TMA reads
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

// One-way counterpart of grid_sync: every CTA bumps `counter` once all its threads' writes are out.  After the decoder phase of the
// barrier-free schedule only the finalize step depends on the other CTAs (it merges their partial race results): its warps wait for
// this counter themselves (sb_lstm_layers), while the producer and the MMA issuer go straight on to layer 0's recurrent half -- the
// stragglers of the decoder phase and the barrier's latency no longer hold the next phase up.  (A counter of its own: CTAs without
// LSTM work reach the next grid_sync at once, so grid_sync's counter cannot tell how many CTAs are through with the decoder.)
__device__ __forceinline__ void grid_arrive(unsigned* counter) {
    asm volatile("fence.proxy.async.global;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(counter, 1u);
    }
}

// Tiles of one CTA in one phase of the persistent kernel.
//   LSTM phases: all (M tile, N tile) pairs, dealt round-robin over the CTAs.
//   Decoder phase (row-stationary): a CTA keeps its M tile; the tiles_m CTAs of a group walk the same columns.  With g.col_w the
//   vocabulary is cut into ONE contiguous range of col_w columns per group -- whole tiles of BN, then one narrower tile -- so that
//   every group multiplies the same number of columns: dealing whole 256-column tiles leaves 235 tiles on 148 / 74 / 37 / 18
//   groups at 2 / 4 / 7 / 14 tiles for the slowest group against 1.59 / 3.18 / 6.35 / 13.06 on average.
struct PhaseTiles {
    int tm, n0, bw;                         // M tile, first column, width (multiple of 32, <= BN)
    int cur, step, end, tiles_m, BN, N, col_w, grp0;
    int xe, xE;                             // left-over CTA: its index among them and their number (xE = 0: a CTA of a whole group)
    bool res;
    __device__ __forceinline__ PhaseTiles(const GemmShape& g, int cta, int ncta, int bn) {
        res = g.row_stat != 0; tiles_m = g.tiles_m; BN = bn; N = g.N; col_w = res ? g.col_w : 0; xE = 0; xe = 0;
        if (res) {
            tm = cta % tiles_m;
            if (col_w && cta >= ncta) {         // (only reached with g.x_w > 0)
                xe = cta - ncta; xE = (int)gridDim.x - ncta;
                cur = 0; step = 1; col_w = g.x_w; grp0 = g.x_n0;
                end = ((tiles_m + xE - 1) / xE) * ((col_w + BN - 1) / BN);      // passes x tiles per pass
            } else if (col_w) { cur = 0; step = 1; end = (col_w + BN - 1) / BN; grp0 = (cta / tiles_m) * col_w; }
            else { cur = cta / tiles_m; step = ncta / tiles_m; end = g.tiles_n; grp0 = 0; }
        } else { cur = cta; step = ncta; end = tiles_m * g.tiles_n; grp0 = 0; }
    }
    __device__ __forceinline__ bool valid() {
        if (cur >= end) return false;
        if (!res) { tm = cur % tiles_m; n0 = (cur / tiles_m) * BN; bw = BN; return true; }
        int t = cur;
        if (xE) {
            // The CTAs left over after the last whole group (148 - 18 x 8 = 4 at eight M tiles) take the tail of the vocabulary
            // for all M tiles in turn: pass p gives CTA xe the M tile p * xE + xe.  Their range is shorter by the number of passes.
            const int tpp = (col_w + BN - 1) / BN;
            tm = (cur / tpp) * xE + xe;
            if (tm >= tiles_m) return false;    // (later passes only have larger M tiles)
            t = cur % tpp;
        }
        n0 = grp0 + t * BN;
        bw = col_w ? min(BN, col_w - t * BN) : BN;
        return n0 < N;
    }
    __device__ __forceinline__ void next() { cur += step; }
};

// One GEMM phase of the persistent kernel (single CTA per tile, no cluster, A and B streamed).
// BN_T > 0: tile width fixed at compile time (decoder).  BN_T == 0: width `bn_rt` in {32, 64, 128, 256} chosen
// per layer on the host (LSTM phases) -- ONE instantiation serves all widths, so that the small-batch tile
// shapes cost the persistent kernel no registers.  Ring depth: 4 stages of 48 KB at 256, 6 of 32 KB at 128,
// 8 of (16 + bn/8) KB below; the barriers sit behind the largest ring (192 KB) in every case.
template <int BN_T, class Epi, bool kConv = false>
__device__ __forceinline__ void mega_gemm_phase(uint8_t* gen_base, uint32_t base, uint32_t tmem_base, bool first,
                                                const CUtensorMap* tmA0, const CUtensorMap* tmA1, const CUtensorMap* tmB,
                                                const GemmShape& g, const Epi& epi, int cta, int ncta, uint64_t b_policy,
                                                int bn_rt = 0, long long* tr = nullptr, const CUtensorMap* tmB_r = nullptr, long long* tr_t = nullptr,
                                                const unsigned* a_wait = nullptr, int a_wait_n = 0, unsigned a_wait_target = 0) {
    const int BN = BN_T ? BN_T : bn_rt;
    if (tr && threadIdx.x == 0) tr[0] = clock64();
    const int STAGES = BN >= 256 ? 4 : (BN >= 128 ? 6 : 8);
    const uint32_t B_STAGE = (uint32_t)BN * BK * 2;
    const uint32_t sA = base, sB = base + STAGES * A_STAGE_BYTES;
    const uint32_t bars = base + 4 * 49152;                      // fixed place for every phase (after the largest ring)
    const uint32_t full0 = bars, empty0 = bars + 64, tfull0 = bars + 128, tempty0 = bars + 144;
    float* sbias = reinterpret_cast<float*>(gen_base + 4 * 49152 + 256);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int num_kb = g.kb0 + g.kb1;

    // Arm the 20 phase barriers (full[8], empty[8], tfull[2], tempty[2]) with 20 threads at once (one thread doing the
    // 40 operations in turn cost ~0.5 us at the head of every phase).  All invalidations first, then all
    // initialisations: interleaving them across threads hangs the kernel (an mbarrier.inval next to a freshly
    // initialised neighbour seems to undo it).
    {
        const int i = threadIdx.x - 64;                       // 20 lanes of the first epilogue warp
        const bool mine = i >= 0 && i < 20;
        const uint32_t bar = bars + (i < 16 ? 8 * i : 128 + 8 * (i - 16));
        if (!first) {
            if (mine) asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
            __syncthreads();
        }
        if (mine) {
            mbar_init(bar, i < 18 ? 1 : NUM_EPI_THREADS / 32);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
    }
    __syncthreads();
    tc_fence_after();
    // A CTA without tiles in this phase leaves -- but with a_wait not before the phase's inputs are complete: the counters of this
    // schedule count across tokens (targets grow with the token), and a CTA that ran ahead would bump the counter behind this phase
    // (grid_arrive) for the next token while others still wait for this token's count.
    const bool has_tiles = !(cta >= ncta && !(g.row_stat && g.x_w > 0));
    if (!has_tiles && a_wait == nullptr) return;

    if (warp == 0) {
        // a_wait: the A operand is written by other CTAs in the phase before, which no grid barrier separates from this one; its
        // producers bump the a_wait_n counters (to a_wait_target when complete).  The weight tiles of the first ring turn do not
        // depend on anything and are requested first; then all 32 lanes poll the counters; then the activations follow.
        int pre = 0;
        if (a_wait != nullptr) {
            if (lane == 0 && has_tiles) {
                PhaseTiles it(g, cta, ncta, BN);
                if (it.valid()) {
                    const CUtensorMap* tb = it.bw == BN ? tmB : tmB_r;
                    for (; pre < STAGES && pre < num_kb; ++pre) {
                        mbar_expect_tx(full0 + 8 * pre, A_STAGE_BYTES + (uint32_t)it.bw * BK * 2);
                        tma_load_2d_hint(sB + pre * B_STAGE, tb, g.b_k0 + pre * BK, it.n0, full0 + 8 * pre, b_policy);
                    }
                }
            }
            for (unsigned spin = 0;; ++spin) {
                bool ok = true;
                for (int i = lane; i < a_wait_n; i += 32) {
                    unsigned c;
                    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(c) : "l"(a_wait + i) : "memory");
                    ok = ok && (int)(c - a_wait_target) >= 0;
                }
                if (__all_sync(0xffffffffu, ok)) break;
                if (spin > (1u << 26)) __trap();
            }
            asm volatile("fence.proxy.async.global;" ::: "memory");
            pre = __shfl_sync(0xffffffffu, pre, 0);
        }
        if (lane == 0 && has_tiles) {
            int stage = 0;
            uint32_t phase = 0;
            for (PhaseTiles it(g, cta, ncta, BN); it.valid(); it.next()) {
                const int m0 = it.tm * BM, n0 = it.n0;
                // (the narrower last tile of a column range comes through tmB_r, whose box has exactly its rows)
                const CUtensorMap* tb = it.bw == BN ? tmB : tmB_r;
                const uint32_t tx = A_STAGE_BYTES + (uint32_t)it.bw * BK * 2;
                for (int kb = 0; kb < num_kb; ++kb) {
                    if (pre > 0) {              // first ring turn of the phase: the barrier is armed and the weight tile on its way
                        --pre;
                        const uint32_t fb = full0 + 8 * stage;
                        if (kb < g.kb0) tma_load_2d(sA + stage * A_STAGE_BYTES, tmA0, kb * BK, g.a0_row + m0, fb);
                        else tma_load_2d(sA + stage * A_STAGE_BYTES, tmA1, (kb - g.kb0) * BK, g.a1_row + m0, fb);
                        if (++stage == STAGES) { stage = 0; phase ^= 1; }
                        continue;
                    }
                    mbar_wait(empty0 + 8 * stage, phase ^ 1);
                    const uint32_t fb = full0 + 8 * stage;
                    mbar_expect_tx(fb, tx);
                    if (kb < g.kb0) tma_load_2d(sA + stage * A_STAGE_BYTES, tmA0, kb * BK, g.a0_row + m0, fb);
                    else tma_load_2d(sA + stage * A_STAGE_BYTES, tmA1, (kb - g.kb0) * BK, g.a1_row + m0, fb);
                    tma_load_2d_hint(sB + stage * B_STAGE, tb, g.b_k0 + kb * BK, n0, fb, b_policy);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (!has_tiles) {
    } else if (warp == 1) {
        if (kConv || lane == 0) {       // kConv: all 32 lanes, one elected issuer per instruction (see umma_bf16_elect)
            int stage = 0, acc = 0;
            uint32_t phase = 0, acc_phase = 0;
            int tile_no = 0;
            for (PhaseTiles it(g, cta, ncta, BN); it.valid(); it.next(), ++tile_no) {
                const uint32_t idesc = umma_idesc_bf16(it.bw);
                mbar_wait(tempty0 + 8 * acc, acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(full0 + 8 * stage, phase);
                    if (tr && lane == 0 && kb == 0 && tile_no < 2) tr[1 + tile_no] = clock64();
                    if (tr_t && lane == 0 && kb == 0 && tile_no < 16) tr_t[tile_no * 4] = clock64();
                    tc_fence_after();
                    const uint64_t adesc = umma_desc_sw128(sA + stage * A_STAGE_BYTES);
                    const uint64_t bdesc = umma_desc_sw128(sB + stage * B_STAGE);
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k)
                        if constexpr (kConv) umma_bf16_elect(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
                        else umma_bf16(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
                    if constexpr (kConv) umma_commit_elect(empty0 + 8 * stage); else umma_commit(empty0 + 8 * stage);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                if constexpr (kConv) umma_commit_elect(tfull0 + 8 * acc); else umma_commit(tfull0 + 8 * acc);
                if (tr && lane == 0 && tile_no < 2) tr[3 + tile_no] = clock64();
                if (tr_t && lane == 0 && tile_no < 16) tr_t[tile_no * 4 + 1] = clock64();
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else {
        const int ew = warp - 2;
        const int q = warp & 3, part = ew >> 2;                 // TMEM lane quarter, column part
        const int et = ew * 32 + lane;
        int acc = 0;
        uint32_t acc_phase = 0;
        typename Epi::State es;
        epi.init(es);
        int tile_no = 0;
        int cur_tm = -1;                                        // row-stationary: the M tile whose running result `es` holds
        auto finish_rows = [&](int tm_done) {
            const int row = tm_done * BM + q * 32 + lane;
            epi.finish(es, row, row < g.M, (cta / g.tiles_m) * EPI_PARTS + part);
        };
        for (PhaseTiles it(g, cta, ncta, BN); it.valid(); it.next(), ++tile_no) {
            const int tm = it.tm, n0 = it.n0;
            if (g.row_stat && tm != cur_tm) {                   // (a left-over CTA moving on to its next M tile)
                if (cur_tm >= 0) { finish_rows(cur_tm); epi.init(es); }
                cur_tm = tm;
            }
            const int c0 = ((it.bw / 32) * part / EPI_PARTS) * 32, c1 = ((it.bw / 32) * (part + 1) / EPI_PARTS) * 32;
            const int row = tm * BM + q * 32 + lane;
            const bool row_ok = row < g.M;
            float* sb = sbias + acc * BN;
            if (epi.stage_bias()) {                       // uniform over the CTA
                if (et < BN) {
                    const int n = n0 + et;
                    sb[et] = (epi.bias != nullptr && n < g.N) ? __ldg(epi.bias + n) * epi.bias_scale() : 0.f;
                }
                epi_bar_sync();
            }
            epi.tile_begin(es, row, row_ok, n0 + c0, (c1 - c0) / 32);
            if (tr && et == 0 && tile_no == 0) tr[9] = clock64();
            mbar_wait(tfull0 + 8 * acc, acc_phase);
            if (tr && et == 0 && tile_no == 0) tr[10] = clock64();
            if (tr_t && et == 0 && tile_no < 16) tr_t[tile_no * 4 + 2] = clock64();
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN;
#pragma unroll 1
            for (int c = c0; c < c1; c += 32) {
                float v[32];
                tmem_ld32(taddr + c, v);
                if (tr && et == 0 && tile_no == 0 && c == c0) tr[15] = clock64();
                if (n0 + c < g.N) epi.chunk(es, row, row_ok, n0 + c, g.N, v, sb + c, nullptr);
                if (tr && et == 0 && tile_no == 0) tr[11 + ((c - c0) >> 5)] = clock64();
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty0 + 8 * acc);
            epi.tile_end(es, row, row_ok, (n0 / BN) * EPI_PARTS + part);
            if (tr && et == 0 && tile_no < 2) tr[5 + tile_no] = clock64();
            if (tr_t && et == 0 && tile_no < 16) tr_t[tile_no * 4 + 3] = clock64();
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if (g.row_stat) finish_rows(cur_tm >= 0 ? cur_tm : cta % g.tiles_m);    // (a CTA without columns still owns its partial slot)
    }
    tc_fence_before();
}


// =============================================================================================
// One LSTM layer step for <= 128 streams (ONE M tile) as a 2-D split.
//
// A tcgen05.mma of K = 16 occupies the SM's tensor pipe for ~128 clocks whatever its N (<= 256) is
// (profiles/micro/mma_probe.cu: 142 clocks per instruction for every M in {64,128}, N in {16..256}), so with one
// M tile the time of a layer is (MMA instructions per CTA) x 128 clocks.  Splitting only N over the SMs (tiles
// 32 wide, 144 CTAs) leaves every CTA the whole K loop: 144 dependent-rate instructions = 9.4 us for layer 1,
// independent of operand traffic, clusters or multicast (profiles/micro/ingest_probe.cu).  Here a layer is cut
// into N tiles of 128 gate columns x gk K slices (36 x 4 CTAs for 4 x 1150 gates): K/16/gk instructions per CTA.
// The gk partial tiles of an N tile are summed through L2: CTA (tile, slice) stores the 32-column chunks that
// other slices finish into `scratch`, bumps the tile's counter (release), waits until all gk slices have arrived
// (acquire), adds the others' partials to the chunk(s) it finishes -- always in slice order, so the sum does not
// depend on timing -- and applies bias, gates and the cell update to them.
// =============================================================================================
__device__ __forceinline__ unsigned long long globaltimer_ns();
constexpr int SB_BN = 128;
constexpr int kSbCnt = 160;          // counters per layer: (N tile, M tile) pairs
constexpr int SB_STAGES = 6;

// All LSTM layers of one token run as ONE phase of the persistent kernel: the barriers are armed once, the operand ring and its
// parities run through the layers, the accumulator alternates between two TMEM buffers, and the three warp roles walk the layers
// independently.  There is no grid barrier between layers: a CTA that has finished its pieces of an N tile (32 hidden units)
// bumps done[layer][tile], and the TMA producer of the next layer waits, K block by K block, only for the two tiles that wrote
// that block of h.  A CTA's K blocks are interleaved (kb = slice mod gk) and ordered recurrent half first -- h(t-1) of the layer
// itself has been complete since the previous token -- so the producer and the MMA issuer run ahead into the next layer's
// recurrent half (and request its weight tiles) while the epilogue warps are still exchanging partials of this one; only the
// activation loads of the input half wait for the layer below.
struct SbLayerView {
    const CUtensorMap *tmA0, *tmA1, *tmB;
    int kb0, kb1, N, n_tiles, gk;       // N tiles of bn gate columns x gk K slices, for each of the tiles_m M tiles
    int bn;                             // 128 (gk = 4, 2, 1), or 256 with gk = 1 (full-width MMAs, no partial exchange)
    int cta0;                           // the layer's n_tiles * tiles_m * gk CTAs start here (see launch_mega_generate: with 2..4 M tiles
                                        //  neighbouring layers sit on disjoint CTAs)
    const float* bias;
    float* c;                    // blocked [Hpad / 8][c_rows][8] (TcEpiCell::cell_at)
    int c_rows;
    __nv_bfloat16* hb_out;
    int ld;                      // row pitch of hb_out (bf16): Hpad
    float4* scratch;
    unsigned* part_cnt;          // partials stored, per tile
    unsigned* done_cnt;          // pieces finished, per tile
    const unsigned* wait_cnt;    // what the input half waits for: the done counters of the layer below, one per tile of 32 units
    int wait_tiles;              //   (wait_tiles of them), or ONE counter for the whole input (wait_tiles == 0: layer 0's gathered rows);
    int wait_per_token;          //   arrivals per counter and token; nullptr: the input is complete (behind a grid barrier)
    int wait_tpk;                //   tiles of the layer below per 64-unit K block: 2 (128-column tiles) or 1 (256-column tiles)
};

// enumeration of a CTA's K blocks in one layer (shared by the producer and the MMA issuer)
struct SbBlocks {
    int hh_first, n_hh, n_ih, n_blk, slice, gk;
    __device__ __forceinline__ SbBlocks(const SbLayerView& v, int slice_) : slice(slice_), gk(v.gk) {
        const int nkb = v.kb0 + v.kb1;
        hh_first = v.kb0 + ((slice - v.kb0) % gk + gk) % gk;
        n_hh = hh_first < nkb ? (nkb - 1 - hh_first) / gk + 1 : 0;
        n_ih = slice < v.kb0 ? (v.kb0 - 1 - slice) / gk + 1 : 0;
        n_blk = n_hh + n_ih;
    }
    __device__ __forceinline__ int kb(int idx) const { return idx < n_hh ? hh_first + idx * gk : slice + (idx - n_hh) * gk; }
};

// kGk: 0 = any mix of K-slice counts, 1 = every layer has one K slice (no exchange code), 2 = every layer has 2 or 4 (no plain epilogue)
template <int kGk, class ViewFn>
__device__ __forceinline__ void sb_lstm_layers(uint8_t* gen_base, uint32_t base, uint32_t tmem_base, int n_layers, int M, int tiles_m, unsigned token,
                                               ViewFn view, int cta, uint64_t b_policy, const FinalizeArgs* fin, const SampleParams* fin_sp, const unsigned* dec_cnt, unsigned dec_target, unsigned* x_cnt,
                                               const CUtensorMap* tm_next, int next_kb, int next_tiles, int next_step, int next_col_w, int next_depth, int STAGES,
                                               long long* tr, long long* gtr = nullptr) {
    // ring: STAGES slots of (16 KB activations, b_slot weights): 6 x (16 + 16) KB when every layer uses 128-wide tiles, 4 x (16 + 32) KB otherwise
    const uint32_t b_slot = STAGES == 6 ? 16384u : 32768u;
    const uint32_t sA = base, sB = base + STAGES * A_STAGE_BYTES;
    const uint32_t bars = base + 4 * 49152;
    const uint32_t full0 = bars, empty0 = bars + 64, tfull0 = bars + 128, tempty0 = bars + 144;
    float* sbias = reinterpret_cast<float*>(gen_base + 4 * 49152 + 256);          // [2][256]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (tr && threadIdx.x == 0) tr[0] = clock64();
    if (gtr && threadIdx.x == 0) gtr[0] = (long long)globaltimer_ns();
    {   // arm the phase barriers (see mega_gemm_phase; this is never the first phase of the kernel)
        const int i = threadIdx.x - 64;
        const bool mine = i >= 0 && i < 20;
        const uint32_t bar = bars + (i < 16 ? 8 * i : 128 + 8 * (i - 16));
        if (mine) asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
        __syncthreads();
        if (mine) {
            mbar_init(bar, i < 18 ? 1 : NUM_EPI_THREADS / 32);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
    }
    __syncthreads();
    tc_fence_after();

    if (warp == 0) {
        // TMA producer.  All 32 lanes run the loop, one elected lane issues each instruction (AWD_SB_PCONV; see umma_bf16_elect); all
        // lanes poll the counters of the layer below together (one L2 round trip for all tiles this CTA depends on, instead of one per
        // K block in turn: ~1400 clocks each, measured).
        int g = 0;                                              // K blocks loaded so far in this phase: stage g % STAGES, use g / STAGES
        for (int l = 0; l < n_layers; ++l) {
            const SbLayerView v = view(l);
            const int lc = cta - v.cta0;                          // index among the CTAs of this layer
            if (lc < 0 || lc >= v.n_tiles * v.gk * tiles_m) continue;
            // CTA -> (N tile, M tile, K slice): the gk slices of a (tile, M tile) pair are neighbours
            const int pair = lc / v.gk, tile = pair / tiles_m, mt = pair - tile * tiles_m, n0 = tile * v.bn, m0 = mt * BM;
            const uint32_t B_STAGE = (uint32_t)v.bn * BK * 2, STAGE_TX = A_STAGE_BYTES + B_STAGE;
            const SbBlocks blk(v, lc - pair * v.gk);
            int issued_b = 0;                                   // weight tiles of this layer already requested (lane 0)
            auto issue_b = [&](int idx) {
                const int gg = g + idx, st = gg % STAGES;
                mbar_wait(empty0 + 8 * st, ((gg / STAGES) & 1) ^ 1);
                mbar_expect_tx_elect(full0 + 8 * st, STAGE_TX);
                tma_load_2d_hint_elect(sB + st * b_slot, v.tmB, blk.kb(idx) * BK, n0, full0 + 8 * st, b_policy);
            };
            {
                // recurrent half: h(t-1) of this layer is complete, nothing to wait for
                for (int idx = 0; idx < blk.n_hh; ++idx) {
                    issue_b(idx);
                    tma_load_2d_elect(sA + ((g + idx) % STAGES) * A_STAGE_BYTES, v.tmA1, (blk.kb(idx) - v.kb0) * BK, m0, full0 + 8 * ((g + idx) % STAGES));
                }
                issued_b = blk.n_hh;
                // weight tiles of the input half whose ring slot is free or held by a recurrent block (all of whose loads are out)
                for (; issued_b < blk.n_blk && issued_b - STAGES < blk.n_hh; ++issued_b) issue_b(issued_b);
            }
            if (v.wait_cnt != nullptr && blk.n_ih > 0) {
                const unsigned target = (unsigned)v.wait_per_token * (token + 1);
                for (unsigned spin = 0;; ++spin) {
                    bool ok = true;
                    for (int i = lane; i < (v.wait_tiles ? v.wait_tpk * blk.n_ih : 1); i += 32) {      // the tiles that write units [64 kb, 64 kb + 64)
                        const int t = v.wait_tiles ? v.wait_tpk * blk.kb(blk.n_hh + i / v.wait_tpk) + i % v.wait_tpk : 0;
                        if (t < v.wait_tiles || v.wait_tiles == 0) {
                            unsigned c;
                            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(c) : "l"(v.wait_cnt + (v.wait_tiles ? t * tiles_m + mt : 0)) : "memory");
                            ok = ok && (int)(c - target) >= 0;
                        }
                    }
                    if (__all_sync(0xffffffffu, ok)) break;
                    if (spin > (1u << 26)) __trap();
                }
                asm volatile("fence.proxy.async.global;" ::: "memory");
                if (tr && lane == 0 && l == 0) tr[15] = clock64();
            }
            __syncwarp();
            {
                for (int idx = blk.n_hh; idx < blk.n_blk; ++idx) {
                    if (issued_b <= idx) { issue_b(idx); issued_b = idx + 1; }
                    tma_load_2d_elect(sA + ((g + idx) % STAGES) * A_STAGE_BYTES, v.tmA0, blk.kb(idx) * BK, m0, full0 + 8 * ((g + idx) % STAGES));
                }
            }
            g += blk.n_blk;
        }
        // HBM idles while the last layer finishes (the LSTM weights live in L2): the first K blocks of the decoder tile this CTA
        // multiplies next are asked into L2 now, so that the decoder phase does not start from a cold 28 MB burst
        if (lane == 0 && tm_next != nullptr && cta / tiles_m < next_tiles)
            for (int kb = 0; kb < next_depth && kb < 2 * next_kb; ++kb)      // first tile, then the second one (PhaseTiles order)
                tma_prefetch_l2_2d(tm_next, (kb % next_kb) * BK,
                                   next_col_w ? (cta / tiles_m) * next_col_w + (kb / next_kb) * 256 : (cta / tiles_m + (kb / next_kb) * next_step) * 256);
    } else if (warp == 1) {
        {   // all 32 lanes, one elected issuer per instruction (see umma_bf16_elect)
            int g = 0, used = 0;                                // `used`: layers this CTA took part in so far (accumulator ping-pong)
            for (int l = 0; l < n_layers; ++l) {
                const SbLayerView v = view(l);
                const int lc = cta - v.cta0;
                if (lc < 0 || lc >= v.n_tiles * v.gk * tiles_m) continue;
                const SbBlocks blk(v, lc % v.gk);
                const uint32_t idesc = umma_idesc_bf16(v.bn);
                const int buf = used & 1;
                mbar_wait(tempty0 + 8 * buf, ((used >> 1) & 1) ^ 1);      // the epilogue has drained this buffer (two layers ago)
                tc_fence_after();
                for (int idx = 0; idx < blk.n_blk; ++idx) {
                    const int gg = g + idx, st = gg % STAGES;
                    mbar_wait(full0 + 8 * st, (gg / STAGES) & 1);
                    if (tr && lane == 0 && idx == 0) tr[16 * l + 2] = clock64();
                    if (tr && lane == 0 && idx == blk.n_hh) tr[16 * l + 11] = clock64();
                    if (gtr && lane == 0 && idx == blk.n_hh && l < 3) gtr[5 + l] = (long long)globaltimer_ns();
                    tc_fence_after();
                    const uint64_t adesc = umma_desc_sw128(sA + st * A_STAGE_BYTES);
                    const uint64_t bdesc = umma_desc_sw128(sB + st * b_slot);
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k)
                        umma_bf16_elect(tmem_base + buf * 256, adesc + 2 * k, bdesc + 2 * k, idesc, idx > 0 || k != 0);
                    umma_commit_elect(empty0 + 8 * st);
                }
                umma_commit_elect(tfull0 + 8 * buf);
                if (tr && lane == 0) tr[16 * l + 3] = clock64();
                g += blk.n_blk;
                ++used;
            }
        }
    } else {
        const int ew = warp - 2;
        const int q = warp & 3, part = ew >> 2;
        const int et = ew * 32 + lane;
        // The token itself (main.py:63-71): CTA r resolves stream r -- merge of the decoder's partial race results, first allowed
        // candidate, id out, embedding row gathered as layer 0's input -- in its first epilogue warp, while the producer and the
        // MMA issuer already work on layer 0's recurrent half.  Layer 0's input-half loads wait on x_cnt instead of a grid barrier.
        if (fin != nullptr && cta + (int)gridDim.x * ew < M) {       // (several M tiles: epilogue warp w takes stream r + w x grid size)
            // every CTA's partial race results of this token must be there (grid_arrive after the decoder phase)
            for (unsigned spin = 0;; ++spin) {
                unsigned c;
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(c) : "l"(dec_cnt) : "memory");
                if ((int)(c - dec_target) >= 0) break;
                if (spin > (1u << 26)) __trap();
            }
            if (tr && lane == 0) tr[12] = clock64();
            finalize_row(*fin, *fin_sp, token, cta + (int)gridDim.x * ew, lane);
            if (tr && lane == 0) tr[13] = clock64();
            asm volatile("fence.proxy.async.global;" ::: "memory");
            __syncwarp();
            if (lane == 0) asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(x_cnt) : "memory");
            if (tr && lane == 0) tr[14] = clock64();
        }
        int used = 0;
        for (int l = 0; l < n_layers; ++l) {
            const SbLayerView v = view(l);
            const int lc = cta - v.cta0;
            if (lc < 0 || lc >= v.n_tiles * v.gk * tiles_m) continue;
            const int gk = v.gk, N = v.N;
            const int BN = v.bn;
            const int pair = lc / gk, slice = lc - pair * gk, tile = pair / tiles_m, mt = pair - tile * tiles_m, n0 = tile * BN;
            const int row = mt * BM + q * 32 + lane, lrow = q * 32 + lane;          // stream; row inside the M tile (scratch index)
            const bool row_ok = row < M;
            const int buf = used & 1;
            float* sb_l = sbias + buf * 256;              // (the buffer's previous user, two layers ago, is past its last read: see the
            if (et < BN) {                                //  epi_bar_sync calls of the layer in between)
                const int n = n0 + et;
                sb_l[et] = (v.bias != nullptr && n < N) ? __ldg(v.bias + n) : 0.f;
            }
            epi_bar_sync();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + buf * 256;
            if (kGk == 1 || (kGk == 0 && gk == 1)) {
                // ---- one K slice: the accumulator is complete, the plain cell epilogue (old cell state prefetched while the MMAs run)
                TcEpiCell epi{};
                epi.addend = nullptr; epi.ld_add = 0; epi.bias = v.bias; epi.c = v.c; epi.c_rows = v.c_rows; epi.h = nullptr; epi.ld_state = v.ld;
                epi.hb = v.hb_out; epi.ld_hb = v.ld; epi.raw_out = nullptr; epi.ld_raw = 0; epi.start = nullptr; epi.t = 0;
                const int nch = BN / 32, c0 = (nch * part / EPI_PARTS) * 32, c1 = (nch * (part + 1) / EPI_PARTS) * 32;
                TcEpiCell::State es;
                epi.tile_begin(es, row, row_ok, n0 + c0, (c1 - c0) / 32);
                if (tr && et == 0) tr[16 * l + 4] = clock64();
                mbar_wait(tfull0 + 8 * buf, (used >> 1) & 1);
                if (tr && et == 0) tr[16 * l + 5] = clock64();
                tc_fence_after();
#pragma unroll 1
                for (int c = c0; c < c1; c += 32) {
                    float x[32];
                    tmem_ld32(taddr + c, x);
                    if (n0 + c < N) epi.chunk(es, row, row_ok, n0 + c, N, x, sb_l + c, nullptr);
                }
                epi.tile_end(es, row, row_ok, 0);
            } else if constexpr (kGk != 1) {
                // ---- gk = 2 or 4 K slices of a 128-column tile: exchange partial chunks through L2, finish the own ones
                const int own_per = 4 / gk;               // 32-column chunks of the tile this CTA finishes: [slice*own_per, +own_per)
                // What this warp finishes, in 16-column pieces (4 hidden units): with gk = 4 the single own chunk is shared by the two
                // column parts (one piece each), with gk = 2 a part takes one whole chunk (two pieces).
                auto piece_col = [&](int k) -> int {      // first tile column of piece k (0, 1), -1: none
                    const int col = own_per == 1 ? (k == 0 ? slice * 32 + part * 16 : -1) : (slice * 2 + part) * 32 + k * 16;
                    return (col >= 0 && n0 + col < N) ? col : -1;
                };
                float4 c_old[2];                          // old cell state of those units, fetched while the MMAs run
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    const int col = piece_col(k);
                    c_old[k] = (col >= 0 && row_ok) ? *reinterpret_cast<const float4*>(v.c + ((int64_t)((n0 + col) >> 5) * v.c_rows + row) * 8 + (((n0 + col) >> 2) & 7))
                                                   : make_float4(0.f, 0.f, 0.f, 0.f);
                }
                if (tr && et == 0) tr[16 * l + 4] = clock64();
                mbar_wait(tfull0 + 8 * buf, (used >> 1) & 1);
                if (tr && et == 0) tr[16 * l + 5] = clock64();
                tc_fence_after();
                // foreign chunks -> scratch[pair][chunk][src slice][8 float4 columns][128 rows]
                {
                    int fidx = 0;
                    for (int ch = 0; ch < 4; ++ch) {
                        if (ch / own_per == slice) continue;
                        const bool mine = (gk == 4) ? (part == 1 ? fidx != 0 : fidx == 0) : ((fidx & 1) == part);
                        ++fidx;
                        if (!mine || n0 + ch * 32 >= N) continue;
                        float x[32];
                        tmem_ld32(taddr + ch * 32, x);
                        float4* dst = v.scratch + ((((int64_t)pair * 4 + ch) * 4 + slice) * 8) * 128 + lrow;
#pragma unroll
                        for (int j = 0; j < 8; ++j) dst[j * 128] = make_float4(x[4 * j], x[4 * j + 1], x[4 * j + 2], x[4 * j + 3]);
                    }
                    epi_bar_sync();                           // every epilogue thread's stores are ordered before the release below
                    if (et == 0) {
                        if (tr) tr[16 * l + 6] = clock64();
                        const unsigned target = (unsigned)gk * (token + 1);
                        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(v.part_cnt + pair) : "memory");
                        unsigned c = 0;
                        for (unsigned spin = 0;; ++spin) {
                            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(c) : "l"(v.part_cnt + pair) : "memory");
                            if ((int)(c - target) >= 0) break;
                            if (spin > (1u << 26)) __trap();
                        }
                        if (tr) tr[16 * l + 7] = clock64();
                    }
                    epi_bar_sync();
                }
                // own pieces: all partial loads first, then the sum in slice order, then bias + gates + cell update
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    const int col = piece_col(k);
                    if (col < 0) continue;
                    const int ch = col >> 5, j0 = (col & 31) >> 2;       // chunk, first float4 column inside it
                    float4 t[4][4];
#pragma unroll
                    for (int sl = 0; sl < 4; ++sl) {
                        const float4* src = v.scratch + ((((int64_t)pair * 4 + ch) * 4 + sl) * 8 + j0) * 128 + lrow;
#pragma unroll
                        for (int j = 0; j < 4; ++j) t[sl][j] = (sl < gk && sl != slice) ? __ldcg(src + j * 128) : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                    float mine_v[16], x[16];
                    tmem_ld16(taddr + col, mine_v);
#pragma unroll
                    for (int i = 0; i < 16; ++i) x[i] = 0.f;            // 0 + x is exact: the sum is p0 + p1 + ... in slice order
#pragma unroll
                    for (int sl = 0; sl < 4; ++sl) {
                        if (sl >= gk) break;
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            x[4 * j] += sl == slice ? mine_v[4 * j] : t[sl][j].x;
                            x[4 * j + 1] += sl == slice ? mine_v[4 * j + 1] : t[sl][j].y;
                            x[4 * j + 2] += sl == slice ? mine_v[4 * j + 2] : t[sl][j].z;
                            x[4 * j + 3] += sl == slice ? mine_v[4 * j + 3] : t[sl][j].w;
                        }
                    }
                    if (!row_ok) continue;
                    // four hidden units: columns 4u..4u+3 are their i, f, g, o pre-activations
                    const int n = n0 + col, unit0 = n >> 2;
                    const float* sb = sb_l + col;
                    float cv[4] = {c_old[k].x, c_old[k].y, c_old[k].z, c_old[k].w}, hv[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const float ig = fast_sigmoid(x[4 * u] + sb[4 * u]), fg = fast_sigmoid(x[4 * u + 1] + sb[4 * u + 1]);
                        const float gg = fast_tanh(x[4 * u + 2] + sb[4 * u + 2]), og = fast_sigmoid(x[4 * u + 3] + sb[4 * u + 3]);
                        if (n + 4 * u >= N) { cv[u] = 0.f; hv[u] = 0.f; }  // zero K-padding units stay 0
                        else { cv[u] = fg * cv[u] + ig * gg; hv[u] = og * fast_tanh(cv[u]); }
                    }
                    *reinterpret_cast<float4*>(v.c + ((int64_t)(unit0 >> 3) * v.c_rows + row) * 8 + (unit0 & 7)) = make_float4(cv[0], cv[1], cv[2], cv[3]);
                    __nv_bfloat162 p0 = __floats2bfloat162_rn(hv[0], hv[1]), p1 = __floats2bfloat162_rn(hv[2], hv[3]);
                    *reinterpret_cast<uint2*>(v.hb_out + (int64_t)row * v.ld + unit0) = make_uint2(*reinterpret_cast<uint32_t*>(&p0), *reinterpret_cast<uint32_t*>(&p1));
                }
            }
            // the accumulator buffer is drained: the MMA issuer may reuse it two layers on
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty0 + 8 * buf);
            if (tr && et == 0) tr[16 * l + 8] = clock64();
            if (gtr && et == 0) gtr[1 + l] = (long long)globaltimer_ns();
            // h of this CTA's pieces is written: tell the next layer's producers (generic-proxy stores -> their TMA loads)
            asm volatile("fence.proxy.async.global;" ::: "memory");
            epi_bar_sync();
            if (et == 0) asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(v.done_cnt + pair) : "memory");
            ++used;
        }
    }
    tc_fence_before();
}

__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// Two instantiations, one per LSTM schedule (kSb: the barrier-free phase of sb_lstm_layers, <= 896 streams; else a phase per layer):
// each carries only its own code.  (With both schedules in one kernel, adding code to either slowed the other's phases by 1-3 %.)
// kTrace: the AWD_MEGA_TRACE stamps; the production instantiations carry none of that code either.
template <int kSched, bool kTrace>
__global__ void __launch_bounds__(NUM_THREADS, 1) mega_generate_kernel(const __grid_constant__ MegaParams pp) {
    constexpr bool kSb = kSched != 0;       // kSched: 0 a phase per layer; 1 / 2 / 3 the barrier-free phase with kGk = 0 / 1 / 2 (sb_lstm_layers)
    MegaParams const& p = pp;
    long long* const ptrace = kTrace ? pp.trace : nullptr;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* gen_base = smem_raw + (base - smem_u32(smem_raw));
    const uint32_t tmem_slot = base + 4 * 49152 + 200;
    uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(gen_base + 4 * 49152 + 200);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 1) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;
    unsigned epoch = 0;
    bool first = true;
    const int cta = blockIdx.x, ncta = gridDim.x;
    const int tiles_m = (p.bs + BM - 1) / BM;
    unsigned long long t_prev = 0;
    const bool timing = p.phase_ns != nullptr && cta == 0 && threadIdx.x == 0;
    if (timing) t_prev = globaltimer_ns();
    auto stamp = [&](int slot) {            // straight to global memory: no per-thread accumulators kept live across the phases
        if (timing) {
            unsigned long long t = globaltimer_ns();
            atomicAdd(p.phase_ns + slot, t - t_prev);
            t_prev = t;
        }
    };

    if (threadIdx.x == 0) {                 // descriptors of every phase, fetched once while TMEM is being set up
        prefetch_tmap(&p.tm_x); prefetch_tmap(&p.tm_dec); prefetch_tmap(&p.tm_dec_r); prefetch_tmap(&p.tm_dec_x);
        for (int l = 0; l < p.n_layers; ++l) { prefetch_tmap(&p.tm_w[l]); prefetch_tmap(&p.tm_hb[l][0]); prefetch_tmap(&p.tm_hb[l][1]); }
    }
    for (int i = 0; i < p.n_steps; ++i) {
        const int par = p.p0 ^ (i & 1);
        const int L = p.n_layers;
        // ---- decoder GEMM + sampling race on the last layer's output (reference main.py:61)
        {
            GemmShape g{};
            g.M = p.bs; g.N = p.V; g.kb0 = p.Epad / BK; g.kb1 = 0; g.tiles_m = tiles_m; g.tiles_n = (p.V + 255) / 256;
            g.row_stat = 1; g.col_w = p.dec_col_w; g.x_n0 = p.dec_x_n0; g.x_w = p.dec_x_w;
            TcEpiSampleT<true> epi{};
            epi.sp.inv_temperature = p.sp.inv_temperature; epi.sp.greedy = p.sp.greedy; epi.sp.n_draws = p.sp.n_draws;
            epi.sp.n_banned = p.sp.n_banned; epi.sp.seed_lo = p.sp.seed_lo; epi.sp.seed_hi = p.sp.seed_hi;
            epi.sp.step_lo = (uint32_t)i; epi.sp.step_hi = 0;
            epi.banned_list = p.sp.banned;
            epi.ban_lo = p.ban_lo; epi.ban_hi = p.ban_hi;
            epi.bias = p.dec_bias; epi.part = p.part; epi.zwin = p.zwin; epi.banned_keys = p.banned_keys;
            epi.slots = p.slots; epi.tiles_m = tiles_m; epi.M = p.bs;
            mega_gemm_phase<256, TcEpiSampleT<true>>(gen_base, base, tmem_base, first, &p.tm_hb[L - 1][par], &p.tm_hb[L - 1][par],
                                              &p.tm_dec, g, epi, cta, p.dec_ctas, p.dec_policy, 0,
                                              (ptrace && i == 5 && (cta == 0 || cta == 77)) ? ptrace + 48 + (cta ? 64 : 0) : nullptr, cta < p.dec_ctas ? &p.tm_dec_r : &p.tm_dec_x,
                                              (ptrace && i == 5 && cta == 8) ? ptrace + kTraceTiles : nullptr,
                                              kSb ? p.sb_counter + (MEGA_MAXL + L - 1) * kSbCnt : nullptr, kSb ? p.sb_tiles[L - 1] * tiles_m : 0,
                                              (unsigned)(p.sb_gk[L - 1] * i));
            first = false;
        }
        if (ptrace && i == 5 && cta == 0 && threadIdx.x == 0) ptrace[48 + 7] = clock64();
        if constexpr (kSb) grid_arrive(p.sb_counter + 2 * MEGA_MAXL * kSbCnt + 32);
        else grid_sync(p.sync_counter, epoch);
        if (ptrace && i == 5 && cta == 0 && threadIdx.x == 0) ptrace[48 + 8] = clock64();
        stamp(0);
        // ---- first allowed candidate, emit id, gather its embedding row (main.py:63-71)
        FinalizeArgs fa{};
        fa.part = p.part; fa.zwin = p.zwin; fa.banned_keys = p.banned_keys; fa.tiles_n = p.slots; fa.bs = p.bs; fa.V = p.V;
        fa.out = p.gen + (int64_t)i * p.bs; fa.table = p.table; fa.row_vec4 = p.Epad / 8; fa.xnext = p.xb;
        fa.forced = (p.forced != nullptr && i < p.forced_F) ? p.forced : nullptr; fa.forced_len = p.forced_len; fa.step = i;
        if constexpr (!kSb) {   // (kSb: done inside the LSTM phase, see sb_lstm_layers)
            for (int row = cta + ncta * warp; row < p.bs; row += ncta * (NUM_THREADS / 32)) finalize_row(fa, p.sp, (uint32_t)i, row, lane);
            grid_sync(p.sync_counter, epoch);
            stamp(1);
        }
        // ---- one LSTM step per layer: [x ; h] * [W_ih | W_hh]^T with the cell update in the epilogue
        if constexpr (kSb) {
            // all layers in one phase, 2-D split, layers linked by per-tile counters (sb_lstm_layers)
            auto view = [&](int l) {
                SbLayerView v;
                v.tmA0 = (l == 0) ? &p.tm_x : &p.tm_hb[l - 1][par ^ 1];
                v.tmA1 = &p.tm_hb[l][par];
                v.tmB = &p.tm_w[l];
                v.kb0 = p.Ipad[l] / BK; v.kb1 = p.Hpad[l] / BK; v.N = 4 * p.H[l];
                v.n_tiles = p.sb_tiles[l]; v.gk = p.sb_gk[l]; v.bn = p.bn[l]; v.cta0 = p.sb_cta0[l];
                v.bias = p.bias[l]; v.c = p.c[l]; v.c_rows = p.bs; v.hb_out = p.hb[l][par ^ 1]; v.ld = p.Hpad[l];
                v.scratch = p.sb_scratch + p.sb_scratch_off[l];
                v.part_cnt = p.sb_counter + l * kSbCnt;
                v.done_cnt = p.sb_counter + (MEGA_MAXL + l) * kSbCnt;
                // input half: layer l-1's tiles, or (layer 0) the bs rows gathered by the finalize step of this phase
                v.wait_cnt = l > 0 ? p.sb_counter + (MEGA_MAXL + l - 1) * kSbCnt : p.sb_counter + 2 * MEGA_MAXL * kSbCnt;
                v.wait_tiles = l > 0 ? p.sb_tiles[l - 1] : 0;
                v.wait_per_token = l > 0 ? p.sb_gk[l - 1] : p.bs;
                v.wait_tpk = l > 0 ? 256 / p.bn[l - 1] : 1;
                return v;
            };
            long long* tr = (ptrace && i == 5 && (cta == 0 || cta == 77)) ? ptrace + (cta ? 64 : 0) : nullptr;
            sb_lstm_layers<kSched == 0 ? 0 : kSched - 1>(gen_base, base, tmem_base, L, p.bs, tiles_m, (unsigned)i, view, cta, p.lstm_policy, &fa, &p.sp, p.sb_counter + 2 * MEGA_MAXL * kSbCnt + 32, (unsigned)(i + 1) * gridDim.x, p.sb_counter + 2 * MEGA_MAXL * kSbCnt,
                           (i + 1 < p.n_steps && cta < p.dec_ctas) ? &p.tm_dec : nullptr, p.Epad / BK, (p.V + 255) / 256, p.dec_ctas / tiles_m, p.dec_col_w, p.sb_prefetch, p.sb_stages, tr,
                           (ptrace && i == 5) ? ptrace + 128 + cta * 8 : nullptr);
            // (no grid barrier: the next decoder phase waits for the last layer's `done` counters before its first activation load)
            if (tr && threadIdx.x == 0) tr[9] = tr[10] = clock64();
            stamp(2);
        } else {
            for (int l = 0; l < L; ++l) {
                GemmShape g{};
                g.M = p.bs; g.N = 4 * p.H[l]; g.kb0 = p.Ipad[l] / BK; g.kb1 = p.Hpad[l] / BK; g.tiles_m = tiles_m;
                g.tiles_n = (g.N + p.bn[l] - 1) / p.bn[l];
                TcEpiCell epi{};
                epi.addend = nullptr; epi.ld_add = 0; epi.bias = p.bias[l]; epi.c = p.c[l]; epi.c_rows = p.bs; epi.h = nullptr; epi.ld_state = p.Hpad[l];
                epi.hb = p.hb[l][par ^ 1]; epi.ld_hb = p.Hpad[l]; epi.raw_out = nullptr; epi.ld_raw = 0; epi.start = nullptr; epi.t = 0;
                const CUtensorMap* a0 = (l == 0) ? &p.tm_x : &p.tm_hb[l - 1][par ^ 1];
                // tile width chosen on the host so that the layer's tiles cover (nearly) all SMs: 256 at 1024 streams,
                // down to 32 at two M tiles (256 streams per GPU, the 4-GPU split of BASELINE config 3)
                mega_gemm_phase<0, TcEpiCell>(gen_base, base, tmem_base, false, a0, &p.tm_hb[l][par], &p.tm_w[l], g, epi, cta, ncta, p.lstm_policy, p.bn[l]);
                grid_sync(p.sync_counter, epoch);
                stamp(2 + (l < 2 ? l : 2));
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// =============================================================================================
// Persistent prompt kernel: the T x L recurrent steps of one awd_forward time chunk in ONE launch.
//
// As separate launches the recurrence of a prompt is T*L dependent kernels of a few microseconds of
// work each (C2: 70 x 3 launches, ~18 us apiece).  Here the steps run as a WAVEFRONT: wave w executes
// layer l at timestep t = w - l for every l at once -- the three (layer, step) problems of a wave are
// independent, so they go to disjoint CTA ranges and share one grid barrier.  Layer 0 adds its
// pre-computed input projection (one big GEMM over the whole chunk); layers >= 1 contract [h_below ; h]
// directly, because the layer below finished that timestep in the previous wave.
// =============================================================================================
struct FwdParams {
    CUtensorMap tm_y[MEGA_MAXL];            // layer outputs of the chunk, bf16 [T*bs][Hpad]
    CUtensorMap tm_hb[MEGA_MAXL];           // carried hidden operand (state before the chunk), [bs][Hpad]
    CUtensorMap tm_w[MEGA_MAXL];            // [W_ih | W_hh], box rows = bn[l]
    int n_layers, bs, T, t_base;
    int H[MEGA_MAXL], Ipad[MEGA_MAXL], Hpad[MEGA_MAXL], bn[MEGA_MAXL], cta_off[MEGA_MAXL];
    const float* bias[MEGA_MAXL];
    float* c[MEGA_MAXL];
    float* h[MEGA_MAXL];
    __nv_bfloat16* y[MEGA_MAXL];
    float* raw[MEGA_MAXL];                  // optional caller tensors (T, bs, H)
    const float* xproj0;                    // [T*bs][4*H0] fp32, bias included
    const int32_t* start;                   // ragged prompts
    unsigned* sync_counter;
};

__global__ void __launch_bounds__(NUM_THREADS, 1) mega_forward_kernel(const __grid_constant__ FwdParams p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* gen_base = smem_raw + (base - smem_u32(smem_raw));
    const uint32_t tmem_slot = base + 4 * 49152 + 200;
    uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(gen_base + 4 * 49152 + 200);
    const int warp = threadIdx.x >> 5;
    if (warp == 1) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;
    unsigned epoch = 0;
    bool first = true;
    const int ncta = gridDim.x, L = p.n_layers;
    const int tiles_m = (p.bs + BM - 1) / BM;

    for (int w = 0; w < p.T + L - 1; ++w) {
        for (int l = 0; l < L; ++l) {
            const int t = w - l;
            if (t < 0 || t >= p.T) continue;
            GemmShape g{};
            g.M = p.bs; g.N = 4 * p.H[l]; g.tiles_m = tiles_m; g.tiles_n = (g.N + p.bn[l] - 1) / p.bn[l];
            TcEpiCell epi{};
            epi.c = p.c[l]; epi.h = (t == p.T - 1) ? p.h[l] : nullptr; epi.ld_state = p.Hpad[l];
            epi.hb = p.y[l] + (int64_t)t * p.bs * p.Hpad[l]; epi.ld_hb = p.Hpad[l];
            epi.raw_out = p.raw[l] ? p.raw[l] + (int64_t)t * p.bs * p.H[l] : nullptr; epi.ld_raw = p.H[l];
            epi.start = p.start; epi.t = p.t_base + t;
            const CUtensorMap* hprev = (t == 0) ? &p.tm_hb[l] : &p.tm_y[l];
            const int hrow = (t == 0) ? 0 : (t - 1) * p.bs;
            const CUtensorMap *a0, *a1;
            if (l == 0) {       // h * W_hh^T + (x * W_ih^T + b), the bracket pre-computed for the whole chunk
                g.kb0 = p.Hpad[0] / BK; g.kb1 = 0; g.a0_row = hrow; g.b_k0 = p.Ipad[0];
                a0 = a1 = hprev;
                epi.addend = p.xproj0 + (int64_t)t * p.bs * 4 * p.H[0]; epi.ld_add = 4 * p.H[0]; epi.bias = nullptr;
            } else {            // [h_below(t) ; h(t-1)] * [W_ih | W_hh]^T + b
                g.kb0 = p.Ipad[l] / BK; g.kb1 = p.Hpad[l] / BK; g.a0_row = t * p.bs; g.a1_row = hrow; g.b_k0 = 0;
                a0 = &p.tm_y[l - 1]; a1 = hprev;
                epi.addend = nullptr; epi.ld_add = 0; epi.bias = p.bias[l];
            }
            const int cta = (blockIdx.x + ncta - p.cta_off[l] % ncta) % ncta;     // each layer starts on its own CTA range
            if (p.bn[l] == 256)
                mega_gemm_phase<256, TcEpiCell, true>(gen_base, base, tmem_base, first, a0, a1, &p.tm_w[l], g, epi, cta, ncta, 0x1000000000000000ull);
            else
                mega_gemm_phase<128, TcEpiCell, true>(gen_base, base, tmem_base, first, a0, a1, &p.tm_w[l], g, epi, cta, ncta, 0x1000000000000000ull);
            first = false;
            __syncthreads();        // the next sub-phase re-initialises this CTA's barriers
        }
        grid_sync(p.sync_counter, epoch);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

int launch_mega_forward(const FwdArgs& m, cudaStream_t st) {
    static std::atomic<bool> attr_set_dev[kMaxDev];
    std::atomic<bool>& attr_set = attr_set_dev[cur_dev()];
    constexpr int kSmem = 4 * 49152 + 256 + 2048 + 1024;
    if (!attr_set) {
        AWD_CUDA(cudaFuncSetAttribute(mega_forward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
        attr_set = true;
    }
    if (m.n_layers < 1 || m.n_layers > MEGA_MAXL) return set_error(AWD_ELIMIT, "persistent prompt kernel: 1..%d layers", MEGA_MAXL);
    const int tiles_m = (int)cdiv(m.bs, BM);
    FwdParams p{};
    const uint64_t R = (uint64_t)m.T * m.bs;
    int off = 0, rc;
    // Tile width per layer.  The (layer, step) problems of a wave run side by side, one CTA per tile; a tcgen05.mma occupies the
    // tensor pipe just as long at N = 128 as at 256, so when 128-wide tiles of all layers do not fit the SMs in one round the
    // layers with the most tiles go to 256 (C2, 70 x 256: 72 + 72 + 26 tiles -> 36 + 36 + 26).
    {
        int total = 0;
        for (int l = 0; l < m.n_layers; ++l) { p.bn[l] = 128; total += tiles_m * (int)cdiv(4 * m.H[l], 128); }
        while (total > num_sms()) {
            int best = -1, best_tiles = 0;
            for (int l = 0; l < m.n_layers; ++l) {
                const int t = tiles_m * (int)cdiv(4 * m.H[l], 128);
                if (p.bn[l] == 128 && t > best_tiles) { best = l; best_tiles = t; }
            }
            if (best < 0) break;
            p.bn[best] = 256;
            total -= best_tiles - tiles_m * (int)cdiv(4 * m.H[best], 256);
        }
    }
    for (int l = 0; l < m.n_layers; ++l) {
        p.H[l] = m.H[l]; p.Ipad[l] = m.Ipad[l]; p.Hpad[l] = m.Hpad[l];
        const int N = 4 * m.H[l];
        p.cta_off[l] = off % num_sms();
        off += tiles_m * (int)cdiv(N, p.bn[l]);
        if ((rc = make_tmap_bf16(&p.tm_y[l], m.y[l], R, (uint64_t)m.Hpad[l], (uint64_t)m.Hpad[l], BM))) return rc;
        if ((rc = make_tmap_bf16(&p.tm_hb[l], m.hb[l], (uint64_t)m.bs, (uint64_t)m.Hpad[l], (uint64_t)m.Hpad[l], BM))) return rc;
        const uint64_t kcat = (uint64_t)(m.Ipad[l] + m.Hpad[l]);
        if ((rc = make_tmap_bf16(&p.tm_w[l], m.wcat[l], (uint64_t)N, kcat, kcat, p.bn[l]))) return rc;
        p.bias[l] = m.bias[l]; p.c[l] = m.c[l]; p.h[l] = m.h[l]; p.y[l] = (__nv_bfloat16*)m.y[l]; p.raw[l] = m.raw[l];
    }
    p.n_layers = m.n_layers; p.bs = m.bs; p.T = m.T; p.t_base = m.t_base; p.xproj0 = m.xproj0; p.start = m.start;
    p.sync_counter = m.sync_counter;
    AWD_CUDA(cudaMemsetAsync(m.sync_counter, 0, sizeof(unsigned), st));

    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(num_sms());
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = kSmem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    AWD_CUDA(cudaLaunchKernelEx(&cfg, mega_forward_kernel, p));
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

// scratch[layer][tile][chunk 4][src slice 4][8 float4 columns][128 rows]: one region per layer, because no grid barrier
// separates the layers of a token (a CTA may store layer l+1 partials while a peer still reads layer l's)
static size_t sb_tile_float4() { return (size_t)4 * 4 * 8 * 128; }
size_t sb_scratch_bytes(int n_layers, const int* H, int bs) {
    size_t tiles = 0;
    for (int l = 0; l < n_layers; ++l) tiles += (size_t)cdiv(4 * H[l], SB_BN);
    return tiles * (size_t)cdiv(bs, BM) * sb_tile_float4() * sizeof(float4);
}

// engine cell state [bs][Hpad] <-> the persistent kernel's blocked copy [Hpad / 8][bs][8]; one thread per 8 states
__global__ void cell_block_kernel(const float* __restrict__ src, float* __restrict__ dst, int bs, int Hpad, int to_blocked) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;       // i = ub * bs + row
    if (i >= (int64_t)bs * (Hpad / 8)) return;
    const int ub = (int)(i / bs), row = (int)(i - (int64_t)ub * bs);
    const float4* s4 = reinterpret_cast<const float4*>(to_blocked ? src + (int64_t)row * Hpad + ub * 8 : src + i * 8);
    float4* d4 = reinterpret_cast<float4*>(to_blocked ? dst + i * 8 : dst + (int64_t)row * Hpad + ub * 8);
    d4[0] = s4[0]; d4[1] = s4[1];
}

int launch_mega_generate(const MegaArgs& m, cudaStream_t st) {
    static std::atomic<bool> attr_set_dev[kMaxDev];
    std::atomic<bool>& attr_set = attr_set_dev[cur_dev()];
    constexpr int kSmem = 4 * 49152 + 256 + 2048 + 1024;
    if (!attr_set) {
        AWD_CUDA(cudaFuncSetAttribute(mega_generate_kernel<0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
        AWD_CUDA(cudaFuncSetAttribute(mega_generate_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
        AWD_CUDA(cudaFuncSetAttribute(mega_generate_kernel<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
        AWD_CUDA(cudaFuncSetAttribute(mega_generate_kernel<3, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
        AWD_CUDA(cudaFuncSetAttribute(mega_generate_kernel<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
        AWD_CUDA(cudaFuncSetAttribute(mega_generate_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
        attr_set = true;
    }
    if (m.n_layers < 1 || m.n_layers > MEGA_MAXL) return set_error(AWD_ELIMIT, "persistent generation kernel: 1..%d layers", MEGA_MAXL);
    const int tiles_m = (int)cdiv(m.bs, BM);
    if (tiles_m > num_sms()) return set_error(AWD_ELIMIT, "persistent generation kernel: at most %d streams", num_sms() * BM);
    MegaParams p{};
    int rc = make_tmap_bf16(&p.tm_x, m.xb, (uint64_t)m.bs, (uint64_t)m.Epad, (uint64_t)m.Epad, BM);
    if (rc) return rc;
    // one M tile: the 2-D split of the LSTM layers, if the caller provided its scratch (AWD_SB=0 switches it off)
    static int sb_env = -1;
    if (sb_env < 0) { const char* sv = getenv("AWD_SB"); sb_env = sv ? atoi(sv) : 1; }
    // The barrier-free LSTM phase (sb_lstm_layers) needs one (N tile, M tile, K slice) per CTA and layer: tiles of 128 gate columns
    // with 4 / 2 K slices at one / two M tiles, one K slice above (256-column tiles where 128-column ones do not fit the SMs).  Used
    // up to eight M tiles (AWD_SB_MAX_MTILES caps it): with the per-schedule kernel instantiations it measured, per token, 1024
    // streams 110.1 -> 103-105 us against the phase-per-layer schedule (round-2 start, both schedules in one kernel: no difference).
    static int sb_max_tm = -1;
    if (sb_max_tm < 0) { const char* sv = getenv("AWD_SB_MAX_MTILES"); sb_max_tm = sv ? atoi(sv) : 8; }
    bool sb = sb_env && tiles_m <= sb_max_tm && tiles_m <= 8 && m.sb_scratch != nullptr && m.sb_counter != nullptr;
    // tile width per layer: 128 when the layer's (N tile, M tile) pairs fit the SMs, else 256
    int sb_bn[MEGA_MAXL], sb_cta0[MEGA_MAXL] = {0, 0, 0, 0};
    for (int l = 0; l < m.n_layers; ++l) sb_bn[l] = cdiv(4 * m.H[l], SB_BN) * tiles_m <= num_sms() ? SB_BN : 256;
    // Two to four M tiles: NEIGHBOURING LAYERS ON DISJOINT CTAs, one K slice each.  When every CTA walks all layers in turn, the
    // recurrent half of layer l+1 (half of its K blocks, independent of layer l) only starts once the CTA is through with layer l;
    // at 512 streams layer 1's input half started 10 us after layer 0 was complete, and with 2 x 144 tiles of 128 columns in flight
    // the phase was bound by the L2 -> SM traffic of the narrow tiles (~800 clocks per 32 KB K block on all SMs).  Placing layer l+1
    // on other CTAs lets its recurrent half run while layer l is on the critical path, and needs no exchange of partial tiles:
    // 256 streams: 72 + 72 CTAs with 128-wide tiles; 384 / 512 streams: 256-wide tiles (54 + 54 + 39 resp. 72 + 72, layer 2 back on
    // layer 0's CTAs).  AWD_SB_LP=0 switches it off.
    static int lp_env = -1;
    if (lp_env < 0) { const char* sv = getenv("AWD_SB_LP"); lp_env = sv ? atoi(sv) : 1; }
    bool lp = sb && lp_env && tiles_m >= 3 && tiles_m <= 4 && m.n_layers >= 2;
    if (lp) {
        int bn[MEGA_MAXL];
        for (int l = 0; l < m.n_layers; ++l) bn[l] = SB_BN;
        auto cnt = [&](int l) { return (int)cdiv(4 * m.H[l], bn[l]) * tiles_m; };
        for (int l = 0; l + 1 < m.n_layers && lp; ++l)
            while (cnt(l) + cnt(l + 1) > num_sms()) {
                const int wl = (bn[l] < 256 && (bn[l + 1] == 256 || cnt(l) >= cnt(l + 1))) ? l : l + 1;
                if (bn[wl] == 256) { lp = false; break; }
                bn[wl] = 256;
            }
        int off[MEGA_MAXL] = {0, 0, 0, 0};
        for (int l = 1; l < m.n_layers && lp; ++l) {
            if (off[l - 1] + cnt(l - 1) + cnt(l) <= num_sms()) off[l] = off[l - 1] + cnt(l - 1);
            else if (cnt(l) <= off[l - 1]) off[l] = 0;
            else lp = false;
        }
        for (int l = 0; l < m.n_layers && lp; ++l) { sb_bn[l] = bn[l]; sb_cta0[l] = off[l]; }
    }
    bool sb_wide = false;
    for (int l = 0; l < m.n_layers && sb; ++l) {
        const int bn = sb_bn[l];
        sb_wide = sb_wide || bn == 256;
        if (cdiv(4 * m.H[l], bn) * tiles_m > kSbCnt || cdiv(4 * m.H[l], bn) * tiles_m > num_sms()) sb = false;
    }
    p.sb = sb ? 1 : 0;
    p.sb_stages = sb_wide ? 4 : SB_STAGES;
    { static int pf = -1; if (pf < 0) { const char* sv = getenv("AWD_SB_PREFETCH"); pf = sv ? atoi(sv) : 4; } p.sb_prefetch = pf; }
    p.sb_scratch = (float4*)m.sb_scratch; p.sb_counter = m.sb_counter; p.trace = m.trace;
    if (sb) AWD_CUDA(cudaMemsetAsync(m.sb_counter, 0, kSbCounterBytes, st));
    for (int l = 0; l < m.n_layers; ++l) {
        p.H[l] = m.H[l]; p.Ipad[l] = m.Ipad[l]; p.Hpad[l] = m.Hpad[l];
        const int N = 4 * m.H[l];
        // Tile width per layer.  >= 3 M tiles (384+ streams): 256 wide when that still fills the SMs, else 128.
        // 1-2 M tiles (the 128- and 256-stream shards of an 8- or 4-GPU split of 1024 streams): the narrowest
        // width (>= 32) whose tiles still fit one wave -- measured 8-14 % faster per phase there, and slower
        // than 128 from 4 M tiles up (profiles/r01_bf16_summary.md).
        {
            static int small = -1;      // AWD_MEGA_SMALL_BN=0: never below 128
            if (small < 0) { const char* sv = getenv("AWD_MEGA_SMALL_BN"); small = sv ? atoi(sv) : 1; }
            const int tm = (int)cdiv(m.bs, BM);
            p.bn[l] = (tm * cdiv(N, 256) >= num_sms() - 8) ? 256 : 128;
            if (small && tm <= 2)
                for (int bn : {32, 64, 128})
                    if (tm * cdiv(N, bn) <= num_sms()) { p.bn[l] = bn; break; }
        }
        if (sb) {        // 2-D split (sb_lstm_phase): tiles of 128 gate columns x the most K slices (4, 2 or 1) that fit the SMs
            p.bn[l] = sb_bn[l];
            p.sb_cta0[l] = sb_cta0[l];
            p.sb_tiles[l] = (int)cdiv(N, p.bn[l]);
            p.sb_scratch_off[l] = l ? p.sb_scratch_off[l - 1] + (int64_t)p.sb_tiles[l - 1] * tiles_m * (int64_t)sb_tile_float4() : 0;
            p.sb_gk[l] = 1;
            static int gk_max = -1;
            if (gk_max < 0) { const char* sv = getenv("AWD_SB_GK_MAX"); gk_max = sv ? atoi(sv) : 4; }
            for (int gk : {4, 2})
                if (gk <= gk_max && !lp && tiles_m <= 4 && p.bn[l] == SB_BN && p.sb_tiles[l] * gk * tiles_m <= num_sms() && (m.Ipad[l] + m.Hpad[l]) / BK >= gk) { p.sb_gk[l] = gk; break; }
        }
        for (int k = 0; k < 2; ++k) {
            if ((rc = make_tmap_bf16(&p.tm_hb[l][k], m.hb[l][k], (uint64_t)m.bs, (uint64_t)m.Hpad[l], (uint64_t)m.Hpad[l], BM))) return rc;
            p.hb[l][k] = (__nv_bfloat16*)m.hb[l][k];
        }
        const uint64_t kcat = (uint64_t)(m.Ipad[l] + m.Hpad[l]);
        if ((rc = make_tmap_bf16(&p.tm_w[l], m.wcat[l], (uint64_t)N, kcat, kcat, p.bn[l]))) return rc;
        p.bias[l] = m.bias[l]; p.c[l] = m.c_blk[l]; p.h[l] = m.h[l];
        {
            const int64_t n8 = (int64_t)m.bs * (m.Hpad[l] / 8);
            cell_block_kernel<<<(unsigned)cdiv(n8, 256), 256, 0, st>>>(m.c[l], m.c_blk[l], m.bs, m.Hpad[l], 1);
            AWD_LAUNCH_CHECK();
        }
    }
    if ((rc = make_tmap_bf16(&p.tm_dec, m.dec_w, (uint64_t)m.V, (uint64_t)m.Epad, (uint64_t)m.Epad, 256))) return rc;
    p.n_layers = m.n_layers; p.bs = m.bs; p.V = m.V; p.n_steps = m.n_steps; p.p0 = m.p0; p.Epad = m.Epad;
    p.ban_lo = m.sp.n_banned ? m.sp.banned[0] : 0x7fffffff; p.ban_hi = m.sp.n_banned ? m.sp.banned[m.sp.n_banned - 1] : -1;
    p.dec_bias = m.dec_bias; p.sp = m.sp; p.part = (float2*)m.part; p.zwin = m.zwin; p.banned_keys = m.banned_keys;
    const int groups = std::min((int)cdiv(m.V, 256), std::max(1, num_sms() / tiles_m));
    p.slots = EPI_PARTS * groups; p.dec_ctas = groups * tiles_m;
    {
        // Columns per group (see PhaseTiles): ceil(V / groups) rounded up to the 32-column chunks of the sampling race; 60 000 columns
        // on 148 / 74 / 37 / 18 groups = 416 / 832 / 1632 / 3360, i.e. 1.63 / 3.25 / 6.38 / 13.1 tiles for every group instead of
        // 2 / 4 / 7 / 14 for the slowest.  AWD_DEC_BALANCE=0: whole 256-column tiles dealt round-robin.
        static int bal = -1;
        if (bal < 0) { const char* sv = getenv("AWD_DEC_BALANCE"); bal = sv ? atoi(sv) : 1; }
        p.dec_col_w = bal ? (int)(cdiv(cdiv(m.V, groups), 32) * 32) : 0;
        // CTAs left over after the last whole group (4 at eight M tiles, 1 at seven): they take the tail of the vocabulary for every
        // M tile in `passes` turns, so the groups' ranges shrink to V / (groups + 1 / passes) columns -- 3264 instead of 3360 at 1024
        // streams, 13 tiles per group instead of 14.  Their partials go to one more pair of slots.
        const int left = num_sms() - groups * tiles_m;
        if (bal && left > 0 && groups * 256 < m.V) {
            const int passes = (int)cdiv(tiles_m, left);
            const int w = (int)(cdiv((int64_t)m.V * passes, (int64_t)groups * passes + 1) + 31) / 32 * 32;
            const int x_w = (int)(cdiv(m.V - (int64_t)groups * w, 32) * 32);
            // Only where it saves the groups a tile, and the left-over CTAs finish no later than the groups: the first tile of every
            // pass costs about two (no row bound yet, every chunk is evaluated in full).  Measured per token: 1024 streams 122.6 ->
            // 121.3 us; at 768 and 896 streams, where the tile count stays, the extra slots and passes cost 1 %.
            if (x_w > 0 && cdiv(w, 256) < cdiv(p.dec_col_w, 256) && (int64_t)passes * (cdiv(x_w, 256) + 1) <= cdiv(w, 256) + 1) {
                p.dec_col_w = w; p.dec_x_n0 = groups * w; p.dec_x_w = x_w;
                p.slots = EPI_PARTS * (groups + 1);
            }
        }
        const int rem = p.dec_col_w % 256, rem_x = p.dec_x_w % 256;
        if ((rc = make_tmap_bf16(&p.tm_dec_r, m.dec_w, (uint64_t)m.V, (uint64_t)m.Epad, (uint64_t)m.Epad, rem ? rem : 256))) return rc;
        if ((rc = make_tmap_bf16(&p.tm_dec_x, m.dec_w, (uint64_t)m.V, (uint64_t)m.Epad, (uint64_t)m.Epad, rem_x ? rem_x : 256))) return rc;
    }
    p.gen = m.gen; p.table = (const uint4*)m.table; p.xb = (uint4*)m.xb;
    p.forced = m.forced; p.forced_len = m.forced_len; p.forced_F = m.forced_F;
    p.sync_counter = m.sync_counter; p.phase_ns = m.phase_ns;
    {
        // L2 residency.  The 40 MB of LSTM weights are asked to stay (evict_last), the 54 MB decoder matrix streams through
        // (evict_first).  With few M tiles the LSTM phases are latency chains (a few MMAs per CTA) and what they wait for is the first
        // weight tile -- ~2200 clocks from HBM, ~1000 from L2 (profiles/micro/ramp_probe.cu); at 1024 streams the same choice measured
        // 123.2 vs 126.6 us per token against round 1's opposite one.  AWD_MEGA_L2: 0 no hints, 1 decoder resident / LSTM streaming,
        // 2 this default, 3 both evict_last.
        static int pol = -1;
        if (pol < 0) { const char* sv = getenv("AWD_MEGA_L2"); pol = sv ? atoi(sv) : 2; }
        const uint64_t normal = 0x1000000000000000ull;
        p.dec_policy = pol == 0 ? normal : (pol == 2 ? L2_EVICT_FIRST : L2_EVICT_LAST);
        p.lstm_policy = pol == 0 ? normal : (pol == 1 ? L2_EVICT_FIRST : L2_EVICT_LAST);
    }
    AWD_CUDA(cudaMemsetAsync(m.sync_counter, 0, sizeof(unsigned), st));
    AWD_CUDA(cudaMemsetAsync(m.banned_keys + (size_t)m.bs * 64, 0x80, (size_t)m.bs * kRowBoundStride * sizeof(int), st));   // row bounds: "no key yet"
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(num_sms());
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = kSmem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    // instantiation: schedule (and, inside the barrier-free phase, whether any / every layer exchanges partial tiles), tracing
    int sched = 0;
    if (p.sb) {
        bool all_one = true, all_split = true;
        for (int l = 0; l < m.n_layers; ++l) { all_one = all_one && p.sb_gk[l] == 1; all_split = all_split && p.sb_gk[l] > 1; }
        sched = all_one ? 2 : (all_split ? 3 : 1);
    }
    if (p.trace != nullptr) {
        if (sched) AWD_CUDA(cudaLaunchKernelEx(&cfg, mega_generate_kernel<1, true>, p));
        else AWD_CUDA(cudaLaunchKernelEx(&cfg, mega_generate_kernel<0, true>, p));
    } else if (sched == 0) AWD_CUDA(cudaLaunchKernelEx(&cfg, mega_generate_kernel<0, false>, p));
    else if (sched == 1) AWD_CUDA(cudaLaunchKernelEx(&cfg, mega_generate_kernel<1, false>, p));
    else if (sched == 2) AWD_CUDA(cudaLaunchKernelEx(&cfg, mega_generate_kernel<2, false>, p));
    else AWD_CUDA(cudaLaunchKernelEx(&cfg, mega_generate_kernel<3, false>, p));
    AWD_LAUNCH_CHECK();
    for (int l = 0; l < m.n_layers; ++l) {
        const int64_t n8 = (int64_t)m.bs * (m.Hpad[l] / 8);
        cell_block_kernel<<<(unsigned)cdiv(n8, 256), 256, 0, st>>>(m.c_blk[l], m.c[l], m.bs, m.Hpad[l], 0);
    }
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

}  // namespace tc
}  // namespace awd
