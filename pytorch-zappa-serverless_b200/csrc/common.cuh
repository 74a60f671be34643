// Shared declarations of the AWD-LSTM engine (internal; the public surface is include/awd_b200.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <atomic>

#include "awd_b200.h"

namespace awd {

int set_error(int code, const char* fmt, ...);
extern std::atomic<int64_t> g_launches;
inline void count_launch(int n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }

#define AWD_CUDA(x)                                                                            \
    do {                                                                                       \
        cudaError_t _e = (x);                                                                  \
        if (_e != cudaSuccess)                                                                 \
            return ::awd::set_error(_e == cudaErrorMemoryAllocation ? AWD_ENOMEM : AWD_ECUDA,  \
                                    "%s: %s (%s:%d)", #x, cudaGetErrorString(_e), __FILE__, __LINE__); \
    } while (0)

#define AWD_LAUNCH_CHECK()                                                                     \
    do {                                                                                       \
        ::awd::count_launch();                                                                 \
        cudaError_t _e = cudaGetLastError();                                                   \
        if (_e != cudaSuccess)                                                                 \
            return ::awd::set_error(AWD_ECUDA, "kernel launch: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, __LINE__); \
    } while (0)

// Uniform in the OPEN interval (0,1) from random bits, for -log(-log(u)) Gumbel noise: 23 bits + 1/2 ulp, so
// 2^-24 <= u <= 1 - 2^-24 exactly.  (24 bits + 0.5 rounds to 1.0f for the all-ones word, and a key of +inf.)
__host__ __device__ inline float uniform_open01(uint32_t bits) { return ((float)(bits >> 9) + 0.5f) * (1.0f / 8388608.0f); }

#ifdef __CUDACC__
// Gumbel noise -log(-log(u)) with the fast (MUFU) logarithms of the bf16 samplers.  lg2.approx has an ABSOLUTE error
// of ~2^-22, so -log(u) is worthless (even of the wrong sign -> NaN key) when u is within ~1e-6 of 1; there
// v = 1 - u is taken exactly from the complemented bits and -log(1 - v) = v + v^2/2 + O(v^3).
__device__ __forceinline__ float gumbel_from_bits(uint32_t bits) {
    const float u = uniform_open01(bits), v = uniform_open01(~bits);
    const float e = v < (1.0f / 1024.0f) ? fmaf(0.5f * v, v, v) : -__logf(u);
    return -__logf(e);
}
#endif

static inline int round_up(int x, int m) { return (x + m - 1) / m * m; }
static inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }

// ---------------------------------------------------------------------------------------------
// fp32 path (kernels_f32.cu).  All K extents and leading dimensions are multiples of 4 floats so
// that every row starts 16-byte aligned and can be read with 128-bit loads.
// ---------------------------------------------------------------------------------------------

// One A*W^T term.  A is [M][lda] (K-contiguous) or, when idx != nullptr, a row table that is
// gathered on the fly: row m of A is table row idx[m] (the embedding gather fused into the
// first layer's projection).  W is [N][ldw], K-contiguous (torch's weight layout).
struct Operand {
    const float* A;
    const int64_t* idx;
    const float* W;
    int lda, ldw, K;
    int idx_limit;        // valid ids are [0, idx_limit); others raise *err_flag and read row 0
    int* err_flag;
};

// C[m][n] = acc + bias[n]
struct EpiStore {
    float* C;
    int64_t ldc;
    const float* bias;    // may be nullptr
};

// LSTM cell.  Weight rows are packed gate-interleaved: column n = 4*j + g, g in (i,f,g,o), so a
// float4 of accumulators is the four gates of hidden unit j (nn.LSTM semantics at
// pytorch_models/awd_lstm.py:162).
struct EpiCell {
    const float* addend;  // [M][ld_add] pre-computed x*W_ih^T + b (same column order) or nullptr
    int64_t ld_add;
    const float* bias;    // [4H] interleaved, b_ih + b_hh; used only when addend == nullptr
    float* c;             // [M][ld_state] cell state, updated in place
    float* h_out;         // [M][ld_state] new hidden state
    int ld_state;
    float* raw_out;       // optional caller tensor [M][ld_raw] (ld_raw = H, unpadded)
    int64_t ld_raw;
    int H;
    // ragged prompts: stream m is fed from timestep start[m] on; before that its state stays at reset
    const int32_t* start; // [M] or nullptr
    int t;                // timestep of this launch
};

struct SampleParams {
    float inv_temperature;
    int greedy, n_draws, n_banned;   // banned[] sorted ascending, includes 0 when ban_zero
    int banned[64];
    uint32_t seed_lo, seed_hi;
    uint32_t step_lo, step_hi;
};

// Decoder fused with the sampling race (fp32 mode, more than 8 streams): every 128- or 64-column tile of the decoder GEMM reduces
// its logits to ONE candidate pair per stream -- (best key over all ids, best key over allowed ids), the same keys sample_f32_kernel
// computes from stored logits -- so the (bs, vocab_sz) logits never reach HBM (reference main.py:61-68 needs only the winner).
struct EpiRace {
    const float* bias;        // decoder bias [N] or nullptr
    float4* part;             // [M][tiles_n]: (key_all, id_all, key_allowed, id_allowed), ids as int bits
    float* banned_keys;       // [M][64]: race key of each disallowed id (for the "else the first candidate" rule)
    int tiles_n;
    SampleParams sp;
};


int launch_gather_f32(const float* table, int ld, int E, int V, const int64_t* tok, int64_t n,
                      float* out, int64_t ldo, int* err_flag, cudaStream_t st);
int launch_gemm_store_f32(const Operand& a, int M, int N, const EpiStore& epi, cudaStream_t st);
int launch_gemm_cell_f32(const Operand& a, const Operand* b, int M, int N, const EpiCell& epi, cudaStream_t st);
int launch_sample_f32(const float* logits, int64_t ld, int V, int bs, const SampleParams& sp,
                      int64_t* out_a, int64_t* out_b, cudaStream_t st);
// fused decoder + race for M > 8 streams (`part` holds M * race_tiles_f32(M, N) float4), then the merge into out[M]
int race_tiles_f32(int M, int N);
int launch_gemm_race_f32(const Operand& a, int M, int N, const float* bias, const SampleParams& sp, float4* part, float* banned_keys,
                         int64_t* out, cudaStream_t st);
int launch_pack_rows_f32(const float* src, int rows, int cols, float* dst, int ld_dst, int H_interleave,
                         cudaStream_t st);
int launch_bias_sum_f32(const float* b_ih, const float* b_hh, int H, float* dst, cudaStream_t st);

}  // namespace awd
