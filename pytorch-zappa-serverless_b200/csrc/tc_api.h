// Launchers of the tcgen05 bf16 path (gemm_tc.cu), callable from the engine without pulling in
// the kernel header.
#pragma once
#include "common.cuh"

namespace awd {
namespace tc {

// D = [A0 | A1] * W[:, b_k0 : b_k0+k0+k1]^T.  All pointers bf16, K-contiguous, device memory.
struct TcOperands {
    const void* a0; uint64_t a0_rows; int lda0; int k0; int a0_row;   // first K segment: rows a0_row.. of A0
    const void* a1; uint64_t a1_rows; int lda1; int k1; int a1_row;   // optional second K segment (k1 == 0: none)
    const void* w;  uint64_t w_rows;  int ldw;  uint64_t w_cols; int b_k0;
};

constexpr int kRowBoundStride = 1;                           // ints between the row bounds of consecutive streams (see TcEpiSample::row_bound)
constexpr int kBannedKeysFloatsPerRow = 64 + kRowBoundStride;

struct TcCellArgs {
    const float* addend; int64_t ld_add; const float* bias;
    float* c; float* h; int ld_state;
    void* hb; int ld_hb;
    float* raw_out; int64_t ld_raw;
    const int32_t* start; int t;      // ragged prompts (see TcEpiCell)
};

int launch_tc_store(const TcOperands& o, int M, int N, float* C, int64_t ldc, const float* bias, cudaStream_t st);
int launch_tc_cell(const TcOperands& o, int M, int N, const TcCellArgs& a, cudaStream_t st);
int tc_sample_slots(int M, int N);
int tc_mega_slots(int M, int N);       // upper bound of the slots the persistent generation kernel uses
// banned_keys: M*64 floats (race keys of the disallowed ids) followed by M ints (row bounds of the sampling
// epilogue, TcEpiSample::row_bound, one per 128-byte line): callers size the buffer M * kBannedKeysFloatsPerRow * 4 bytes.
int launch_tc_sample(const TcOperands& o, int M, int N, const SampleParams& sp, const float* bias, float* part,
                     float* zwin, float* banned_keys, cudaStream_t st);
int launch_sample_finalize(const float* part, const float* zwin, const float* banned_keys, int slots, int bs, int V,
                           const SampleParams& sp, int64_t* out, const void* table, int ld_table, void* xnext,
                           cudaStream_t st, const int64_t* forced = nullptr, const int32_t* forced_len = nullptr, int step = 0);
int launch_tc_sample_shard(const TcOperands& o, int M, int n_local, int v0, const SampleParams& sp, const float* bias,
                           float* part, float* zwin, float* banned_keys, cudaStream_t st);
int launch_shard_partial(const float* part, const float* zwin, const float* banned_keys, int slots, int bs, int v0, int v1,
                         const SampleParams& sp, float* out_key, int64_t* out_id, float* out_banned, cudaStream_t st);
int launch_shard_merge(const float* keys, const int64_t* ids, const float* banned, int R, int bs, const SampleParams& sp,
                       int64_t* out, cudaStream_t st, int64_t stride_k = 0, int64_t stride_i = 0, int64_t stride_b = 0);
int launch_gather_bf16(const void* table, int ld, int V, const int64_t* tok, int64_t n, void* out, int* err_flag,
                       cudaStream_t st);
int launch_pack_rows_bf16(const float* src, int rows, int cols, void* dst, int ld_dst, int k0, int kpad, int H,
                          cudaStream_t st);
int launch_f32_to_bf16_rows(const float* src, int64_t ld_src, int rows, int cols, void* dst, int ld_dst, cudaStream_t st);

// Whole generation loop in one persistent kernel (see mega_generate_kernel).
struct MegaArgs {
    int n_layers, bs, V, n_steps, p0, Epad;
    int H[4], Ipad[4], Hpad[4];
    const void* wcat[4]; const float* bias[4]; float* c[4]; float* h[4]; void* hb[4][2];
    float* c_blk[4];             // scratch of the size of c[l]: the kernel's blocked copy of the cell state
    const void* dec_w; const float* dec_bias;
    SampleParams sp;
    float* part; float* zwin; float* banned_keys;
    int64_t* gen; const void* table; void* xb;
    unsigned* sync_counter; unsigned long long* phase_ns;
    const int64_t* forced; const int32_t* forced_len; int forced_F;   // awd_generate_resume (nullptr / 0 otherwise)
    // <= 256 streams (one or two M tiles): scratch of the 2-D split LSTM phases (sb_scratch_bytes(), kSbCounterBytes of counters), or nullptr
    void* sb_scratch; unsigned* sb_counter;
    long long* trace;                  // debugging: 128 clock64 stamps (AWD_MEGA_TRACE), or nullptr
};
size_t sb_scratch_bytes(int n_layers, const int* H, int bs);
constexpr size_t kSbCounterBytes = (2 * 4 * 160 + 64) * sizeof(unsigned);
int launch_mega_generate(const MegaArgs& m, cudaStream_t st);

// The T x L recurrent steps of one forward() time chunk as a wavefront in one persistent kernel.
struct FwdArgs {
    int n_layers, bs, T, t_base;
    int H[4], Ipad[4], Hpad[4];
    const void* wcat[4]; const float* bias[4]; float* c[4]; float* h[4];
    const void* hb[4];                 // carried hidden operand of each layer (state before the chunk)
    void* y[4];                        // bf16 [T*bs][Hpad] outputs of each layer
    float* raw[4];                     // optional caller tensors
    const float* xproj0; const int32_t* start; unsigned* sync_counter;
};
int launch_mega_forward(const FwdArgs& m, cudaStream_t st);

// Whole request (prompt + generation) for <= 8 streams in one persistent kernel with SMEM-resident
// W_hh slices (lat_kernel.cu).
struct LatArgs {
    int n_layers, bs, V, P, n_new, Epad, par0;
    int H[4], Ipad[4], Hpad[4];
    const void* wcat[4]; const float* bias[4]; float* c[4]; void* hb[4][2];
    const void* table; const void* dec_w; const float* dec_bias;
    const int64_t* prompt; int64_t* gen;
    SampleParams sp;
    float* part; float* banned_keys; unsigned* sync_counter; int* err_flag;
    const int32_t* start;             // ragged prompts: first fed timestep per stream, or nullptr
};
int lat_max_streams();
int launch_lat_generate(const LatArgs& a, cudaStream_t st);

}  // namespace tc
}  // namespace awd
