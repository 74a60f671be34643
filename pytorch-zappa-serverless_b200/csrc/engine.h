// Engine state behind the opaque awd_engine handle (internal).
#pragma once
#include <map>
#include <string>
#include <vector>

#include "common.cuh"

namespace awd {

struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    int ensure(size_t need);   // grow-only; contents are NOT preserved; new memory is zero-filled
    void release();
    template <typename T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct Staged {
    DevBuf buf;
    std::vector<int64_t> shape;
};

// K extents are padded so that every row is a whole number of 128-bit loads (fp32 mode: x4 floats)
// or of 128-byte TMA/UMMA swizzle rows (bf16 mode: x64 bf16).  Pads are zero.
struct Layer {
    int I = 0, H = 0, Ipad = 0, Hpad = 0;
    // weights, rows gate-interleaved (4*j+g).  fp32 mode: w_ih [4H][Ipad], w_hh [4H][Hpad] fp32.
    // bf16 mode: wcat [4H][Ipad+Hpad] bf16 = [W_ih | W_hh], so one step is ONE contraction over [x ; h].
    DevBuf w_ih, w_hh, wcat, bias;
    // carried state: c fp32 [bs][Hpad].  fp32 mode: h[cur] fp32 is the hidden state (ping-pong).
    // bf16 mode: h[0] fp32 (what get_state returns) and hb[cur] bf16 (the next step's GEMM operand).
    DevBuf h[2], c, hb[2];
    DevBuf c_blk;                // persistent generation kernel: its blocked copy of c (tc::MegaArgs::c_blk)
    int cur = 0;
};

// Optional CUDA-event timing of kernel categories (bench.py's roofline numbers).
enum ProfCat { PROF_XPROJ = 0, PROF_STEP = 1, PROF_DECODER = 2, PROF_SAMPLE = 3, PROF_NCAT = 4 };
struct Profiler {
    bool on = false;
    std::vector<cudaEvent_t> pool;       // reused across runs
    std::vector<int> cats;               // category of pair i (events 2i, 2i+1)
    size_t used = 0;
};

}  // namespace awd

struct awd_engine {
    awd_config cfg{};
    bool committed = false;
    std::map<std::string, awd::Staged> staged;
    int Epad = 0;
    awd::DevBuf enc, dec_w, dec_b;          // fp32 mode: [V][Epad], [V][Epad] (alias of enc when tied); dec_b [V] both modes
    awd::DevBuf enc_bf, dec_bf;             // bf16 mode: [V][Epad] bf16
    // bf16 sampled decoder: copy of the decoder matrix whose spare K columns E, E+1 hold the bias (bf16 hi + lo);
    // the last layer's hidden operand carries 1.0 there, so the GEMM adds the bias itself (engine_bf16.cu ensure_fold)
    awd::DevBuf dec_fold;
    bool fold_ready = false;                 // dec_fold matches the committed weights
    bool fold_ones = false;                  // the 1.0 columns are in place in both hb buffers of the last layer
    std::vector<awd::Layer> layers;
    int bs = 0;                              // current number of streams (RNNCore.bs)
    int state_cap = 0;
    // workspaces
    awd::DevBuf tok, gen_out, xproj, ybuf[4], logits_ws;
    awd::DevBuf xb, part, zwin, banned_keys; // bf16 mode: gathered layer-0 input, sampler partials
    awd::DevBuf err_flag;
    awd::DevBuf start_buf;                   // ragged prompts: first fed timestep per stream (int32)
    const int32_t* cur_start = nullptr;      // non-null only while a ragged prompt is being fed
    // awd_generate_resume: prompt tokens that override the drawn ids of streams still inside their prompts
    awd::DevBuf forced_buf, forced_len_buf, reset_mask_buf;
    // host arguments of awd_generate_resume are copied into pinned staging (two, used alternately) so that the call
    // never has to wait for the stream: a serving loop can enqueue round k+1 while round k is still running
    struct PinnedStage { void* p = nullptr; size_t bytes = 0; cudaEvent_t ev = nullptr; } resume_stage[2];
    int resume_stage_idx = 0;
    PinnedStage start_stage[2];              // awd_generate_ragged: the per-stream start timesteps, same scheme
    int start_stage_idx = 0;
    const int64_t* cur_forced = nullptr;     // non-null only inside awd_generate_resume
    const int32_t* cur_forced_len = nullptr;
    int cur_forced_F = 0;
    awd::DevBuf mega_sync, mega_ns;          // persistent generation kernel: grid-barrier counter, per-phase ns
    awd::DevBuf sb_scratch, sb_counter;      // <= 128 streams: partial-tile exchange of the 2-D split LSTM phases
    int64_t mega_steps = 0;                  // token steps run by it while profiling was on
    awd::Profiler prof;
    bool bf16() const { return cfg.precision == AWD_PREC_BF16; }
    const float* dec_weight() const { return cfg.tie_weights ? enc.as<float>() : dec_w.as<float>(); }
    const void* dec_weight_bf() const { return cfg.tie_weights ? enc_bf.p : dec_bf.p; }
};

namespace awd {

struct ProfScope {
    awd_engine* e; cudaStream_t st; cudaEvent_t stop = nullptr;
    ProfScope(awd_engine* e_, int cat, cudaStream_t st_) : e(e_), st(st_) {
        Profiler& p = e->prof;
        if (!p.on) return;
        while (p.pool.size() < 2 * (p.used + 1)) {
            cudaEvent_t ev;
            if (cudaEventCreate(&ev) != cudaSuccess) return;
            p.pool.push_back(ev);
        }
        cudaEventRecord(p.pool[2 * p.used], st);
        stop = p.pool[2 * p.used + 1];
        p.cats.push_back(cat);
        p.used++;
    }
    ~ProfScope() { if (stop) cudaEventRecord(stop, st); }
};

// bf16 / tcgen05 path (engine_bf16.cu)
int bf16_commit(awd_engine* e, const float* enc_src, const float* dec_src, const float* const* w_ih,
                const float* const* w_hh, cudaStream_t st);
int bf16_forward(awd_engine* e, const int64_t* tok, int sl, int bs, float* logits, float* const* raw_outputs,
                 int last_step_only, cudaStream_t st);
int bf16_generate_loop(awd_engine* e, int bs, int n_new, SampleParams sp, int64_t* gen, cudaStream_t st);
int bf16_decode(awd_engine* e, const float* x, int64_t rows, float* logits, cudaStream_t st);
int bf16_gather(awd_engine* e, const int64_t* tok, int64_t n, float* out, cudaStream_t st);
int bf16_sync_hidden_operand(awd_engine* e, int layer, cudaStream_t st);
int bf16_generate_few_streams(awd_engine* e, int P, int bs, int n_new, SampleParams sp, int64_t* gen, cudaStream_t st);
bool bf16_use_few_streams_kernel(const awd_engine* e, int bs);
int bf16_decode_race_shard(awd_engine* e, int v0, int v1, SampleParams sp, float* out_key, int64_t* out_id, float* out_banned,
                           cudaStream_t st);

}  // namespace awd
