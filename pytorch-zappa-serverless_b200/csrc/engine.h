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
    int ensure(size_t need);   // grow-only; contents are NOT preserved
    void release();
    template <typename T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct Staged {
    DevBuf buf;
    std::vector<int64_t> shape;
};

struct Layer {
    int I = 0, H = 0, Ipad = 0, Hpad = 0;
    // fp32 mode: rows gate-interleaved (4*j+g), K padded to a multiple of 4
    DevBuf w_ih, w_hh, bias;
    // carried state, fp32: h[cur] is RNNCore.hidden[l][0], c is hidden[l][1]; rows are Hpad wide
    DevBuf h[2], c;
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
    awd::DevBuf enc, dec_w, dec_b;          // fp32 [V][Epad], [V][Epad] (alias of enc when tied), [V]
    std::vector<awd::Layer> layers;
    int bs = 0;                              // current number of streams (RNNCore.bs)
    int state_cap = 0;
    // workspaces
    awd::DevBuf tok, gen_out, xproj, ybuf[2], logits_ws;
    awd::DevBuf err_flag;
    awd::Profiler prof;
    const float* dec_weight() const { return cfg.tie_weights ? enc.as<float>() : dec_w.as<float>(); }
};
