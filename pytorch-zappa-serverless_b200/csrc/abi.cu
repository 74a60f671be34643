// C ABI of the AWD-LSTM engine: handle lifecycle, strict state_dict ingestion, weight packing,
// forward and the device-side generation loop.  See include/awd_b200.h for the contract.
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "engine.h"
#include "tc_api.h"

namespace awd {

thread_local char g_err[512] = "";
std::atomic<int64_t> g_launches{0};

int set_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

int DevBuf::ensure(size_t need) {
    if (need <= bytes) return AWD_OK;
    release();
    cudaError_t e = cudaMalloc(&p, need);
    if (e != cudaSuccess) {
        p = nullptr;
        cudaGetLastError();
        return set_error(AWD_ENOMEM, "cudaMalloc(%zu bytes): %s", need, cudaGetErrorString(e));
    }
    bytes = need;
    e = cudaMemset(p, 0, need);
    if (e != cudaSuccess) return set_error(AWD_ECUDA, "cudaMemset: %s", cudaGetErrorString(e));
    return AWD_OK;
}
void DevBuf::release() {
    if (p) cudaFree(p);
    p = nullptr;
    bytes = 0;
}

static std::vector<std::pair<std::string, std::vector<int64_t>>> expected_keys(const awd_config& c) {
    std::vector<std::pair<std::string, std::vector<int64_t>>> k;
    k.push_back({"0.encoder.weight", {c.vocab_sz, c.emb_sz}});
    k.push_back({"0.encoder_dp.emb.weight", {c.vocab_sz, c.emb_sz}});
    for (int l = 0; l < c.n_layers; ++l) {
        int64_t I = l == 0 ? c.emb_sz : c.n_hid;
        int64_t H = l != c.n_layers - 1 ? c.n_hid : c.emb_sz;
        std::string p = "0.rnns." + std::to_string(l) + ".";
        k.push_back({p + "weight_hh_l0_raw", {4 * H, H}});
        k.push_back({p + "module.weight_ih_l0", {4 * H, I}});
        k.push_back({p + "module.weight_hh_l0", {4 * H, H}});
        k.push_back({p + "module.bias_ih_l0", {4 * H}});
        k.push_back({p + "module.bias_hh_l0", {4 * H}});
    }
    k.push_back({"1.decoder.weight", {c.vocab_sz, c.emb_sz}});
    if (c.has_decoder_bias) k.push_back({"1.decoder.bias", {c.vocab_sz}});
    return k;
}

struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
        if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

static int ensure_state(awd_engine* e, int bs) {
    if (bs <= e->state_cap) return AWD_OK;
    for (auto& L : e->layers) {
        size_t need = (size_t)bs * L.Hpad * sizeof(float);
        for (int i = 0; i < 2; ++i) {
            int rc = L.h[i].ensure(need);
            if (rc) return rc;
        }
        int rc = L.c.ensure(need);
        if (rc) return rc;
        if (e->bf16())
            for (int i = 0; i < 2; ++i)
                if ((rc = L.hb[i].ensure((size_t)bs * L.Hpad * 2))) return rc;
    }
    e->state_cap = bs;
    return AWD_OK;
}

static int reset_impl(awd_engine* e, int bs, cudaStream_t st) {
    if (bs <= 0) return set_error(AWD_EINVAL, "reset: bs must be positive (got %d)", bs);
    // growing reallocates (and zero-fills) synchronously; otherwise clear on the stream
    int rc = ensure_state(e, bs);
    if (rc) return rc;
    for (auto& L : e->layers) {
        size_t n = (size_t)bs * L.Hpad * sizeof(float);
        AWD_CUDA(cudaMemsetAsync(L.h[0].p, 0, n, st));
        AWD_CUDA(cudaMemsetAsync(L.h[1].p, 0, n, st));
        AWD_CUDA(cudaMemsetAsync(L.c.p, 0, n, st));
        if (e->bf16()) {
            AWD_CUDA(cudaMemsetAsync(L.hb[0].p, 0, n / 2, st));
            AWD_CUDA(cudaMemsetAsync(L.hb[1].p, 0, n / 2, st));
        }
        L.cur = 0;
    }
    e->bs = bs;
    e->fold_ones = false;
    return AWD_OK;
}

// One pass of RNNCore.forward + LinearDecoder.forward over device-resident token ids.
static int forward_impl(awd_engine* e, const int64_t* tok, int sl, int bs, float* logits, float* const* raw_outputs,
                        int last_step_only, cudaStream_t st) {
    if (e->bf16()) return bf16_forward(e, tok, sl, bs, logits, raw_outputs, last_step_only, st);
    const awd_config& c = e->cfg;
    const int V = c.vocab_sz;
    const int nl = c.n_layers;
    int* flag = e->err_flag.as<int>();
    const int chunk_T = std::max(1, 32768 / bs);
    int maxHpad = 0, maxH = 0;
    for (auto& L : e->layers) { maxHpad = std::max(maxHpad, L.Hpad); maxH = std::max(maxH, L.H); }
    if (sl > 1) {
        size_t rows = (size_t)std::min(sl, chunk_T) * bs;
        int rc = e->xproj.ensure(rows * 4 * maxH * sizeof(float));
        if (rc) return rc;
        for (int i = 0; i < 2; ++i) {
            rc = e->ybuf[i].ensure(rows * maxHpad * sizeof(float));
            if (rc) return rc;
        }
    }
    for (int t0 = 0; t0 < sl; t0 += chunk_T) {
        const int T = std::min(chunk_T, sl - t0);
        const int64_t R = (int64_t)T * bs;
        const float* prev_y = nullptr;   // output rows of the layer below for this chunk
        int prev_ld = 0;
        for (int l = 0; l < nl; ++l) {
            Layer& L = e->layers[l];
            Operand x{};
            if (l == 0) {
                x.A = e->enc.as<float>(); x.idx = tok + (int64_t)t0 * bs; x.lda = e->Epad;
                x.idx_limit = V; x.err_flag = flag;
            } else {
                x.A = prev_y; x.idx = nullptr; x.lda = prev_ld;
            }
            x.W = L.w_ih.as<float>(); x.ldw = L.Ipad; x.K = L.Ipad;
            float* raw = (raw_outputs && raw_outputs[l]) ? raw_outputs[l] + (int64_t)t0 * bs * L.H : nullptr;
            if (T == 1) {
                Operand h{};
                h.A = L.h[L.cur].as<float>(); h.lda = L.Hpad; h.W = L.w_hh.as<float>(); h.ldw = L.Hpad; h.K = L.Hpad;
                EpiCell epi{};
                epi.addend = nullptr; epi.bias = L.bias.as<float>();
                epi.c = L.c.as<float>(); epi.h_out = L.h[L.cur ^ 1].as<float>(); epi.ld_state = L.Hpad;
                epi.raw_out = raw; epi.ld_raw = L.H; epi.H = L.H;
                epi.start = e->cur_start; epi.t = t0;
                int rc;
                { ProfScope ps(e, PROF_STEP, st); rc = launch_gemm_cell_f32(x, &h, bs, 4 * L.H, epi, st); }
                if (rc) return rc;
                L.cur ^= 1;
                prev_y = L.h[L.cur].as<float>();
                prev_ld = L.Hpad;
            } else {
                float* xp = e->xproj.as<float>();
                float* Y = e->ybuf[l & 1].as<float>();
                EpiStore es{xp, (int64_t)4 * L.H, L.bias.as<float>()};
                int rc;
                { ProfScope ps(e, PROF_XPROJ, st); rc = launch_gemm_store_f32(x, (int)R, 4 * L.H, es, st); }
                if (rc) return rc;
                for (int t = 0; t < T; ++t) {
                    Operand h{};
                    h.A = (t == 0) ? L.h[L.cur].as<float>() : Y + (int64_t)(t - 1) * bs * L.Hpad;
                    h.lda = L.Hpad; h.W = L.w_hh.as<float>(); h.ldw = L.Hpad; h.K = L.Hpad;
                    EpiCell epi{};
                    epi.addend = xp + (int64_t)t * bs * 4 * L.H; epi.ld_add = 4 * L.H; epi.bias = nullptr;
                    epi.c = L.c.as<float>(); epi.h_out = Y + (int64_t)t * bs * L.Hpad; epi.ld_state = L.Hpad;
                    epi.raw_out = raw ? raw + (int64_t)t * bs * L.H : nullptr; epi.ld_raw = L.H; epi.H = L.H;
                    epi.start = e->cur_start; epi.t = t0 + t;
                    { ProfScope ps(e, PROF_STEP, st); rc = launch_gemm_cell_f32(h, nullptr, bs, 4 * L.H, epi, st); }
                    if (rc) return rc;
                }
                AWD_CUDA(cudaMemcpyAsync(L.h[L.cur].p, Y + (int64_t)(T - 1) * bs * L.Hpad,
                                         (size_t)bs * L.Hpad * sizeof(float), cudaMemcpyDeviceToDevice, st));
                prev_y = Y;
                prev_ld = L.Hpad;
            }
        }
        if (logits != nullptr) {
            Operand d{};
            d.W = e->dec_weight(); d.ldw = e->Epad; d.K = e->Epad; d.lda = prev_ld;
            const float* bias = c.has_decoder_bias ? e->dec_b.as<float>() : nullptr;
            if (!last_step_only) {
                d.A = prev_y;
                EpiStore es{logits + (int64_t)t0 * bs * V, (int64_t)V, bias};
                int rc;
                { ProfScope ps(e, PROF_DECODER, st); rc = launch_gemm_store_f32(d, (int)R, V, es, st); }
                if (rc) return rc;
            } else if (t0 + T == sl) {
                d.A = prev_y + (int64_t)(T - 1) * bs * prev_ld;
                EpiStore es{logits, (int64_t)V, bias};
                int rc;
                { ProfScope ps(e, PROF_DECODER, st); rc = launch_gemm_store_f32(d, bs, V, es, st); }
                if (rc) return rc;
            }
        }
    }
    return AWD_OK;
}

// fp32 mode, more than 8 streams: decoder + sampling race on rows `x` (bs rows of pitch ldx, the last layer's output) -> ids, the logits
// never stored (kernels_f32.cu EpiRace).  Opt-in (AWD_F32_FUSE=1, read at every call): measured at 1024 streams x V = 60 000 the fused
// kernel takes 1630 us per token against 1165 + 230 us for the decoder that writes the logits plus the sampler that reads them -- the
// FMA-bound GEMM's epilogue is serial per tile, the stand-alone sampler is a bandwidth-bound kernel over all SMs, and the 245 MB of
// logits it round-trips are ~75 us of HBM time in a 3 ms step.  The ids of the two forms are identical (tests/test_gpu_parity.py).
static bool f32_fused_race(int bs) {
    const char* sv = getenv("AWD_F32_FUSE");
    return sv != nullptr && atoi(sv) != 0 && bs > 8;
}
static int f32_decode_race(awd_engine* e, const float* x, int ldx, int bs, const SampleParams& sp, int64_t* out, cudaStream_t st) {
    const awd_config& c = e->cfg;
    const int V = c.vocab_sz;
    int rc = e->part.ensure((size_t)bs * race_tiles_f32(bs, V) * sizeof(float4));
    if (rc) return rc;
    if ((rc = e->banned_keys.ensure((size_t)bs * 64 * sizeof(float)))) return rc;
    Operand d{};
    d.A = x; d.lda = ldx; d.W = e->dec_weight(); d.ldw = e->Epad; d.K = e->Epad;
    ProfScope ps(e, PROF_DECODER, st);
    return launch_gemm_race_f32(d, bs, V, c.has_decoder_bias ? e->dec_b.as<float>() : nullptr, sp, e->part.as<float4>(), e->banned_keys.as<float>(), out, st);
}

static int make_sample_params(const awd_sampling* s, int V, SampleParams& sp) {
    if (!s) return set_error(AWD_EINVAL, "sampling parameters missing");
    if (!(s->temperature > 0.f)) return set_error(AWD_EINVAL, "temperature must be > 0");
    if (s->n_draws < 1) return set_error(AWD_EINVAL, "n_draws must be >= 1");
    if (s->n_banned < 0 || s->n_banned > 63) return set_error(AWD_ELIMIT, "at most 63 banned ids (got %d)", s->n_banned);
    std::vector<int> b;
    if (s->ban_zero) b.push_back(0);
    for (int i = 0; i < s->n_banned; ++i) {
        int64_t id = s->banned[i];
        if (id < 0 || id >= V) return set_error(AWD_EINVAL, "banned id %lld outside the vocabulary", (long long)id);
        b.push_back((int)id);
    }
    std::sort(b.begin(), b.end());
    b.erase(std::unique(b.begin(), b.end()), b.end());
    memset(&sp, 0, sizeof(sp));
    sp.inv_temperature = 1.0f / s->temperature;
    sp.greedy = s->greedy ? 1 : 0;
    sp.n_draws = s->n_draws;
    sp.n_banned = (int)b.size();
    for (size_t i = 0; i < b.size(); ++i) sp.banned[i] = b[i];
    sp.seed_lo = (uint32_t)(s->seed & 0xffffffffu);
    sp.seed_hi = (uint32_t)(s->seed >> 32);
    return AWD_OK;
}

}  // namespace awd

using namespace awd;

#define AWD_ENGINE_ENTER(e)                                                                       \
    if (!(e)) return set_error(AWD_EINVAL, "null engine");                                        \
    DeviceGuard _guard((e)->cfg.device);                                                          \
    if (!_guard.ok) return set_error(AWD_ECUDA, "cannot select device %d", (e)->cfg.device)

extern "C" {

const char* awd_last_error(void) { return g_err; }
const char* awd_version(void) { return "awd_b200 0.1 (sm_100a)"; }
int64_t awd_launch_count(void) { return g_launches.load(); }

int awd_create(const awd_config* cfg, awd_engine** out) {
    if (!cfg || !out) return set_error(AWD_EINVAL, "null argument");
    *out = nullptr;
    if (cfg->vocab_sz < 1 || cfg->emb_sz < 1 || cfg->n_hid < 1 || cfg->n_layers < 1)
        return set_error(AWD_EINVAL, "vocab_sz, emb_sz, n_hid, n_layers must be positive");
    if (cfg->precision != AWD_PREC_FP32 && cfg->precision != AWD_PREC_BF16)
        return set_error(AWD_EINVAL, "unknown precision %d", cfg->precision);
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return set_error(AWD_ENODEV, "no CUDA device: this engine has no CPU fallback");
    }
    if (cfg->device < 0 || cfg->device >= ndev) return set_error(AWD_EINVAL, "device %d of %d", cfg->device, ndev);
    cudaDeviceProp prop;
    AWD_CUDA(cudaGetDeviceProperties(&prop, cfg->device));
    if (prop.major != 10)
        return set_error(AWD_ENODEV, "device %d is sm_%d%d; this library is built for sm_100a only", cfg->device,
                         prop.major, prop.minor);
    DeviceGuard guard(cfg->device);
    if (!guard.ok) return set_error(AWD_ECUDA, "cannot select device %d", cfg->device);
    awd_engine* e = new awd_engine();
    e->cfg = *cfg;
    const int kpad = cfg->precision == AWD_PREC_BF16 ? 64 : 4;
    e->Epad = round_up(cfg->emb_sz, kpad);
    e->layers.resize(cfg->n_layers);
    for (int l = 0; l < cfg->n_layers; ++l) {
        Layer& L = e->layers[l];
        L.I = l == 0 ? cfg->emb_sz : cfg->n_hid;
        L.H = l != cfg->n_layers - 1 ? cfg->n_hid : cfg->emb_sz;
        L.Ipad = round_up(L.I, kpad);
        L.Hpad = round_up(L.H, kpad);
    }
    int rc = e->err_flag.ensure(sizeof(int));
    if (rc) { delete e; return rc; }
    *out = e;
    return AWD_OK;
}

int awd_destroy(awd_engine* e) {
    if (!e) return AWD_OK;
    {
        DeviceGuard guard(e->cfg.device);
        cudaDeviceSynchronize();
        for (auto& kv : e->staged) kv.second.buf.release();
        e->enc.release(); e->dec_w.release(); e->dec_b.release();
        for (auto* arr : {e->resume_stage, e->start_stage})
            for (int i = 0; i < 2; ++i) { auto& S = arr[i]; if (S.p) cudaFreeHost(S.p); if (S.ev) cudaEventDestroy(S.ev); S = {}; }
        e->enc_bf.release(); e->dec_bf.release(); e->dec_fold.release(); e->xb.release(); e->part.release(); e->zwin.release(); e->banned_keys.release(); e->mega_sync.release(); e->mega_ns.release(); e->start_buf.release();
        e->forced_buf.release(); e->forced_len_buf.release(); e->reset_mask_buf.release(); e->sb_scratch.release(); e->sb_counter.release();
        for (auto& L : e->layers) {
            L.w_ih.release(); L.w_hh.release(); L.bias.release();
            L.h[0].release(); L.h[1].release(); L.c.release(); L.c_blk.release();
            L.wcat.release(); L.hb[0].release(); L.hb[1].release();
        }
        e->tok.release(); e->gen_out.release(); e->xproj.release();
        e->ybuf[0].release(); e->ybuf[1].release(); e->ybuf[2].release(); e->ybuf[3].release(); e->logits_ws.release(); e->err_flag.release();
        for (auto ev : e->prof.pool) cudaEventDestroy(ev);
    }
    delete e;
    return AWD_OK;
}

int awd_load_param(awd_engine* e, const char* key, const float* data, const int64_t* shape, int32_t ndim) {
    AWD_ENGINE_ENTER(e);
    if (!key || !data || !shape) return set_error(AWD_EINVAL, "null argument");
    for (auto& kv : expected_keys(e->cfg)) {
        if (kv.first != key) continue;
        if ((size_t)ndim != kv.second.size()) return set_error(AWD_EINVAL, "%s: expected %zu dims, got %d", key, kv.second.size(), ndim);
        size_t n = 1;
        for (int i = 0; i < ndim; ++i) {
            if (shape[i] != kv.second[i])
                return set_error(AWD_EINVAL, "size mismatch for %s: dim %d is %lld, expected %lld", key, i,
                                 (long long)shape[i], (long long)kv.second[i]);
            n *= (size_t)shape[i];
        }
        Staged& s = e->staged[key];
        int rc = s.buf.ensure(n * sizeof(float));
        if (rc) return rc;
        s.shape.assign(shape, shape + ndim);
        AWD_CUDA(cudaMemcpy(s.buf.p, data, n * sizeof(float), cudaMemcpyDefault));
        e->committed = false;
        return AWD_OK;
    }
    return set_error(AWD_EINVAL, "unexpected key in state_dict: %s", key);
}

int awd_commit_weights(awd_engine* e) {
    AWD_ENGINE_ENTER(e);
    const awd_config& c = e->cfg;
    std::string missing;
    for (auto& kv : expected_keys(c))
        if (!e->staged.count(kv.first)) missing += (missing.empty() ? "" : ", ") + kv.first;
    if (!missing.empty()) return set_error(AWD_EMISSING, "missing key(s) in state_dict: %s", missing.c_str());
    cudaStream_t st = 0;
    auto src = [&](const std::string& k) { return e->staged[k].buf.as<float>(); };
    // alias resolution: the key load_state_dict copies last is the surviving value
    const float* enc_src = c.tie_weights ? src("1.decoder.weight") : src("0.encoder_dp.emb.weight");
    int rc = AWD_OK;
    if (c.has_decoder_bias) {
        rc = e->dec_b.ensure((size_t)c.vocab_sz * sizeof(float));
        if (rc) return rc;
        AWD_CUDA(cudaMemcpyAsync(e->dec_b.p, src("1.decoder.bias"), (size_t)c.vocab_sz * sizeof(float),
                                 cudaMemcpyDeviceToDevice, st));
    }
    std::vector<const float*> wih(c.n_layers), whh(c.n_layers);
    for (int l = 0; l < c.n_layers; ++l) {
        Layer& L = e->layers[l];
        std::string p = "0.rnns." + std::to_string(l) + ".module.";
        wih[l] = src(p + "weight_ih_l0");
        whh[l] = src(p + "weight_hh_l0");
        if ((rc = L.bias.ensure((size_t)4 * L.H * sizeof(float)))) return rc;
        if ((rc = launch_bias_sum_f32(src(p + "bias_ih_l0"), src(p + "bias_hh_l0"), L.H, L.bias.as<float>(), st))) return rc;
    }
    if (e->bf16()) {
        if ((rc = bf16_commit(e, enc_src, src("1.decoder.weight"), wih.data(), whh.data(), st))) return rc;
    } else {
        if ((rc = e->enc.ensure((size_t)c.vocab_sz * e->Epad * sizeof(float)))) return rc;
        if ((rc = launch_pack_rows_f32(enc_src, c.vocab_sz, c.emb_sz, e->enc.as<float>(), e->Epad, 0, st))) return rc;
        if (!c.tie_weights) {
            if ((rc = e->dec_w.ensure((size_t)c.vocab_sz * e->Epad * sizeof(float)))) return rc;
            if ((rc = launch_pack_rows_f32(src("1.decoder.weight"), c.vocab_sz, c.emb_sz, e->dec_w.as<float>(), e->Epad, 0, st))) return rc;
        }
        for (int l = 0; l < c.n_layers; ++l) {
            Layer& L = e->layers[l];
            if ((rc = L.w_ih.ensure((size_t)4 * L.H * L.Ipad * sizeof(float)))) return rc;
            if ((rc = L.w_hh.ensure((size_t)4 * L.H * L.Hpad * sizeof(float)))) return rc;
            if ((rc = launch_pack_rows_f32(wih[l], 4 * L.H, L.I, L.w_ih.as<float>(), L.Ipad, L.H, st))) return rc;
            if ((rc = launch_pack_rows_f32(whh[l], 4 * L.H, L.H, L.w_hh.as<float>(), L.Hpad, L.H, st))) return rc;
        }
    }
    AWD_CUDA(cudaStreamSynchronize(st));
    for (auto& kv : e->staged) kv.second.buf.release();   // staging is only needed until packed
    e->staged.clear();
    e->committed = true;
    if (e->bs == 0) return reset_impl(e, 1, st);          // RNNCore.__init__: bs = 1 (awd_lstm.py:57)
    return AWD_OK;
}

int awd_reset(awd_engine* e, int32_t bs, void* stream) {
    AWD_ENGINE_ENTER(e);
    return reset_impl(e, bs, (cudaStream_t)stream);
}

int32_t awd_current_bs(const awd_engine* e) { return e ? e->bs : 0; }

static int state_copy(awd_engine* e, int layer, float* h, float* c, bool to_user, cudaStream_t st) {
    if (layer < 0 || layer >= e->cfg.n_layers) return set_error(AWD_EINVAL, "layer %d out of range", layer);
    if (e->bs <= 0) return set_error(AWD_ESTATE, "no state yet: call awd_reset first");
    Layer& L = e->layers[layer];
    size_t w = (size_t)L.H * sizeof(float), pitch = (size_t)L.Hpad * sizeof(float);
    void* hstate = e->bf16() ? L.h[0].p : L.h[L.cur].p;
    if (h) {
        if (to_user) AWD_CUDA(cudaMemcpy2DAsync(h, w, hstate, pitch, w, e->bs, cudaMemcpyDefault, st));
        else {
            AWD_CUDA(cudaMemcpy2DAsync(hstate, pitch, h, w, w, e->bs, cudaMemcpyDefault, st));
            if (e->bf16()) {
                int rc = bf16_sync_hidden_operand(e, layer, st);
                if (rc) return rc;
            }
        }
    }
    if (c) {
        if (to_user) AWD_CUDA(cudaMemcpy2DAsync(c, w, L.c.p, pitch, w, e->bs, cudaMemcpyDefault, st));
        else AWD_CUDA(cudaMemcpy2DAsync(L.c.p, pitch, c, w, w, e->bs, cudaMemcpyDefault, st));
    }
    return AWD_OK;
}

int awd_get_state(awd_engine* e, int32_t layer, float* h, float* c, void* stream) {
    AWD_ENGINE_ENTER(e);
    return state_copy(e, layer, h, c, true, (cudaStream_t)stream);
}
int awd_set_state(awd_engine* e, int32_t layer, const float* h, const float* c, void* stream) {
    AWD_ENGINE_ENTER(e);
    return state_copy(e, layer, const_cast<float*>(h), const_cast<float*>(c), false, (cudaStream_t)stream);
}

int awd_gather(awd_engine* e, const int64_t* tokens, int64_t n, float* out, void* stream) {
    AWD_ENGINE_ENTER(e);
    if (!e->committed) return set_error(AWD_ESTATE, "weights not committed");
    if (n < 0 || (n > 0 && (!tokens || !out))) return set_error(AWD_EINVAL, "bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    int rc = e->tok.ensure((size_t)std::max<int64_t>(n, 1) * sizeof(int64_t));
    if (rc) return rc;
    AWD_CUDA(cudaMemcpyAsync(e->tok.p, tokens, (size_t)n * sizeof(int64_t), cudaMemcpyDefault, st));
    if (e->bf16()) return bf16_gather(e, e->tok.as<int64_t>(), n, out, st);
    return launch_gather_f32(e->enc.as<float>(), e->Epad, e->cfg.emb_sz, e->cfg.vocab_sz, e->tok.as<int64_t>(), n, out,
                             e->cfg.emb_sz, e->err_flag.as<int>(), st);
}

int awd_forward(awd_engine* e, const int64_t* tokens, int32_t sl, int32_t bs, float* logits,
                float* const* raw_outputs, int32_t last_step_only, void* stream) {
    AWD_ENGINE_ENTER(e);
    if (!e->committed) return set_error(AWD_ESTATE, "weights not committed");
    if (!tokens || sl < 1 || bs < 1) return set_error(AWD_EINVAL, "forward needs tokens of shape (sl>=1, bs>=1)");
    cudaStream_t st = (cudaStream_t)stream;
    if (bs != e->bs) {                       // awd_lstm.py:79-81
        int rc = reset_impl(e, bs, st);
        if (rc) return rc;
    }
    size_t n = (size_t)sl * bs;
    int rc = e->tok.ensure(n * sizeof(int64_t));
    if (rc) return rc;
    AWD_CUDA(cudaMemcpyAsync(e->tok.p, tokens, n * sizeof(int64_t), cudaMemcpyDefault, st));
    return forward_impl(e, e->tok.as<int64_t>(), sl, bs, logits, raw_outputs, last_step_only, st);
}

int awd_decode(awd_engine* e, const float* x, int64_t rows, float* logits, void* stream) {
    AWD_ENGINE_ENTER(e);
    if (!e->committed) return set_error(AWD_ESTATE, "weights not committed");
    if (rows < 0 || (rows > 0 && (!x || !logits))) return set_error(AWD_EINVAL, "bad arguments");
    if (rows == 0) return AWD_OK;
    if (rows > INT32_MAX) return set_error(AWD_ELIMIT, "too many rows");
    cudaStream_t st = (cudaStream_t)stream;
    const awd_config& c = e->cfg;
    if (e->bf16()) return bf16_decode(e, x, rows, logits, st);
    Operand d{};
    d.W = e->dec_weight(); d.ldw = e->Epad; d.K = e->Epad;
    if (c.emb_sz == e->Epad && (((uintptr_t)x) & 15) == 0) {
        d.A = x; d.lda = c.emb_sz;
    } else {   // re-pitch rows to a 16-byte multiple; the pad columns of the workspace stay zero
        int rc = e->ybuf[0].ensure((size_t)rows * e->Epad * sizeof(float));
        if (rc) return rc;
        AWD_CUDA(cudaMemcpy2DAsync(e->ybuf[0].p, (size_t)e->Epad * sizeof(float), x, (size_t)c.emb_sz * sizeof(float),
                                   (size_t)c.emb_sz * sizeof(float), (size_t)rows, cudaMemcpyDeviceToDevice, st));
        d.A = e->ybuf[0].as<float>(); d.lda = e->Epad;
    }
    EpiStore es{logits, (int64_t)c.vocab_sz, c.has_decoder_bias ? e->dec_b.as<float>() : nullptr};
    return launch_gemm_store_f32(d, (int)rows, c.vocab_sz, es, st);
}

int awd_sample(awd_engine* e, const float* logits, int32_t bs, const awd_sampling* s, int64_t step, int64_t* out,
               void* stream) {
    AWD_ENGINE_ENTER(e);
    if (!logits || !out || bs < 1) return set_error(AWD_EINVAL, "bad arguments");
    SampleParams sp;
    int rc = make_sample_params(s, e->cfg.vocab_sz, sp);
    if (rc) return rc;
    sp.step_lo = (uint32_t)((uint64_t)step & 0xffffffffu);
    sp.step_hi = (uint32_t)((uint64_t)step >> 32);
    return launch_sample_f32(logits, e->cfg.vocab_sz, e->cfg.vocab_sz, bs, sp, out, nullptr, (cudaStream_t)stream);
}

static int generate_impl(awd_engine* e, const int64_t* prompt, const int32_t* prompt_len, int32_t P, int32_t bs,
                         int32_t n_new, const awd_sampling* s, int64_t* out_tokens, void* stream);

// Pinned staging for host arguments (used alternately in pairs): wait until the copies enqueued out of it two calls
// ago are done, grow it if needed.  The caller records S.ev once everything that reads the staging is enqueued.
static int acquire_stage(awd_engine::PinnedStage& S, size_t need) {
    if (S.ev == nullptr) AWD_CUDA(cudaEventCreateWithFlags(&S.ev, cudaEventDisableTiming));
    else AWD_CUDA(cudaEventSynchronize(S.ev));
    if (S.bytes < need) {
        if (S.p) cudaFreeHost(S.p);
        S.p = nullptr; S.bytes = 0;
        AWD_CUDA(cudaMallocHost(&S.p, need));
        S.bytes = need;
    }
    return AWD_OK;
}

int awd_generate(awd_engine* e, const int64_t* prompt, int32_t P, int32_t bs, int32_t n_new, const awd_sampling* s,
                 int64_t* out_tokens, void* stream) {
    return generate_impl(e, prompt, nullptr, P, bs, n_new, s, out_tokens, stream);
}

int awd_generate_ragged(awd_engine* e, const int64_t* prompt, const int32_t* prompt_len, int32_t P, int32_t bs,
                        int32_t n_new, const awd_sampling* s, int64_t* out_tokens, void* stream) {
    if (!prompt_len) return set_error(AWD_EINVAL, "prompt_len missing");
    return generate_impl(e, prompt, prompt_len, P, bs, n_new, s, out_tokens, stream);
}

static int generate_impl(awd_engine* e, const int64_t* prompt, const int32_t* prompt_len, int32_t P, int32_t bs,
                         int32_t n_new, const awd_sampling* s, int64_t* out_tokens, void* stream) {
    AWD_ENGINE_ENTER(e);
    if (!e->committed) return set_error(AWD_ESTATE, "weights not committed");
    if (!prompt || P < 1 || bs < 1 || n_new < 0 || (n_new > 0 && !out_tokens))
        return set_error(AWD_EINVAL, "generate needs a prompt of shape (P>=1, bs>=1) and n_new >= 0");
    cudaStream_t st = (cudaStream_t)stream;
    const int V = e->cfg.vocab_sz;
    SampleParams sp;
    int rc = make_sample_params(s, V, sp);
    if (rc) return rc;
    if ((rc = reset_impl(e, bs, st))) return rc;         // main.py:51 model.reset()
    if ((rc = e->tok.ensure((size_t)P * bs * sizeof(int64_t)))) return rc;
    if ((rc = e->gen_out.ensure((size_t)std::max(n_new, 1) * bs * sizeof(int64_t)))) return rc;
    AWD_CUDA(cudaMemcpyAsync(e->tok.p, prompt, (size_t)P * bs * sizeof(int64_t), cudaMemcpyDefault, st));
    // ragged prompts: stream b only has prompt_len[b] tokens, right-aligned in the (P,bs) block; it joins at
    // timestep P - prompt_len[b] with the reset state (main.py:54-57 feeds each request only its own words)
    struct StartScope {
        awd_engine* e;
        ~StartScope() { e->cur_start = nullptr; }
    } start_scope{e};
    if (prompt_len != nullptr) {
        for (int b = 0; b < bs; ++b)
            if (prompt_len[b] < 1 || prompt_len[b] > P)
                return set_error(AWD_EINVAL, "prompt_len[%d] = %d outside [1, %d]", b, prompt_len[b], P);
        // the start timesteps go through one of two pinned staging areas, so the call never waits for the stream
        awd_engine::PinnedStage& S = e->start_stage[e->start_stage_idx ^= 1];
        if ((rc = acquire_stage(S, (size_t)bs * sizeof(int32_t)))) return rc;
        int32_t* start = static_cast<int32_t*>(S.p);
        for (int b = 0; b < bs; ++b) start[b] = P - prompt_len[b];
        if ((rc = e->start_buf.ensure((size_t)bs * sizeof(int32_t)))) return rc;
        AWD_CUDA(cudaMemcpyAsync(e->start_buf.p, start, (size_t)bs * sizeof(int32_t), cudaMemcpyHostToDevice, st));
        AWD_CUDA(cudaEventRecord(S.ev, st));
        e->cur_start = e->start_buf.as<int32_t>();
    }
    if (e->bf16() && bf16_use_few_streams_kernel(e, bs)) {
        if ((rc = bf16_generate_few_streams(e, P, bs, n_new, sp, e->gen_out.as<int64_t>(), st))) return rc;
        if (n_new > 0)
            AWD_CUDA(cudaMemcpyAsync(out_tokens, e->gen_out.p, (size_t)n_new * bs * sizeof(int64_t), cudaMemcpyDefault, st));
        return AWD_OK;
    }
    if (e->bf16()) {
        // prompt pass without any decoder work, then the fused decoder+sampler loop
        if ((rc = bf16_forward(e, e->tok.as<int64_t>(), P, bs, nullptr, nullptr, 1, st))) return rc;
        e->cur_start = nullptr;              // every stream is live from here on
        if ((rc = bf16_generate_loop(e, bs, n_new, sp, e->gen_out.as<int64_t>(), st))) return rc;
        if (n_new > 0)
            AWD_CUDA(cudaMemcpyAsync(out_tokens, e->gen_out.p, (size_t)n_new * bs * sizeof(int64_t), cudaMemcpyDefault, st));
        return AWD_OK;
    }
    int64_t* gen = e->gen_out.as<int64_t>();
    if (f32_fused_race(bs)) {
        // prompt without any decoder work, then per token {decoder + race on the current state -> id, feed it} (main.py:60-71)
        if ((rc = forward_impl(e, e->tok.as<int64_t>(), P, bs, nullptr, nullptr, 1, st))) return rc;
        e->cur_start = nullptr;              // every stream is live from here on
        for (int i = 0; i < n_new; ++i) {
            sp.step_lo = (uint32_t)i;
            sp.step_hi = 0;
            Layer& LL = e->layers[e->cfg.n_layers - 1];
            if ((rc = f32_decode_race(e, LL.h[LL.cur].as<float>(), LL.Hpad, bs, sp, gen + (int64_t)i * bs, st))) return rc;
            if ((rc = forward_impl(e, gen + (int64_t)i * bs, 1, bs, nullptr, nullptr, 1, st))) return rc;          // main.py:71
        }
    } else {
        if ((rc = e->logits_ws.ensure((size_t)bs * V * sizeof(float)))) return rc;
        float* z = e->logits_ws.as<float>();
        // prompt: one (P,bs) pass, decoder on the last step only (the reference decodes and discards
        // every prompt step, main.py:54-57; only the last result is ever read)
        if ((rc = forward_impl(e, e->tok.as<int64_t>(), P, bs, z, nullptr, 1, st))) return rc;
        e->cur_start = nullptr;                  // every stream is live from here on
        for (int i = 0; i < n_new; ++i) {
            sp.step_lo = (uint32_t)i;
            sp.step_hi = 0;
            { ProfScope ps(e, PROF_SAMPLE, st); rc = launch_sample_f32(z, V, V, bs, sp, gen + (int64_t)i * bs, nullptr, st); }   // main.py:61-68
            if (rc) return rc;
            if ((rc = forward_impl(e, gen + (int64_t)i * bs, 1, bs, z, nullptr, 1, st))) return rc;          // main.py:71
        }
    }
    if (n_new > 0)
        AWD_CUDA(cudaMemcpyAsync(out_tokens, gen, (size_t)n_new * bs * sizeof(int64_t), cudaMemcpyDefault, st));
    return AWD_OK;
}

// ---- continuous batching ------------------------------------------------------------
// rows of a [bs][words] uint32 array whose mask byte is set are cleared
__global__ void zero_rows_kernel(uint32_t* __restrict__ p, int words, const uint8_t* __restrict__ mask, int bs) {
    const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t r = gid / words;
    if (r < bs && mask[r]) p[gid] = 0u;
}
__global__ void check_tokens_kernel(const int64_t* __restrict__ tok, int64_t n, int V, int* err_flag) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && (tok[i] < 0 || tok[i] >= V)) atomicExch(err_flag, 1);
}
// fp32 path: ids drawn at this step are replaced by the stream's own next prompt token while it is inside its prompt
__global__ void apply_forced_kernel(int64_t* __restrict__ ids, const int64_t* __restrict__ forced, const int32_t* __restrict__ forced_len,
                                    int step, int bs) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < bs && step < forced_len[b]) ids[b] = forced[b];
}

int awd_generate_resume(awd_engine* e, const uint8_t* reset_mask, const int64_t* forced, const int32_t* forced_len,
                        int32_t F, int32_t bs, int32_t n_new, const awd_sampling* s, int64_t step_base,
                        int64_t* out_tokens, void* stream) {
    AWD_ENGINE_ENTER(e);
    if (!e->committed) return set_error(AWD_ESTATE, "weights not committed");
    if (bs < 1 || n_new < 0 || (n_new > 0 && !out_tokens) || F < 0 || (F > 0 && (!forced || !forced_len)))
        return set_error(AWD_EINVAL, "generate_resume: bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    const int V = e->cfg.vocab_sz;
    SampleParams sp;
    int rc = make_sample_params(s, V, sp);
    if (rc) return rc;
    // every call draws from its own part of the Philox stream (awd_generate's steps restart at 0 each call)
    sp.seed_lo ^= (uint32_t)((uint64_t)step_base * 0x9E3779B97F4A7C15ull >> 32);
    sp.seed_hi ^= (uint32_t)((uint64_t)step_base * 0xD1B54A32D192ED03ull >> 32);
    // pinned staging for this call's host arguments: [forced ids (if they are in pageable host memory)][forced_len][mask]
    bool forced_pageable = false;
    if (F > 0) {
        cudaPointerAttributes attr{};
        if (cudaPointerGetAttributes(&attr, forced) != cudaSuccess) { cudaGetLastError(); forced_pageable = true; }
        else forced_pageable = attr.type == cudaMemoryTypeUnregistered;
    }
    const size_t off_len = forced_pageable ? (size_t)F * bs * sizeof(int64_t) : 0;
    const size_t off_mask = off_len + (size_t)bs * sizeof(int32_t);
    const size_t stage_need = off_mask + (size_t)bs;
    awd_engine::PinnedStage& S = e->resume_stage[e->resume_stage_idx ^= 1];
    if ((rc = acquire_stage(S, stage_need))) return rc;
    uint8_t* stage = static_cast<uint8_t*>(S.p);
    if (bs != e->bs) {                          // first call (or a different number of slots): everything starts at reset
        if ((rc = reset_impl(e, bs, st))) return rc;
    } else if (reset_mask != nullptr) {
        bool any = false;
        for (int b = 0; b < bs; ++b) any |= reset_mask[b] != 0;
        if (any) {
            if ((rc = e->reset_mask_buf.ensure((size_t)bs))) return rc;
            memcpy(stage + off_mask, reset_mask, (size_t)bs);
            AWD_CUDA(cudaMemcpyAsync(e->reset_mask_buf.p, stage + off_mask, (size_t)bs, cudaMemcpyHostToDevice, st));
            const uint8_t* dm = e->reset_mask_buf.as<uint8_t>();
            auto zero = [&](void* p, int words) {
                const int64_t total = (int64_t)bs * words;
                zero_rows_kernel<<<(unsigned)cdiv(total, 256), 256, 0, st>>>((uint32_t*)p, words, dm, bs);
                count_launch();
            };
            for (auto& L : e->layers) {
                zero(L.h[0].p, L.Hpad); zero(L.h[1].p, L.Hpad); zero(L.c.p, L.Hpad);
                if (e->bf16()) { zero(L.hb[0].p, L.Hpad / 2); zero(L.hb[1].p, L.Hpad / 2); }
            }
            AWD_CUDA(cudaGetLastError());
            e->fold_ones = false;
        }
    }
    if ((rc = e->gen_out.ensure((size_t)std::max(n_new, 1) * bs * sizeof(int64_t)))) return rc;
    struct ForcedScope {
        awd_engine* e;
        ~ForcedScope() { e->cur_forced = nullptr; e->cur_forced_len = nullptr; e->cur_forced_F = 0; }
    } forced_scope{e};
    if (F > 0) {
        for (int b = 0; b < bs; ++b)
            if (forced_len[b] < 0 || forced_len[b] > F) return set_error(AWD_EINVAL, "forced_len[%d] = %d outside [0, %d]", b, forced_len[b], F);
        if ((rc = e->forced_buf.ensure((size_t)F * bs * sizeof(int64_t)))) return rc;
        if ((rc = e->forced_len_buf.ensure((size_t)bs * sizeof(int32_t)))) return rc;
        if (forced_pageable) {
            memcpy(stage, forced, (size_t)F * bs * sizeof(int64_t));
            AWD_CUDA(cudaMemcpyAsync(e->forced_buf.p, stage, (size_t)F * bs * sizeof(int64_t), cudaMemcpyHostToDevice, st));
        } else {
            AWD_CUDA(cudaMemcpyAsync(e->forced_buf.p, forced, (size_t)F * bs * sizeof(int64_t), cudaMemcpyDefault, st));
        }
        memcpy(stage + off_len, forced_len, (size_t)bs * sizeof(int32_t));
        AWD_CUDA(cudaMemcpyAsync(e->forced_len_buf.p, stage + off_len, (size_t)bs * sizeof(int32_t), cudaMemcpyHostToDevice, st));
        check_tokens_kernel<<<(unsigned)cdiv((int64_t)F * bs, 256), 256, 0, st>>>(e->forced_buf.as<int64_t>(), (int64_t)F * bs, V, e->err_flag.as<int>());
        AWD_LAUNCH_CHECK();
        e->cur_forced = e->forced_buf.as<int64_t>(); e->cur_forced_len = e->forced_len_buf.as<int32_t>(); e->cur_forced_F = F;
    }
    AWD_CUDA(cudaEventRecord(S.ev, st));                 // everything that reads the staging has been enqueued
    int64_t* gen = e->gen_out.as<int64_t>();
    // every step: decode the slot's current state -> draw (or take the forced prompt token) -> feed it
    if (e->bf16()) {
        if ((rc = bf16_generate_loop(e, bs, n_new, sp, gen, st))) return rc;      // (always the batched kernels, also for a few slots)
    } else if (f32_fused_race(bs)) {
        for (int i = 0; i < n_new; ++i) {
            sp.step_lo = (uint32_t)i;
            sp.step_hi = 0;
            Layer& LL = e->layers[e->cfg.n_layers - 1];      // the current states (reset slots decode zeros)
            if ((rc = f32_decode_race(e, LL.h[LL.cur].as<float>(), LL.Hpad, bs, sp, gen + (int64_t)i * bs, st))) return rc;
            if (i < F) {
                apply_forced_kernel<<<(unsigned)cdiv(bs, 256), 256, 0, st>>>(gen + (int64_t)i * bs, e->cur_forced + (int64_t)i * bs, e->cur_forced_len, i, bs);
                AWD_LAUNCH_CHECK();
            }
            if ((rc = forward_impl(e, gen + (int64_t)i * bs, 1, bs, nullptr, nullptr, 1, st))) return rc;
        }
    } else {
        if ((rc = e->logits_ws.ensure((size_t)bs * V * sizeof(float)))) return rc;
        float* z = e->logits_ws.as<float>();
        {   // logits of the current states (the previous call left them fed, not decoded; reset slots decode zeros)
            Layer& LL = e->layers[e->cfg.n_layers - 1];
            Operand d{};
            d.A = LL.h[LL.cur].as<float>(); d.lda = LL.Hpad; d.W = e->dec_weight(); d.ldw = e->Epad; d.K = e->Epad;
            EpiStore es{z, (int64_t)V, e->cfg.has_decoder_bias ? e->dec_b.as<float>() : nullptr};
            if ((rc = launch_gemm_store_f32(d, bs, V, es, st))) return rc;
        }
        for (int i = 0; i < n_new; ++i) {
            sp.step_lo = (uint32_t)i;
            sp.step_hi = 0;
            if ((rc = launch_sample_f32(z, V, V, bs, sp, gen + (int64_t)i * bs, nullptr, st))) return rc;
            if (i < F) {
                apply_forced_kernel<<<(unsigned)cdiv(bs, 256), 256, 0, st>>>(gen + (int64_t)i * bs, e->cur_forced + (int64_t)i * bs, e->cur_forced_len, i, bs);
                AWD_LAUNCH_CHECK();
            }
            if ((rc = forward_impl(e, gen + (int64_t)i * bs, 1, bs, z, nullptr, 1, st))) return rc;
        }
    }
    if (n_new > 0)
        AWD_CUDA(cudaMemcpyAsync(out_tokens, gen, (size_t)n_new * bs * sizeof(int64_t), cudaMemcpyDefault, st));
    return AWD_OK;
}

int awd_tc_gemm_bf16(const void* A, int64_t lda, const void* W, int64_t ldw, int32_t M, int32_t N, int32_t K,
                     const float* bias, float* C, int64_t ldc, void* stream) {
    if (!A || !W || !C || M < 1 || N < 1 || K < 64) return set_error(AWD_EINVAL, "bad arguments");
    if (K % 64 || lda % 64 || ldw % 64) return set_error(AWD_EINVAL, "K, lda and ldw must be multiples of 64");
    tc::TcOperands o{};
    o.a0 = A; o.a0_rows = (uint64_t)M; o.lda0 = (int)lda; o.k0 = K; o.a0_row = 0; o.k1 = 0;
    o.w = W; o.w_rows = (uint64_t)N; o.ldw = (int)ldw; o.w_cols = (uint64_t)K; o.b_k0 = 0;
    return tc::launch_tc_store(o, M, N, C, ldc, bias, (cudaStream_t)stream);
}

int awd_profile(awd_engine* e, int32_t enable) {
    AWD_ENGINE_ENTER(e);
    e->prof.on = enable != 0;
    e->prof.used = 0;
    e->prof.cats.clear();
    e->mega_steps = 0;
    if (e->mega_ns.p) AWD_CUDA(cudaMemset(e->mega_ns.p, 0, e->mega_ns.bytes));
    else if (enable) {
        int rc = e->mega_ns.ensure(16 * sizeof(unsigned long long));
        if (rc) return rc;
    }
    return AWD_OK;
}

int awd_profile_read(awd_engine* e, int32_t category, double* total_ms, int64_t* launches) {
    AWD_ENGINE_ENTER(e);
    if (category < 0 || category >= PROF_NCAT + 3 || !total_ms || !launches) return set_error(AWD_EINVAL, "bad arguments");
    AWD_CUDA(cudaDeviceSynchronize());
    double tot = 0; int64_t n = 0;
    for (size_t i = 0; i < e->prof.used; ++i) {
        if (e->prof.cats[i] != category) continue;
        float ms = 0.f;
        AWD_CUDA(cudaEventElapsedTime(&ms, e->prof.pool[2 * i], e->prof.pool[2 * i + 1]));
        tot += ms; ++n;
    }
    if (e->mega_steps > 0 && e->mega_ns.p) {
        // phases of the persistent generation kernel, timed in-kernel with %globaltimer by CTA 0:
        // [decoder, finalize, layer 0, layer 1, layers >= 2]
        unsigned long long ns[16] = {0};
        AWD_CUDA(cudaMemcpy(ns, e->mega_ns.p, sizeof(ns), cudaMemcpyDeviceToHost));
        if (category == PROF_DECODER) { tot += ns[0] * 1e-6; n += e->mega_steps; }
        if (category == PROF_SAMPLE) { tot += ns[1] * 1e-6; n += e->mega_steps; }
        if (category == PROF_STEP) { tot += (ns[2] + ns[3] + ns[4]) * 1e-6; n += e->mega_steps * e->cfg.n_layers; }
        if (category >= PROF_NCAT) { tot += ns[2 + category - PROF_NCAT] * 1e-6; n += e->mega_steps; }   // per-layer split
    }
    *total_ms = tot; *launches = n;
    return AWD_OK;
}

int awd_decode_race_shard(awd_engine* e, int32_t v0, int32_t v1, const awd_sampling* s, int64_t step, float* out_key,
                          int64_t* out_id, float* out_banned, void* stream) {
    AWD_ENGINE_ENTER(e);
    if (!e->committed) return set_error(AWD_ESTATE, "weights not committed");
    if (!e->bf16()) return set_error(AWD_ESTATE, "vocabulary-sharded decoding is built for bf16 engines");
    // whole decoder tiles: the race noise of a 32-id chunk is keyed by the first chunk its thread sees in a 256-wide tile
    if (v0 < 0 || v1 > e->cfg.vocab_sz || v0 >= v1 || (v0 % 256)) return set_error(AWD_EINVAL, "shard [%d,%d): need 0 <= v0 < v1 <= vocab_sz and v0 %% 256 == 0", v0, v1);
    if (!out_key || !out_id || !out_banned || e->bs < 1) return set_error(AWD_EINVAL, "bad arguments");
    SampleParams sp;
    int rc = make_sample_params(s, e->cfg.vocab_sz, sp);
    if (rc) return rc;
    sp.step_lo = (uint32_t)((uint64_t)step & 0xffffffffu);
    sp.step_hi = (uint32_t)((uint64_t)step >> 32);
    return bf16_decode_race_shard(e, v0, v1, sp, out_key, out_id, out_banned, (cudaStream_t)stream);
}

int awd_merge_race(const float* keys, const int64_t* ids, const float* banned, int32_t n_shards, int32_t bs, int32_t vocab_sz,
                   const awd_sampling* s, int64_t* out, void* stream) {
    if (!keys || !ids || !banned || !out || n_shards < 1 || bs < 1) return set_error(AWD_EINVAL, "bad arguments");
    SampleParams sp;
    int rc = make_sample_params(s, vocab_sz, sp);
    if (rc) return rc;
    return tc::launch_shard_merge(keys, ids, banned, n_shards, bs, sp, out, (cudaStream_t)stream);
}

__global__ void gumbel_probe_kernel(const uint32_t* __restrict__ bits, int n, float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = gumbel_from_bits(bits[i]);
}

float awd_uniform_from_bits(uint32_t bits) { return uniform_open01(bits); }

int awd_gumbel_from_bits(const uint32_t* bits, int32_t n, float* out, void* stream) {
    if (!bits || !out || n < 0) return set_error(AWD_EINVAL, "bad arguments");
    if (n == 0) return AWD_OK;
    gumbel_probe_kernel<<<(unsigned)cdiv(n, 256), 256, 0, (cudaStream_t)stream>>>(bits, n, out);
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

// One rank's block of the packed exchange buffer: ids int64 [bs] | keys fp32 [bs] | banned fp32 [bs][64]  | pad fp32 [bs] = 68 floats per stream (even: the next block stays 8-byte aligned)
int awd_decode_race_shard_packed(awd_engine* e, int32_t v0, int32_t v1, const awd_sampling* s, int64_t step, void* block, void* stream) {
    if (!e || !block) return set_error(AWD_EINVAL, "bad arguments");
    const int bs = e->bs;
    int64_t* ids = static_cast<int64_t*>(block);
    float* keys = reinterpret_cast<float*>(ids + bs);
    return awd_decode_race_shard(e, v0, v1, s, step, keys, ids, keys + bs, stream);
}

int awd_merge_race_packed(const void* blocks, int32_t n_shards, int32_t bs, int32_t vocab_sz, const awd_sampling* s, int64_t* out,
                          void* stream) {
    if (!blocks || !out || n_shards < 1 || bs < 1) return set_error(AWD_EINVAL, "bad arguments");
    SampleParams sp;
    int rc = make_sample_params(s, vocab_sz, sp);
    if (rc) return rc;
    const int64_t block_floats = (int64_t)bs * AWD_RACE_BLOCK_FLOATS_PER_STREAM;
    const int64_t* ids = static_cast<const int64_t*>(blocks);
    const float* keys = reinterpret_cast<const float*>(ids + bs);
    return tc::launch_shard_merge(keys, ids, keys + bs, n_shards, bs, sp, out, (cudaStream_t)stream, block_floats, block_floats / 2, block_floats);
}

int awd_check_device_errors(awd_engine* e, void* stream) {
    AWD_ENGINE_ENTER(e);
    int flag = 0;
    cudaStream_t st = (cudaStream_t)stream;
    AWD_CUDA(cudaMemcpyAsync(&flag, e->err_flag.p, sizeof(int), cudaMemcpyDeviceToHost, st));
    AWD_CUDA(cudaStreamSynchronize(st));
    if (flag) {
        AWD_CUDA(cudaMemsetAsync(e->err_flag.p, 0, sizeof(int), st));
        return set_error(AWD_ETOKEN, "index out of range in self: a token id is outside [0, %d)", e->cfg.vocab_sz);
    }
    return AWD_OK;
}

}  // extern "C"
