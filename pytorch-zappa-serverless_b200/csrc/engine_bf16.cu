// bf16 precision mode of the engine: every contraction runs on tcgen05 tensor cores (gemm_tc.cuh),
// bf16 weights/activations, fp32 accumulation in TMEM, fp32 cell state.  North-star tolerance for
// this mode: 2e-2 relative on logits and states.
//
// Per generated token (awd_generate): finalize+gather -> 3 x {[x;h]*[W_ih|W_hh]^T with the LSTM cell in
// the epilogue} -> decoder GEMM with the sampling race in the epilogue.  Logits never reach HBM.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include <cuda_bf16.h>

#include "engine.h"
#include "tc_api.h"

namespace awd {

using tc::TcCellArgs;
using tc::TcOperands;

__global__ void bf16_to_f32_rows_kernel(const __nv_bfloat16* __restrict__ src, int ld_src, int64_t rows, int cols,
                                        float* __restrict__ dst, int64_t ld_dst) {
    int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t r = gid / cols;
    int k = (int)(gid - r * cols);
    if (r >= rows) return;
    dst[r * ld_dst + k] = __bfloat162float(src[r * ld_src + k]);
}

// ---- decoder bias folded into the decoder GEMM (sampled decoding only)
// The vocabulary matrix is K-padded to a multiple of 64 (400 -> 448 in the shipped model), so the decoder GEMM
// multiplies 48 columns of zeros.  Two of them carry the bias instead: W'[v][F] = bf16(b[v]), W'[v][F+1] =
// bf16(b[v] - W'[v][F]) (error <= 2^-17 |b|), and the hidden operand holds 1.0 in those columns.  The sampling
// epilogue then needs no bias staging, no shared-memory reads and no epilogue-wide barrier per tile.
// F = round_up(E, 8), not E: the cell epilogue stores h in 16-byte groups of 8 units (TcEpiCell::chunk), so the
// columns up to the next multiple of 8 are rewritten (with zeros) on every LSTM step; from F on nothing writes.
__global__ void fold_bias_kernel(__nv_bfloat16* __restrict__ w, int ld, int V, int E, const float* __restrict__ bias) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= V) return;
    const float b = bias[v];
    const __nv_bfloat16 hi = __float2bfloat16_rn(b);
    w[(int64_t)v * ld + E] = hi;
    w[(int64_t)v * ld + E + 1] = __float2bfloat16_rn(b - __bfloat162float(hi));
}
__global__ void fold_ones_kernel(__nv_bfloat16* __restrict__ h0, __nv_bfloat16* __restrict__ h1, int ld, int bs, int E) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= bs) return;
    const __nv_bfloat16 one = __float2bfloat16_rn(1.f);
    h0[(int64_t)r * ld + E] = one; h0[(int64_t)r * ld + E + 1] = one;
    h1[(int64_t)r * ld + E] = one; h1[(int64_t)r * ld + E + 1] = one;
}

// Decoder operands of the sampled paths: (*w, *bias) = (folded matrix, nullptr) when folding applies, else the
// plain matrix and the fp32 bias.  Stream-ordered; call right before enqueueing the decoder.
static int ensure_fold(awd_engine* e, cudaStream_t st, const void** w, const float** bias) {
    const awd_config& c = e->cfg;
    static int use = -1;
    if (use < 0) { const char* s = getenv("AWD_FOLD_BIAS"); use = s ? atoi(s) : 1; }
    *w = e->dec_weight_bf();
    *bias = c.has_decoder_bias ? e->dec_b.as<float>() : nullptr;
    Layer& LL = e->layers[c.n_layers - 1];
    const int F = round_up(c.emb_sz, 8);          // first pad column no cell epilogue ever stores to
    if (!use || !c.has_decoder_bias || e->Epad - F < 2 || LL.Hpad != e->Epad) return AWD_OK;
    if (!e->fold_ready) {
        const size_t bytes = (size_t)c.vocab_sz * e->Epad * 2;
        int rc = e->dec_fold.ensure(bytes);
        if (rc) return rc;
        AWD_CUDA(cudaMemcpyAsync(e->dec_fold.p, e->dec_weight_bf(), bytes, cudaMemcpyDeviceToDevice, st));
        fold_bias_kernel<<<(unsigned)cdiv(c.vocab_sz, 256), 256, 0, st>>>(e->dec_fold.as<__nv_bfloat16>(), e->Epad, c.vocab_sz, F,
                                                                          e->dec_b.as<float>());
        AWD_LAUNCH_CHECK();
        e->fold_ready = true;
    }
    if (!e->fold_ones) {
        fold_ones_kernel<<<(unsigned)cdiv(e->bs, 256), 256, 0, st>>>(LL.hb[0].as<__nv_bfloat16>(), LL.hb[1].as<__nv_bfloat16>(), LL.Hpad,
                                                                     e->bs, F);
        AWD_LAUNCH_CHECK();
        e->fold_ones = true;
    }
    *w = e->dec_fold.p;
    *bias = nullptr;
    return AWD_OK;
}

int bf16_commit(awd_engine* e, const float* enc_src, const float* dec_src, const float* const* w_ih,
                const float* const* w_hh, cudaStream_t st) {
    const awd_config& c = e->cfg;
    e->fold_ready = false;
    int rc = e->enc_bf.ensure((size_t)c.vocab_sz * e->Epad * 2);
    if (rc) return rc;
    if ((rc = tc::launch_pack_rows_bf16(enc_src, c.vocab_sz, c.emb_sz, e->enc_bf.p, e->Epad, 0, e->Epad, 0, st))) return rc;
    if (!c.tie_weights) {
        if ((rc = e->dec_bf.ensure((size_t)c.vocab_sz * e->Epad * 2))) return rc;
        if ((rc = tc::launch_pack_rows_bf16(dec_src, c.vocab_sz, c.emb_sz, e->dec_bf.p, e->Epad, 0, e->Epad, 0, st))) return rc;
    }
    for (int l = 0; l < c.n_layers; ++l) {
        Layer& L = e->layers[l];
        const int kcat = L.Ipad + L.Hpad;
        if ((rc = L.wcat.ensure((size_t)4 * L.H * kcat * 2))) return rc;
        if ((rc = tc::launch_pack_rows_bf16(w_ih[l], 4 * L.H, L.I, L.wcat.p, kcat, 0, L.Ipad, L.H, st))) return rc;
        if ((rc = tc::launch_pack_rows_bf16(w_hh[l], 4 * L.H, L.H, L.wcat.p, kcat, L.Ipad, L.Hpad, L.H, st))) return rc;
    }
    return AWD_OK;
}

// set_state wrote fp32 h: refresh the bf16 operand copy.
int bf16_sync_hidden_operand(awd_engine* e, int layer, cudaStream_t st) {
    Layer& L = e->layers[layer];
    e->fold_ones = false;
    return tc::launch_f32_to_bf16_rows(L.h[0].as<float>(), L.Hpad, e->bs, L.H, L.hb[L.cur].p, L.Hpad, st);
}

static TcOperands weight_ops(const Layer& L) {
    TcOperands o{};
    o.w = L.wcat.p; o.w_rows = (uint64_t)4 * L.H; o.ldw = L.Ipad + L.Hpad; o.w_cols = (uint64_t)(L.Ipad + L.Hpad);
    return o;
}

// One timestep of every layer for bs streams.  x: bf16 [bs][Ipad of layer 0] (already gathered).
static int bf16_step(awd_engine* e, const void* x, uint64_t x_rows, int x_row0, int bs, float* const* raw_outputs,
                     int64_t raw_row0, int t_abs, cudaStream_t st) {
    const void* prev = x;
    uint64_t prev_rows = x_rows;
    int prev_row0 = x_row0;
    for (int l = 0; l < e->cfg.n_layers; ++l) {
        Layer& L = e->layers[l];
        TcOperands o = weight_ops(L);
        o.a0 = prev; o.a0_rows = prev_rows; o.lda0 = L.Ipad; o.k0 = L.Ipad; o.a0_row = prev_row0;
        o.a1 = L.hb[L.cur].p; o.a1_rows = (uint64_t)bs; o.lda1 = L.Hpad; o.k1 = L.Hpad; o.a1_row = 0;
        o.b_k0 = 0;
        TcCellArgs a{};
        a.addend = nullptr; a.bias = L.bias.as<float>();
        a.c = L.c.as<float>(); a.h = L.h[0].as<float>(); a.ld_state = L.Hpad;
        a.hb = L.hb[L.cur ^ 1].p; a.ld_hb = L.Hpad;
        a.raw_out = (raw_outputs && raw_outputs[l]) ? raw_outputs[l] + raw_row0 * L.H : nullptr;
        a.ld_raw = L.H;
        a.start = e->cur_start; a.t = t_abs;
        int rc;
        { ProfScope ps(e, PROF_STEP, st); rc = tc::launch_tc_cell(o, bs, 4 * L.H, a, st); }
        if (rc) return rc;
        L.cur ^= 1;
        prev = L.hb[L.cur].p; prev_rows = (uint64_t)bs; prev_row0 = 0;
    }
    return AWD_OK;
}

static TcOperands decoder_ops(awd_engine* e, const void* a, uint64_t a_rows, int a_row0) {
    TcOperands o{};
    o.a0 = a; o.a0_rows = a_rows; o.lda0 = e->Epad; o.k0 = e->Epad; o.a0_row = a_row0;
    o.w = e->dec_weight_bf(); o.w_rows = (uint64_t)e->cfg.vocab_sz; o.ldw = e->Epad; o.w_cols = (uint64_t)e->Epad;
    return o;
}

int bf16_forward(awd_engine* e, const int64_t* tok, int sl, int bs, float* logits, float* const* raw_outputs,
                 int last_step_only, cudaStream_t st) {
    const awd_config& c = e->cfg;
    const int V = c.vocab_sz, nl = c.n_layers;
    int* flag = e->err_flag.as<int>();
    e->fold_ones = false;            // the prompt paths below rewrite whole hidden-operand rows
    const int chunk_T = std::max(1, 32768 / bs);
    int maxHpad = 0, maxH = 0;
    for (auto& L : e->layers) { maxHpad = std::max(maxHpad, L.Hpad); maxH = std::max(maxH, L.H); }
    const size_t rows_cap = (size_t)std::min(sl, chunk_T) * bs;
    int rc = e->xb.ensure(rows_cap * e->Epad * 2);
    if (rc) return rc;
    if (sl > 1) {
        if ((rc = e->xproj.ensure(rows_cap * 4 * maxH * sizeof(float)))) return rc;
        for (int i = 0; i < 2; ++i)
            if ((rc = e->ybuf[i].ensure(rows_cap * maxHpad * 2))) return rc;
    }
    const float* dbias = c.has_decoder_bias ? e->dec_b.as<float>() : nullptr;
    static int use_mega_fwd = -1;
    if (use_mega_fwd < 0) { const char* s = getenv("AWD_MEGA_FWD"); use_mega_fwd = s ? atoi(s) : 1; }
    for (int t0 = 0; t0 < sl; t0 += chunk_T) {
        const int T = std::min(chunk_T, sl - t0);
        const int64_t R = (int64_t)T * bs;
        if ((rc = tc::launch_gather_bf16(e->enc_bf.p, e->Epad, V, tok + (int64_t)t0 * bs, R, e->xb.p, flag, st))) return rc;
        const void* last = nullptr;          // last layer's outputs for this chunk
        uint64_t last_rows = 0;
        if (T == 1) {
            if ((rc = bf16_step(e, e->xb.p, (uint64_t)bs, 0, bs, raw_outputs, (int64_t)t0 * bs, t0, st))) return rc;
            last = e->layers[nl - 1].hb[e->layers[nl - 1].cur].p;
            last_rows = (uint64_t)bs;
        } else if (use_mega_fwd && nl <= 4 && (int)cdiv(bs, 128) <= 148) {
            // input projection of layer 0 for the whole chunk, then every (layer, step) as a wavefront in
            // one persistent launch (mega_forward_kernel)
            for (int l = 0; l < nl; ++l)
                if ((rc = e->ybuf[l].ensure(rows_cap * e->layers[l].Hpad * 2))) return rc;
            if ((rc = e->mega_sync.ensure(sizeof(unsigned)))) return rc;
            Layer& L0 = e->layers[0];
            float* xp = e->xproj.as<float>();
            TcOperands o = weight_ops(L0);
            o.a0 = e->xb.p; o.a0_rows = (uint64_t)R; o.lda0 = L0.Ipad; o.k0 = L0.Ipad; o.a0_row = 0; o.k1 = 0; o.b_k0 = 0;
            { ProfScope ps(e, PROF_XPROJ, st); rc = tc::launch_tc_store(o, (int)R, 4 * L0.H, xp, (int64_t)4 * L0.H, L0.bias.as<float>(), st); }
            if (rc) return rc;
            tc::FwdArgs f{};
            f.n_layers = nl; f.bs = bs; f.T = T; f.t_base = t0;
            for (int l = 0; l < nl; ++l) {
                Layer& L = e->layers[l];
                f.H[l] = L.H; f.Ipad[l] = L.Ipad; f.Hpad[l] = L.Hpad;
                f.wcat[l] = L.wcat.p; f.bias[l] = L.bias.as<float>(); f.c[l] = L.c.as<float>(); f.h[l] = L.h[0].as<float>();
                f.hb[l] = L.hb[L.cur].p; f.y[l] = e->ybuf[l].p;
                f.raw[l] = (raw_outputs && raw_outputs[l]) ? raw_outputs[l] + (int64_t)t0 * bs * L.H : nullptr;
            }
            f.xproj0 = xp; f.start = e->cur_start; f.sync_counter = e->mega_sync.as<unsigned>();
            { ProfScope ps(e, PROF_STEP, st); rc = tc::launch_mega_forward(f, st); }
            if (rc) return rc;
            for (int l = 0; l < nl; ++l) {
                Layer& L = e->layers[l];
                AWD_CUDA(cudaMemcpyAsync(L.hb[L.cur].p, e->ybuf[l].as<__nv_bfloat16>() + (int64_t)(T - 1) * bs * L.Hpad,
                                         (size_t)bs * L.Hpad * 2, cudaMemcpyDeviceToDevice, st));
            }
            last = e->ybuf[nl - 1].p;
            last_rows = (uint64_t)R;
        } else {
            const void* prev = e->xb.p;
            for (int l = 0; l < nl; ++l) {
                Layer& L = e->layers[l];
                float* xp = e->xproj.as<float>();
                __nv_bfloat16* Y = e->ybuf[l & 1].as<__nv_bfloat16>();
                TcOperands o = weight_ops(L);
                o.a0 = prev; o.a0_rows = (uint64_t)R; o.lda0 = L.Ipad; o.k0 = L.Ipad; o.a0_row = 0; o.k1 = 0; o.b_k0 = 0;
                { ProfScope ps(e, PROF_XPROJ, st); rc = tc::launch_tc_store(o, (int)R, 4 * L.H, xp, (int64_t)4 * L.H, L.bias.as<float>(), st); }
                if (rc) return rc;
                float* raw = (raw_outputs && raw_outputs[l]) ? raw_outputs[l] + (int64_t)t0 * bs * L.H : nullptr;
                for (int t = 0; t < T; ++t) {
                    TcOperands h = weight_ops(L);
                    if (t == 0) { h.a0 = L.hb[L.cur].p; h.a0_rows = (uint64_t)bs; h.a0_row = 0; }
                    else { h.a0 = Y; h.a0_rows = (uint64_t)R; h.a0_row = (t - 1) * bs; }
                    h.lda0 = L.Hpad; h.k0 = L.Hpad; h.k1 = 0; h.b_k0 = L.Ipad;
                    TcCellArgs a{};
                    a.addend = xp + (int64_t)t * bs * 4 * L.H; a.ld_add = 4 * L.H; a.bias = nullptr;
                    a.c = L.c.as<float>(); a.h = L.h[0].as<float>(); a.ld_state = L.Hpad;
                    a.hb = Y + (int64_t)t * bs * L.Hpad; a.ld_hb = L.Hpad;
                    a.raw_out = raw ? raw + (int64_t)t * bs * L.H : nullptr; a.ld_raw = L.H;
                    a.start = e->cur_start; a.t = t0 + t;
                    { ProfScope ps(e, PROF_STEP, st); rc = tc::launch_tc_cell(h, bs, 4 * L.H, a, st); }
                    if (rc) return rc;
                }
                AWD_CUDA(cudaMemcpyAsync(L.hb[L.cur].p, Y + (int64_t)(T - 1) * bs * L.Hpad, (size_t)bs * L.Hpad * 2,
                                         cudaMemcpyDeviceToDevice, st));
                prev = Y;
            }
            last = prev;
            last_rows = (uint64_t)R;
        }
        if (logits != nullptr) {
            if (!last_step_only) {
                TcOperands d = decoder_ops(e, last, last_rows, 0);
                { ProfScope ps(e, PROF_DECODER, st); rc = tc::launch_tc_store(d, (int)R, V, logits + (int64_t)t0 * bs * V, (int64_t)V, dbias, st); }
                if (rc) return rc;
            } else if (t0 + T == sl) {
                TcOperands d = decoder_ops(e, last, last_rows, (T - 1) * bs);
                { ProfScope ps(e, PROF_DECODER, st); rc = tc::launch_tc_store(d, bs, V, logits, (int64_t)V, dbias, st); }
                if (rc) return rc;
            }
        }
    }
    return AWD_OK;
}

// The n_new iterations of sample_model's loop (main.py:60-71) after the prompt has been fed.
int bf16_generate_loop(awd_engine* e, int bs, int n_new, SampleParams sp, int64_t* gen, cudaStream_t st) {
    const awd_config& c = e->cfg;
    const int V = c.vocab_sz, nl = c.n_layers;
    const int slots = tc::tc_sample_slots(bs, V), slots_max = tc::tc_mega_slots(bs, V);
    int rc = e->part.ensure((size_t)bs * slots_max * 2 * sizeof(float));
    if (rc) return rc;
    if ((rc = e->zwin.ensure((size_t)bs * slots_max * 32 * sizeof(float)))) return rc;
    if ((rc = e->banned_keys.ensure((size_t)bs * tc::kBannedKeysFloatsPerRow * sizeof(float))   /* keys of the disallowed ids + one row bound per stream */)) return rc;
    if ((rc = e->xb.ensure((size_t)bs * e->Epad * 2))) return rc;
    const void* dec_w = nullptr;
    const float* dbias = nullptr;
    if ((rc = ensure_fold(e, st, &dec_w, &dbias))) return rc;
    static int use_mega = -1;
    if (use_mega < 0) { const char* s = getenv("AWD_MEGA"); use_mega = s ? atoi(s) : 1; }
    if (use_mega && n_new > 0 && nl <= 4 && e->layers[0].cur == e->layers[nl - 1].cur) {
        // the whole token loop in one persistent kernel
        if ((rc = e->mega_sync.ensure(sizeof(unsigned)))) return rc;
        if ((rc = e->mega_ns.ensure(16 * sizeof(unsigned long long)))) return rc;
        tc::MegaArgs m{};
        m.n_layers = nl; m.bs = bs; m.V = V; m.n_steps = n_new; m.p0 = e->layers[0].cur; m.Epad = e->Epad;
        for (int l = 0; l < nl; ++l) {
            Layer& L = e->layers[l];
            m.H[l] = L.H; m.Ipad[l] = L.Ipad; m.Hpad[l] = L.Hpad;
            m.wcat[l] = L.wcat.p; m.bias[l] = L.bias.as<float>(); m.c[l] = L.c.as<float>(); m.h[l] = L.h[0].as<float>();
            if ((rc = L.c_blk.ensure((size_t)bs * L.Hpad * sizeof(float)))) return rc;
            m.c_blk[l] = L.c_blk.as<float>();
            m.hb[l][0] = L.hb[0].p; m.hb[l][1] = L.hb[1].p;
        }
        m.dec_w = dec_w; m.dec_bias = dbias; m.sp = sp;
        m.part = e->part.as<float>(); m.zwin = e->zwin.as<float>(); m.banned_keys = e->banned_keys.as<float>();
        m.gen = gen; m.table = e->enc_bf.p; m.xb = e->xb.p;
        m.forced = e->cur_forced; m.forced_len = e->cur_forced_len; m.forced_F = e->cur_forced_F;
        m.sync_counter = e->mega_sync.as<unsigned>();
        m.phase_ns = e->prof.on ? e->mega_ns.as<unsigned long long>() : nullptr;
        if (bs <= 1024) {
            if ((rc = e->sb_scratch.ensure(tc::sb_scratch_bytes(nl, m.H, bs)))) return rc;
            if ((rc = e->sb_counter.ensure(tc::kSbCounterBytes))) return rc;
            m.sb_scratch = e->sb_scratch.p; m.sb_counter = e->sb_counter.as<unsigned>();
        }
        static const bool trace = getenv("AWD_MEGA_TRACE") != nullptr;
        static long long* trace_buf = nullptr;
        if (trace) {
            if (!trace_buf) AWD_CUDA(cudaMalloc(&trace_buf, (128 + 148 * 8 + 64) * sizeof(long long)));
            AWD_CUDA(cudaMemsetAsync(trace_buf, 0, (128 + 148 * 8 + 64) * sizeof(long long), st));
            m.trace = trace_buf;
        }
        if ((rc = tc::launch_mega_generate(m, st))) return rc;
        if (trace && n_new > 5) {
            static long long h[128 + 148 * 8 + 64];
            AWD_CUDA(cudaMemcpyAsync(h, trace_buf, sizeof(h), cudaMemcpyDeviceToHost, st));
            AWD_CUDA(cudaStreamSynchronize(st));
            for (int c = 0; c < 2; ++c) {
                const long long* t = h + 64 * c;
                for (int l = 0; l < nl; ++l) {
                    const long long* u = t + 16 * l;
                    fprintf(stderr, "[awd trace] cta %d layer %d (clocks from phase entry): first operands %lld, first input-half operands %lld, MMAs issued %lld | "
                            "epilogue waits from %lld, accumulator ready %lld, partials stored %lld, all slices arrived %lld, cell done %lld\n", c ? 77 : 0, l,
                            u[2] - t[0], u[11] - t[0], u[3] - t[0], u[4] - t[0], u[5] - t[0], u[6] - t[0], u[7] - t[0], u[8] - t[0]);
                }
                fprintf(stderr, "[awd trace] cta %d: thread 0 at grid barrier %lld, released %lld | finalize starts %lld, row resolved and x written %lld, released %lld; producer sees all rows %lld\n",
                        c ? 77 : 0, t[9] - t[0], t[10] - t[0], t[12] - t[0], t[13] - t[0], t[14] - t[0], t[15] - t[0]);
            }
            for (int c = 0; c < 2; ++c) {
                const long long* d = h + 48 + 64 * c;
                fprintf(stderr, "[awd trace] cta 0/77 decoder phase (clocks from entry): first operands tile0 %lld tile1 %lld, MMAs issued tile0 %lld tile1 %lld, epilogue done tile0 %lld tile1 %lld, "
                        "thread 0 at grid barrier %lld, released %lld | epilogue warp 0, tile 0: tile_begin done %lld, accumulator ready %lld, first tcgen05.ld back %lld, chunks done %lld %lld %lld %lld\n", d[1] - d[0], d[2] - d[0], d[3] - d[0], d[4] - d[0], d[5] - d[0], d[6] - d[0], d[7] - d[0], d[8] - d[0],
                        d[9] - d[0], d[10] - d[0], d[15] - d[0], d[11] - d[0], d[12] - d[0], d[13] - d[0], d[14] - d[0]);
            }
            {   // CTA 8 (a group with the average number of tiles), decoder phase, tile by tile, relative to its first stamp
                const long long* q = h + 128 + 148 * 8;
                for (int t = 0; t < 16 && q[t * 4 + 3]; ++t)
                    fprintf(stderr, "[awd trace] cta 8 decoder tile %d: first operands %lld, MMAs issued %lld, epilogue sees accumulator %lld, epilogue done %lld\n", t,
                            q[t * 4] - q[0], q[t * 4 + 1] - q[0], q[t * 4 + 2] - q[0], q[t * 4 + 3] - q[0]);
            }
            {   // every CTA's global-timer stamps, relative to the earliest phase entry
                const long long* g = h + 128;
                long long t0 = g[0];
                for (int c = 0; c < 148; ++c) if (g[c * 8] && g[c * 8] < t0) t0 = g[c * 8];
                for (int k = 0; k < 8; ++k) {
                    std::vector<long long> v;
                    for (int c = 0; c < 148; ++c) if (g[c * 8 + k]) v.push_back(g[c * 8 + k] - t0);
                    if (v.empty()) continue;
                    std::sort(v.begin(), v.end());
                    fprintf(stderr, "[awd trace] all CTAs, %s %d: n %zu, min %lld, p25 %lld, median %lld, p75 %lld, max %lld ns\n",
                            k == 0 ? "phase entry" : (k < 5 ? "cell done, layer" : "first input-half operands, layer"), k == 0 ? 0 : (k < 5 ? k - 1 : k - 5),
                            v.size(), v[0], v[v.size() / 4], v[v.size() / 2], v[v.size() * 3 / 4], v.back());
                }
            }
        }
        if (e->prof.on) e->mega_steps += n_new;
        if (n_new & 1)
            for (auto& L : e->layers) L.cur ^= 1;
        // the persistent kernel keeps h only as the bf16 operand; refresh the fp32 copy get_state returns
        for (auto& L : e->layers) {
            const int64_t total = (int64_t)bs * L.H;     // real columns only: the pad may hold the folded-bias ones
            bf16_to_f32_rows_kernel<<<(unsigned)cdiv(total, 256), 256, 0, st>>>(L.hb[L.cur].as<__nv_bfloat16>(), L.Hpad, bs, L.H,
                                                                                L.h[0].as<float>(), L.Hpad);
            AWD_LAUNCH_CHECK();
        }
        return AWD_OK;
    }
    for (int i = 0; i < n_new; ++i) {
        sp.step_lo = (uint32_t)i;
        sp.step_hi = 0;
        Layer& LL = e->layers[nl - 1];
        TcOperands d = decoder_ops(e, LL.hb[LL.cur].p, (uint64_t)bs, 0);
        d.w = dec_w;
        { ProfScope ps(e, PROF_DECODER, st); rc = tc::launch_tc_sample(d, bs, V, sp, dbias, e->part.as<float>(), e->zwin.as<float>(), e->banned_keys.as<float>(), st); }
        if (rc) return rc;
        { ProfScope ps(e, PROF_SAMPLE, st);
          rc = tc::launch_sample_finalize(e->part.as<float>(), e->zwin.as<float>(), e->banned_keys.as<float>(), slots, bs, V, sp, gen + (int64_t)i * bs,
                                          e->enc_bf.p, e->Epad, e->xb.p, st, (e->cur_forced && i < e->cur_forced_F) ? e->cur_forced : nullptr,
                                          e->cur_forced_len, i); }
        if (rc) return rc;
        if ((rc = bf16_step(e, e->xb.p, (uint64_t)bs, 0, bs, nullptr, 0, 1 << 30, st))) return rc;    // main.py:71
    }
    return AWD_OK;
}

// Vocabulary-sharded decoding: decoder + race over rows [v0, v1) for the current state.
int bf16_decode_race_shard(awd_engine* e, int v0, int v1, SampleParams sp, float* out_key, int64_t* out_id, float* out_banned,
                           cudaStream_t st) {
    const awd_config& c = e->cfg;
    const int bs = e->bs, n_local = v1 - v0;
    Layer& LL = e->layers[c.n_layers - 1];
    const int slots = tc::tc_sample_slots(bs, n_local);
    int rc = e->part.ensure((size_t)bs * slots * 2 * sizeof(float));
    if (rc) return rc;
    if ((rc = e->zwin.ensure((size_t)bs * slots * 32 * sizeof(float)))) return rc;
    if ((rc = e->banned_keys.ensure((size_t)bs * tc::kBannedKeysFloatsPerRow * sizeof(float))   /* keys of the disallowed ids + one row bound per stream */)) return rc;
    TcOperands d = decoder_ops(e, LL.hb[LL.cur].p, (uint64_t)bs, 0);
    const void* dec_w = nullptr;
    const float* bias = nullptr;
    if ((rc = ensure_fold(e, st, &dec_w, &bias))) return rc;
    d.w = (const char*)dec_w + (size_t)v0 * e->Epad * 2;
    d.w_rows = (uint64_t)n_local;
    if (bias) bias += v0;
    { ProfScope ps(e, PROF_DECODER, st);
      rc = tc::launch_tc_sample_shard(d, bs, n_local, v0, sp, bias, e->part.as<float>(), e->zwin.as<float>(), e->banned_keys.as<float>(), st); }
    if (rc) return rc;
    return tc::launch_shard_partial(e->part.as<float>(), e->zwin.as<float>(), e->banned_keys.as<float>(), slots, bs, v0, v1, sp,
                                    out_key, out_id, out_banned, st);
}

bool bf16_use_few_streams_kernel(const awd_engine* e, int bs) {
    static int use = -1;
    if (use < 0) { const char* s = getenv("AWD_LAT"); use = s ? atoi(s) : 1; }
    return use && bs <= tc::lat_max_streams() && e->cfg.n_layers <= 4;
}

// awd_generate for a handful of streams: prompt (already in e->tok) and all new tokens in one
// persistent launch (lat_kernel.cu).  State must have been reset by the caller.
int bf16_generate_few_streams(awd_engine* e, int P, int bs, int n_new, SampleParams sp, int64_t* gen, cudaStream_t st) {
    const awd_config& c = e->cfg;
    const int nl = c.n_layers;
    int rc = e->part.ensure((size_t)256 * 8 * 4 * sizeof(float));
    if (rc) return rc;
    if ((rc = e->banned_keys.ensure((size_t)8 * 64 * sizeof(float)))) return rc;
    if ((rc = e->mega_sync.ensure(sizeof(unsigned)))) return rc;
    tc::LatArgs a{};
    a.n_layers = nl; a.bs = bs; a.V = c.vocab_sz; a.P = P; a.n_new = n_new; a.Epad = e->Epad; a.par0 = e->layers[0].cur;
    for (int l = 0; l < nl; ++l) {
        Layer& L = e->layers[l];
        if (L.cur != a.par0) return set_error(AWD_ESTATE, "layer parities diverged");
        a.H[l] = L.H; a.Ipad[l] = L.Ipad; a.Hpad[l] = L.Hpad;
        a.wcat[l] = L.wcat.p; a.bias[l] = L.bias.as<float>(); a.c[l] = L.c.as<float>();
        a.hb[l][0] = L.hb[0].p; a.hb[l][1] = L.hb[1].p;
    }
    a.table = e->enc_bf.p; a.dec_w = e->dec_weight_bf(); a.dec_bias = c.has_decoder_bias ? e->dec_b.as<float>() : nullptr;
    a.prompt = e->tok.as<int64_t>(); a.gen = gen; a.sp = sp;
    a.part = e->part.as<float>(); a.banned_keys = e->banned_keys.as<float>();
    a.sync_counter = e->mega_sync.as<unsigned>(); a.err_flag = e->err_flag.as<int>();
    a.start = e->cur_start;
    if ((rc = tc::launch_lat_generate(a, st))) return rc;
    if ((P + n_new) & 1)
        for (auto& L : e->layers) L.cur ^= 1;
    for (auto& L : e->layers) {      // fp32 copy of h for get_state
        const int64_t total = (int64_t)bs * L.H;
        bf16_to_f32_rows_kernel<<<(unsigned)cdiv(total, 256), 256, 0, st>>>(L.hb[L.cur].as<__nv_bfloat16>(), L.Hpad, bs, L.H,
                                                                            L.h[0].as<float>(), L.Hpad);
        AWD_LAUNCH_CHECK();
    }
    return AWD_OK;
}

int bf16_decode(awd_engine* e, const float* x, int64_t rows, float* logits, cudaStream_t st) {
    const awd_config& c = e->cfg;
    int rc = e->ybuf[0].ensure((size_t)rows * e->Epad * 2);
    if (rc) return rc;
    if ((rc = tc::launch_f32_to_bf16_rows(x, c.emb_sz, (int)rows, c.emb_sz, e->ybuf[0].p, e->Epad, st))) return rc;
    TcOperands d = decoder_ops(e, e->ybuf[0].p, (uint64_t)rows, 0);
    ProfScope ps(e, PROF_DECODER, st);
    return tc::launch_tc_store(d, (int)rows, c.vocab_sz, logits, (int64_t)c.vocab_sz,
                               c.has_decoder_bias ? e->dec_b.as<float>() : nullptr, st);
}

int bf16_gather(awd_engine* e, const int64_t* tok, int64_t n, float* out, cudaStream_t st) {
    const awd_config& c = e->cfg;
    int rc = e->xb.ensure((size_t)std::max<int64_t>(n, 1) * e->Epad * 2);
    if (rc) return rc;
    if ((rc = tc::launch_gather_bf16(e->enc_bf.p, e->Epad, c.vocab_sz, tok, n, e->xb.p, e->err_flag.as<int>(), st))) return rc;
    int64_t total = n * c.emb_sz;
    if (total == 0) return AWD_OK;
    bf16_to_f32_rows_kernel<<<(unsigned)cdiv(total, 256), 256, 0, st>>>(e->xb.as<__nv_bfloat16>(), e->Epad, n, c.emb_sz, out,
                                                                        c.emb_sz);
    AWD_LAUNCH_CHECK();
    return AWD_OK;
}

}  // namespace awd
