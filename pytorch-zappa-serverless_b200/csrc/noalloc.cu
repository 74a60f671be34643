// awd_decode_sample_static: decoder + sampling rule for callers that must not allocate.
//
// Who needs this: a serving loop that has handed the whole GPU to its own memory arena (a
// cudaMalloc in the middle of it can fail, or fragment the arena), a latency-critical decode
// thread that cannot afford the device-wide synchronisation that cudaMalloc/cudaFree imply, and
// hosts that keep the token loop inside a region where the CUDA allocator is off limits.  The
// regular entry points (awd_generate, awd_decode on odd row pitches) grow engine workspaces on
// demand; this one never does.  The (bs, vocab_sz) logits that sit between the decoder GEMM and
// the sampler live in one fixed array inside the library image instead.
//
// Cost and limits: AWD_STATIC_SCRATCH_BYTES (16 MiB) of device memory on every GPU that loads
// the library, held until the process exits; bs * vocab_sz must fit in it (69 streams at the
// shipped 60k vocabulary).  All calls on a GPU share that one array and it keeps whatever the
// previous call left there, so two calls that overlap on the GPU (different streams, or another
// host thread) race on it; serialise them.  Nothing in it has to start at zero: each call
// overwrites every logit row it later reads, and rows beyond bs are never read.
//
// The first call on each GPU looks up the array's device address (cudaGetSymbolAddress); make
// that first call outside any stream capture.
#include <mutex>

#include "engine.h"

namespace {

constexpr size_t kScratchBytes = 16u << 20;   // AWD_STATIC_SCRATCH_BYTES in include/awd_b200.h
constexpr int kMaxDevices = 64;

// Device-resident scratch.  In host code this name is only a handle for the runtime; the device
// address is obtained per GPU below and that pointer is what the kernels receive.
static __device__ __align__(256) unsigned char g_static_scratch[kScratchBytes];

std::mutex g_setup_mutex;
float* g_scratch_ptr[kMaxDevices] = {};

// Once per device, with that device current.
int static_scratch_for(int device, float** out) {
    if (device < 0 || device >= kMaxDevices)
        return awd::set_error(AWD_ELIMIT, "device ordinal %d beyond the %d this entry point tracks", device, kMaxDevices);
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (g_scratch_ptr[device] == nullptr) {
        void* p = nullptr;
        cudaError_t err = cudaGetSymbolAddress(&p, g_static_scratch);
        if (err != cudaSuccess)
            return awd::set_error(AWD_ECUDA, "cudaGetSymbolAddress(static scratch) on device %d: %s", device,
                                  cudaGetErrorString(err));
        g_scratch_ptr[device] = reinterpret_cast<float*>(p);
    }
    *out = g_scratch_ptr[device];
    return AWD_OK;
}

// Same normalisation as the regular sampler entry (sorted, unique, id 0 first when banned),
// done in fixed storage.
int fill_sample_params(const awd_sampling* s, int V, int64_t step, awd::SampleParams& sp) {
    if (!s) return awd::set_error(AWD_EINVAL, "sampling parameters missing");
    if (!(s->temperature > 0.f)) return awd::set_error(AWD_EINVAL, "temperature must be > 0");
    if (s->n_draws < 1) return awd::set_error(AWD_EINVAL, "n_draws must be >= 1");
    if (s->n_banned < 0 || s->n_banned > 63)
        return awd::set_error(AWD_ELIMIT, "at most 63 banned ids (got %d)", s->n_banned);
    int ids[64];
    int n = 0;
    if (s->ban_zero) ids[n++] = 0;
    for (int i = 0; i < s->n_banned; ++i) {
        int64_t id = s->banned[i];
        if (id < 0 || id >= V) return awd::set_error(AWD_EINVAL, "banned id %lld outside the vocabulary", (long long)id);
        int v = (int)id, pos = 0;
        while (pos < n && ids[pos] < v) ++pos;
        if (pos < n && ids[pos] == v) continue;
        for (int j = n; j > pos; --j) ids[j] = ids[j - 1];
        ids[pos] = v;
        ++n;
    }
    sp = awd::SampleParams{};
    sp.inv_temperature = 1.0f / s->temperature;
    sp.greedy = s->greedy ? 1 : 0;
    sp.n_draws = s->n_draws;
    sp.n_banned = n;
    for (int i = 0; i < n; ++i) sp.banned[i] = ids[i];
    sp.seed_lo = (uint32_t)(s->seed & 0xffffffffu);
    sp.seed_hi = (uint32_t)(s->seed >> 32);
    sp.step_lo = (uint32_t)((uint64_t)step & 0xffffffffu);
    sp.step_hi = (uint32_t)((uint64_t)step >> 32);
    return AWD_OK;
}

}  // namespace

extern "C" int awd_decode_sample_static(awd_engine* e, const float* x, int32_t bs, const awd_sampling* s,
                                        int64_t step, int64_t* out, void* stream) {
    using namespace awd;
    if (!e) return set_error(AWD_EINVAL, "null engine");
    if (!e->committed) return set_error(AWD_ESTATE, "weights not committed");
    if (!x || !out || bs < 1) return set_error(AWD_EINVAL, "bad arguments");
    const awd_config& c = e->cfg;
    if (c.precision != AWD_PREC_FP32)
        return set_error(AWD_ESTATE, "awd_decode_sample_static serves fp32 engines only");
    const size_t need = (size_t)bs * (size_t)c.vocab_sz * sizeof(float);
    if (need > kScratchBytes)
        return set_error(AWD_ELIMIT,
                         "awd_decode_sample_static: bs*vocab_sz = %d*%d floats (%zu bytes) exceeds the fixed %zu-byte "
                         "scratch; use at most %zu streams at this vocabulary, or awd_decode + awd_sample",
                         bs, c.vocab_sz, need, kScratchBytes, kScratchBytes / ((size_t)c.vocab_sz * sizeof(float)));
    if (c.emb_sz != e->Epad || (((uintptr_t)x) & 15) != 0)
        return set_error(AWD_EINVAL,
                         "awd_decode_sample_static needs emb_sz to be a multiple of 4 and x 16-byte aligned "
                         "(re-pitching the rows would need a workspace)");
    SampleParams sp;
    int rc = fill_sample_params(s, c.vocab_sz, step, sp);
    if (rc) return rc;

    int prev = -1;
    if (cudaGetDevice(&prev) != cudaSuccess) return set_error(AWD_ECUDA, "cudaGetDevice failed");
    if (prev != c.device && cudaSetDevice(c.device) != cudaSuccess)
        return set_error(AWD_ECUDA, "cannot select device %d", c.device);
    float* logits = nullptr;
    rc = static_scratch_for(c.device, &logits);
    if (rc == AWD_OK) {
        cudaStream_t st = (cudaStream_t)stream;
        Operand d{};
        d.A = x; d.lda = c.emb_sz; d.W = e->dec_weight(); d.ldw = e->Epad; d.K = e->Epad;
        EpiStore es{logits, (int64_t)c.vocab_sz, c.has_decoder_bias ? e->dec_b.as<float>() : nullptr};
        rc = launch_gemm_store_f32(d, bs, c.vocab_sz, es, st);
        if (rc == AWD_OK) rc = launch_sample_f32(logits, c.vocab_sz, c.vocab_sz, bs, sp, out, nullptr, st);
    }
    if (prev != c.device) cudaSetDevice(prev);
    return rc;
}
