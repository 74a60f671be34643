/*
 * awd_b200.h -- C ABI of the B200-native AWD-LSTM inference engine.
 *
 * This is the drop-in boundary for the one hot path of gdoteof/pytorch-zappa-serverless:
 * AWD-LSTM forward and autoregressive generation.  The reference has no FFI (it is pure
 * Python calling PyTorch CPU operators); each entry point below replaces the Python call
 * site cited beside it (paths are inside the reference checkout).  A Python host binds these
 * with ctypes (pytorch-zappa-serverless_b200/awd_b200/_lib.py); see INTEGRATION.md for the
 * stub a maintainer of the reference would add.
 *
 * Conventions
 *   - every function returns 0 on success or a negative AWD_E* code; awd_last_error() gives the
 *     message for the calling thread;
 *   - plain pointers and sizes only; `stream` is a cudaStream_t passed as void* (NULL = default);
 *   - pointers marked [dev] must be device memory of the engine's GPU; pointers marked [any] may
 *     be host (pinned makes the copy asynchronous) or device memory -- they are copied with
 *     cudaMemcpyDefault; nothing is returned through host memory without the caller asking;
 *   - an engine is owned by one host thread at a time (like a reference module instance, whose
 *     `hidden`/`bs` are mutable state: pytorch_models/awd_lstm.py:77-103);
 *   - there is NO CPU fallback: awd_create fails if no sm_100 device is present.
 */
#ifndef AWD_B200_H
#define AWD_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AWD_OK            0
#define AWD_EINVAL       -1   /* bad argument / shape / key                    */
#define AWD_ECUDA        -2   /* CUDA runtime or driver error                  */
#define AWD_ENODEV       -3   /* no usable sm_100 GPU                          */
#define AWD_ESTATE       -4   /* call order (e.g. forward before commit)       */
#define AWD_ENOMEM       -5   /* device allocation failed                      */
#define AWD_EMISSING     -6   /* strict load: a state_dict key was not given   */
#define AWD_ETOKEN       -7   /* token id outside [0, vocab_sz)                */
#define AWD_ELIMIT       -8   /* size beyond a documented fixed limit          */

#define AWD_PREC_FP32     0   /* fp32 weights, fp32 FMA arithmetic (1e-4 parity mode)           */
#define AWD_PREC_BF16     1   /* bf16 weights/activations, fp32 accumulate + fp32 cell state     */

typedef struct awd_engine awd_engine;

/* Arguments of get_language_model that shape the arithmetic (pytorch_models/awd_lstm.py:7-14). */
typedef struct awd_config {
    int32_t vocab_sz;
    int32_t emb_sz;
    int32_t n_hid;
    int32_t n_layers;
    int32_t tie_weights;       /* decoder.weight is encoder.weight (awd_lstm.py:40)             */
    int32_t has_decoder_bias;  /* bias=True (awd_lstm.py:36)                                    */
    int32_t precision;         /* AWD_PREC_*                                                    */
    int32_t device;            /* CUDA ordinal                                                  */
} awd_config;

/* Sampling rule of sample_model (main.py:61-68). */
typedef struct awd_sampling {
    float    temperature;      /* reference: 1.0 (multinomial of exp(logits))                   */
    int32_t  greedy;           /* 1: candidates = the n_draws largest logits     */
    int32_t  n_draws;          /* reference: 10 draws without replacement (main.py:61)          */
    int32_t  ban_zero;         /* reference: 1 (`r != 0`, main.py:64)                           */
    int32_t  n_banned;         /* ids in `banned` (exclude_tokens, main.py:47); <= 63           */
    int32_t  _pad;
    const int64_t* banned;     /* [host]                                                        */
    uint64_t seed;             /* Philox key; stream b, step i are the counter                  */
} awd_sampling;

/* Replaces get_language_model(...) at main.py:96: allocates device state for one model. */
int awd_create(const awd_config* cfg, awd_engine** out);
int awd_destroy(awd_engine* e);

/* Replaces load_state_dict(torch.load(..)) at main.py:99.  `key` is a state_dict key of the
 * reference module, data is fp32 row-major [any].  Load is strict: commit fails
 * with AWD_EMISSING unless all keys were given.  Aliased keys resolve as the reference does:
 * module.weight_hh_l0 over weight_hh_l0_raw, 1.decoder.weight over the encoder keys when tied. */
int awd_load_param(awd_engine* e, const char* key, const float* data, const int64_t* shape, int32_t ndim);
int awd_commit_weights(awd_engine* e);

/* Replaces RNNCore.reset (awd_lstm.py:98-103): zero (h,c) of every layer for `bs` streams. */
int awd_reset(awd_engine* e, int32_t bs, void* stream);
/* Read / write RNNCore.hidden[layer] = (h, c), each fp32 (1,bs,H_layer) contiguous [any]. */
int awd_get_state(awd_engine* e, int32_t layer, float* h, float* c, void* stream);
int awd_set_state(awd_engine* e, int32_t layer, const float* h, const float* c, void* stream);
int32_t awd_current_bs(const awd_engine* e);

/* Replaces EmbeddingDropout.forward in eval mode (awd_lstm.py:114-122): out[n][:] = enc[tok[n]][:].
 * tokens int64 [any] (n), out fp32 [dev] (n, emb_sz). */
int awd_gather(awd_engine* e, const int64_t* tokens, int64_t n, float* out, void* stream);

/* Replaces SequentialRNN.__call__ -> RNNCore.forward + LinearDecoder.forward
 * (awd_lstm.py:77-91, 42-46).  tokens int64 [any] (sl,bs) seq-first.  The carried state is
 * used and updated; a `bs` different from the current one resets first (awd_lstm.py:79-81).
 *   logits      fp32 [dev] (rows, vocab_sz), or NULL to skip the decoder; rows = sl*bs, row = t*bs+b,
 *               or bs when last_step_only != 0 (the rows main.py:56,71 actually reads: res[-1]);
 *   raw_outputs array of n_layers fp32 [dev] pointers (sl,bs,H_layer), or NULL; entries may be NULL. */
int awd_forward(awd_engine* e, const int64_t* tokens, int32_t sl, int32_t bs,
                float* logits, float* const* raw_outputs, int32_t last_step_only, void* stream);

/* Replaces LinearDecoder.forward alone (awd_lstm.py:42-46) for callers that run the two halves
 * of the SequentialRNN separately: x fp32 [dev] (rows, emb_sz) -> logits fp32 [dev] (rows, vocab_sz). */
int awd_decode(awd_engine* e, const float* x, int64_t rows, float* logits, void* stream);

/* The dense contraction underneath nn.LSTM's input half and nn.Linear (awd_lstm.py:162, :45) as a
 * stand-alone operator on the tcgen05 tensor cores: C[M][N] = A[M][K] * W[N][K]^T + bias[N].
 * A, W bf16 [dev], K-contiguous, 16-byte aligned, lda/ldw/K multiples of 64 elements; C fp32 [dev]. */
int awd_tc_gemm_bf16(const void* A, int64_t lda, const void* W, int64_t ldw, int32_t M, int32_t N, int32_t K,
                     const float* bias, float* C, int64_t ldc, void* stream);

/* Replaces the id-level loop of sample_model (main.py:50-71) for bs independent streams:
 * reset; feed the prompt (P,bs); then n_new times {candidates, first allowed id, feed back}.
 * prompt int64 [any]; out_tokens int64 [any] (n_new,bs).  The whole loop runs on the device;
 * the call returns after enqueueing (copies to pageable host memory synchronise). */
int awd_generate(awd_engine* e, const int64_t* prompt, int32_t P, int32_t bs, int32_t n_new,
                 const awd_sampling* s, int64_t* out_tokens, void* stream);

/* The same loop for requests whose prompts differ in length (each reference request feeds only its own
 * words, main.py:54-57): prompt int64 [any] (P,bs) holds the prompts RIGHT-aligned (rows above a shorter
 * prompt are ignored), prompt_len int32 [host] (bs) with 1 <= prompt_len[b] <= P.  Stream b joins at
 * timestep P - prompt_len[b] from the reset state, so its result equals the one it would get alone. */
int awd_generate_ragged(awd_engine* e, const int64_t* prompt, const int32_t* prompt_len, int32_t P, int32_t bs,
                        int32_t n_new, const awd_sampling* s, int64_t* out_tokens, void* stream);

/* Continuous batching behind GET /inference (main.py:105-112): the loop of awd_generate -- n_new times
 * {decode the current state, draw an id by the sampling rule, feed it} -- for bs stream slots whose state the engine
 * keeps between calls, so a serving loop can retire finished requests and admit new ones every n_new tokens.
 * reset_mask uint8 [host] (bs) or NULL: slots whose state is zeroed first (newly admitted requests; main.py:51).
 * forced int64 [any] (F,bs) + forced_len int32 [host] (bs), or NULL/0: a slot still inside its prompt is fed
 * forced[i][b] instead of the id drawn at step i < forced_len[b] (a new request's prompt tokens, main.py:54-57; a
 * prompt longer than n_new continues in the next call); out_tokens int64 [any] (n_new,bs) shows the forced id there.
 * step_base: any value that differs between calls (selects the Philox stream, as the step index does inside a call).
 * The first call, or a call with a different bs, starts every slot from reset; awd_forward / awd_generate / awd_reset
 * in between overwrite the slots' state (reset the slots you care about afterwards).  Slots do not interact: in greedy mode a request gets exactly the ids awd_generate gives it alone. */
int awd_generate_resume(awd_engine* e, const uint8_t* reset_mask, const int64_t* forced, const int32_t* forced_len,
                        int32_t F, int32_t bs, int32_t n_new, const awd_sampling* s, int64_t step_base,
                        int64_t* out_tokens, void* stream);

/* Vocabulary-sharded decoder (no counterpart call site: the reference decodes on one CPU; this is the
 * optional multi-GPU form of awd_lstm.py:45 + main.py:61-68 for very large vocabularies).  Every rank keeps
 * the recurrence replicated and decodes only rows [v0, v1) (v0 a multiple of 256: whole decoder tiles) of the current last-layer
 * state: out_key fp32 [dev] (bs) / out_id int64 [dev] (bs) = its best allowed race key and the id it
 * resolves to (-1: none), out_banned fp32 [dev] (bs,64) = keys of the disallowed ids it owns (-inf
 * otherwise).  Race keys of different shards compare directly, so after ONE all-gather of these three small
 * arrays awd_merge_race gives every rank the token the unsharded engine would have drawn with the same
 * seed.  keys/ids/banned for awd_merge_race are [n_shards][bs], [n_shards][bs], [n_shards][bs][64]. */
int awd_decode_race_shard(awd_engine* e, int32_t v0, int32_t v1, const awd_sampling* s, int64_t step, float* out_key,
                          int64_t* out_id, float* out_banned, void* stream);
int awd_merge_race(const float* keys, const int64_t* ids, const float* banned, int32_t n_shards, int32_t bs, int32_t vocab_sz,
                   const awd_sampling* s, int64_t* out, void* stream);

/* The same exchange with ONE buffer per rank, so that one all-gather per token moves everything: a rank's block is
 * ids int64 [bs] | keys fp32 [bs] | banned fp32 [bs][64] | pad fp32 [bs], AWD_RACE_BLOCK_FLOATS_PER_STREAM * bs floats, 8-byte aligned [dev];
 * awd_merge_race_packed reads n_shards such blocks laid end to end (what all_gather_into_tensor produces). */
#define AWD_RACE_BLOCK_FLOATS_PER_STREAM 68
int awd_decode_race_shard_packed(awd_engine* e, int32_t v0, int32_t v1, const awd_sampling* s, int64_t step, void* block, void* stream);
int awd_merge_race_packed(const void* blocks, int32_t n_shards, int32_t bs, int32_t vocab_sz, const awd_sampling* s, int64_t* out,
                          void* stream);

/* Sampling rule alone (main.py:61-68) on caller logits fp32 [dev] (bs,vocab_sz): writes
 * out int64 [dev] (bs).  `step` selects the Philox counter, as step i of awd_generate does. */
int awd_sample(awd_engine* e, const float* logits, int32_t bs, const awd_sampling* s,
               int64_t step, int64_t* out, void* stream);

/* No counterpart in the reference (it has no allocator to avoid): decoder + sampling rule in one
 * call that allocates nothing, for hosts that cannot call cudaMalloc where they decode -- a
 * server that gave the whole GPU to its own arena, or a latency-critical thread that cannot
 * take the device-wide stall an allocation implies.  x fp32 [dev] (bs, emb_sz), 16-byte aligned,
 * emb_sz a multiple of 4; out int64 [dev] (bs).  The (bs, vocab_sz) logits in between live in a
 * fixed AWD_STATIC_SCRATCH_BYTES array inside the library image (resident on every GPU that
 * loads the library until exit); bs*vocab_sz*4 beyond that returns AWD_ELIMIT.  All calls share
 * the array and it keeps the previous call's contents: calls that overlap on the GPU race, so
 * serialise them.  The first call on a GPU resolves the array's address; issue it before any
 * stream capture.  Results equal awd_decode followed by awd_sample.  fp32 engines only (AWD_PREC_FP32):
 * a bf16 engine returns AWD_ESTATE -- its decoder never materialises logits (awd_generate fuses the rule). */
#define AWD_STATIC_SCRATCH_BYTES (16u << 20)
int awd_decode_sample_static(awd_engine* e, const float* x, int32_t bs, const awd_sampling* s,
                             int64_t step, int64_t* out, void* stream);

/* The noise of the sampling race, exposed for tests (no counterpart in the reference, whose noise is
 * torch.multinomial's): awd_uniform_from_bits maps a random word to u in the OPEN interval (0,1) on the host;
 * awd_gumbel_from_bits evaluates -log(-log(u)) on the device exactly as the bf16 samplers do, for n words
 * bits uint32 [dev] -> out fp32 [dev].  Both must be finite for every word, 0xFFFFFFFF included. */
float awd_uniform_from_bits(uint32_t bits);
int awd_gumbel_from_bits(const uint32_t* bits, int32_t n, float* out, void* stream);

/* Device error flag set by kernels (e.g. token id out of range); reading it synchronises the
 * stream.  Returns AWD_OK or AWD_ETOKEN and clears the flag. */
int awd_check_device_errors(awd_engine* e, void* stream);

/* Optional per-category kernel timing with CUDA events recorded on the launching stream
 * (bench.py's roofline figures).  awd_profile(e,1) starts a fresh recording, awd_profile(e,0)
 * stops it; awd_profile_read synchronises and sums one category:
 * 0 = input projection GEMM, 1 = recurrent step + cell, 2 = decoder, 3 = sampling;
 * 4, 5, 6 = layer 0, layer 1, layers >= 2 of category 1 (persistent generation kernel only). */
int awd_profile(awd_engine* e, int32_t enable);
int awd_profile_read(awd_engine* e, int32_t category, double* total_ms, int64_t* launches);

/* Number of kernels launched by this library in this process (bench.py's gpu_launches). */
int64_t awd_launch_count(void);

const char* awd_last_error(void);
const char* awd_version(void);

#ifdef __cplusplus
}
#endif
#endif /* AWD_B200_H */
