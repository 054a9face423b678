/* qtts_b200.h — C ABI of the B200-native Qwen3-TTS codec-frame engine (libqtts_b200.so).
 *
 * This is the drop-in boundary for ONE hot path of syb0rg/qwen3-tts-jetson: the per-frame
 * codec-token generation loop (Talker decode step -> Code Predictor 5 layers x 16 positions ->
 * sampling).  Every entry point below names the reference interface it replaces.  Signatures
 * carry only plain pointers and sizes (host memory unless stated); no torch / ggml types.
 *
 * Conventions (mirroring the reference, src/tts_transformer.h:164-245):
 *   - return value 1 = success, 0 = failure; qtts_last_error() gives the message
 *     (reference: bool + get_error()).
 *   - caller owns every buffer passed in or out; they are only used during the call.
 *   - one caller thread per engine (reference classes are not thread safe either).
 *   - there is NO CPU fallback: every call fails loudly if CUDA is unavailable.
 */
#ifndef QTTS_B200_H
#define QTTS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct qtts_engine qtts_engine;

/* Hyper-parameters and special ids.  Reference: tts_transformer_config,
 * src/tts_transformer.h:21-69, filled by parse_config src/tts_transformer.cpp:288-442. */
typedef struct qtts_config {
    int32_t hidden_size, n_layers, n_heads, n_kv_heads, intermediate_size, head_dim;
    int32_t codec_vocab_size, n_codebooks;
    int32_t cp_layers, cp_vocab_size, cp_hidden_size, cp_intermediate_size;
    int32_t cp_n_heads, cp_n_kv_heads, cp_head_dim;
    int32_t text_vocab_size, text_embd_dim;
    float   rms_norm_eps, rope_theta;
    int32_t codec_pad_id, codec_bos_id, codec_eos_id;
    int32_t codec_think_id, codec_nothink_id, codec_think_bos_id, codec_think_eos_id;
    int32_t tts_bos_id, tts_eos_id, tts_pad_id;
} qtts_config;

/* tensor element types accepted by qtts_upload_tensor (GGUF stores 2-D as F16, 1-D as F32:
 * scripts/convert_tts_to_gguf.py:289-297) */
enum { QTTS_F32 = 0, QTTS_F16 = 1, QTTS_BF16 = 2 };

/* numerics modes (see DESIGN.md "Numerics"):
 *   REFERENCE: reproduces the rounding points of the reference's ggml-cpu execution, including
 *              the sequential F16 V-accumulator of its flash attention (ops.cpp:8108-8344), and
 *              the reference's codebook-0 top-k threshold = scored_idx[top_k - 1] after
 *              std::nth_element (src/tts_transformer.cpp:2464-2466): whichever member of the top-k
 *              set libstdc++'s introselect leaves there, so fewer than top_k candidates may survive;
 *   FAST:      same rounding points for every projection, KV cache and Q, but the Talker's
 *              decode attention accumulates P.V in FP32 with a parallel split-KV reduction, and the
 *              codebook-0 threshold is the exact top_k-th largest logit (what the reference's
 *              code-predictor sampler does with std::partial_sort, :2211-2218). */
enum { QTTS_NUMERICS_REFERENCE = 0, QTTS_NUMERICS_FAST = 1 };

/* ---- life cycle ------------------------------------------------------------------------- */

/* Create an engine on CUDA device `device` for up to `max_streams` concurrent utterances
 * with `max_ctx` KV slots each.  Replaces TTSTransformer() + backend selection
 * (src/tts_transformer.cpp:51-153, src/gguf_loader.cpp:26-76). */
int qtts_create(const qtts_config * cfg, int device, int max_streams, int max_ctx,
                qtts_engine ** out);
void qtts_destroy(qtts_engine * e);                       /* ~TTSTransformer / unload_model */
const char * qtts_last_error(const qtts_engine * e);      /* get_error() */
/* message of the last failed qtts_create (no engine to ask) */
const char * qtts_create_error(void);

/* Upload one GGUF tensor by its GGUF name (names: scripts/convert_tts_to_gguf.py:42-109,
 * consumed by create_tensors src/tts_transformer.cpp:444-731).  dims are in GGUF/ggml order
 * (dims[0] = innermost = input features).  Replaces load_tensor_data
 * (src/tts_transformer.cpp:733-792).  Projection matrices stay F16, the file's own values (2-D F32 tensors are
 * rounded to F16, as ggml's F16 mat-mul path would); the count of values a BF16 copy could not represent is
 * accumulated for information (qtts_weight_stats; the synthetic fixtures are BF16-exact). */
int qtts_upload_tensor(qtts_engine * e, const char * name, int dtype,
                       const int64_t * dims, int n_dims, const void * host_data);
/* The whole tensor set straight from a GGUF file (replaces load_tensor_data, src/tts_transformer.cpp:733-792): the
 * file is mapped and its F16 / F32 tensors go page cache -> pinned staging ring -> HBM without a per-tensor host
 * copy; quantised tensor types (Q8_0, Q4_0, Q4_K ... which scripts/convert_tts_to_gguf.py can emit) are rejected with
 * the tensor's name and type.  Call qtts_finalize afterwards.  qtts_load_stats: wall ms and bytes of that stage. */
int qtts_load_gguf(qtts_engine * e, const char * gguf_path);
int qtts_load_stats(const qtts_engine * e, double * load_ms, int64_t * bytes);
/* Build the fused device layouts (QKV concatenation, gate/up row interleave), check that every
 * tensor of the path is present.  Must be called once after the uploads. */
int qtts_finalize(qtts_engine * e);
/* out[0] = #projection weights uploaded, out[1] = #of those not BF16-exact,
 * out[2] = device bytes held by weights, out[3] = device bytes held by KV caches */
int qtts_weight_stats(const qtts_engine * e, int64_t out[4]);

int qtts_set_numerics(qtts_engine * e, int mode, int reference_threads);
/* persistent = 1: one cooperative launch per frame group (default); 0: one launch per stage
 * (debug path, same device code). */
int qtts_set_persistent(qtts_engine * e, int persistent);
/* BENCHMARK switch, not reference behaviour: on = 1 makes codebook-0 EOS unselectable for
 * streams begun afterwards, so every utterance runs exactly max_len frames (BASELINE configs
 * quote fixed frame counts; a random-init model otherwise samples EOS at a random frame). */
int qtts_set_ignore_eos(qtts_engine * e, int on);

/* ---- reference-level entry points (TTSTransformer methods, one stream = `slot`) ----------- */

/* init_kv_cache / clear_kv_cache (src/tts_transformer.cpp:794-900).  The cache is F16,
 * token-major [n_ctx][n_kv_heads*head_dim] per layer, like the reference's (:823-833). */
int qtts_clear_kv_cache(qtts_engine * e, int slot);

/* build_prefill_graph + project_text_tokens (src/tts_transformer.cpp:1026-1231): assemble the
 * prefill embeddings, the trailing text rows and the pad row for `slot` on the device.
 * speaker may be NULL (no speaker row).  Optional outputs (may be NULL):
 *   prefill_out [prefill_len*H], trailing_out [trailing_len*H], pad_out [H]. */
int qtts_build_prefill(qtts_engine * e, int slot, const int32_t * text_tokens, int32_t n_tokens,
                       const float * speaker_embd, int32_t language_id,
                       int32_t * prefill_len, int32_t * trailing_len,
                       float * prefill_out, float * trailing_out, float * pad_out);

/* forward_prefill (src/tts_transformer.cpp:1829-1920): n_tokens rows of embeddings at
 * positions n_past.. ; hidden_out [n_tokens*H] = normalised hidden of EVERY token like the
 * reference's `output` (:1893; the last row is what feeds the code predictor), logits_out [V] =
 * logits of the last token.  Either may be NULL. */
int qtts_forward_prefill(qtts_engine * e, int slot, const float * embd, int32_t n_tokens,
                         int32_t n_past, float * hidden_out, float * logits_out);

/* forward_text (src/tts_transformer.cpp:1922-1950): project_text_tokens (+ speaker_embd on every row, may
 * be NULL) then forward_prefill; hidden_out [n_tokens*H]. */
int qtts_forward_text(qtts_engine * e, int slot, const int32_t * text_tokens, int32_t n_tokens,
                      const float * speaker_embd, int32_t n_past, float * hidden_out);
/* forward_codec (src/tts_transformer.cpp:2030-2040): codec_embd row of one token -> forward_step. */
int qtts_forward_codec(qtts_engine * e, int slot, int32_t codec_token, int32_t n_past, float * logits_out);

/* get_hidden_states (src/tts_transformer.cpp:2042-2048): normalised hidden [H] of the last forward call of `slot`
 * (last_hidden_, what predict_codes_autoregressive is fed). */
int qtts_get_hidden_states(qtts_engine * e, int slot, float * hidden_out);

/* forward_step (src/tts_transformer.cpp:1952-2028). */
int qtts_forward_step(qtts_engine * e, int slot, const float * step_embd, int32_t n_past,
                      float * logits_out, float * hidden_out);

/* predict_codes_autoregressive (src/tts_transformer.cpp:2153-2340) and
 * TRTCodePredictor::run_greedy_loop / run_sampling_loop (src/trt_code_predictor.cpp:484-726).
 * temperature <= 0: greedy.  uniforms15: one uniform in [0,1) per code (as run_sampling_loop's
 * rand_vals); NULL draws them from the engine's counter-based generator.
 * forced15 (may be NULL): teacher forcing — the code fed back after step s is forced15[s]
 * while out15[s] is still the engine's own choice.  logits_out (may be NULL): [15*cp_vocab]. */
int qtts_predict_codes(qtts_engine * e, int slot, const float * hidden, int32_t cb0,
                       float temperature, int32_t top_k, const float * uniforms15,
                       const int32_t * forced15, int32_t * out15, float * logits_out);

/* TRTCodePredictor::step / step_with_token (src/trt_code_predictor.h:53-66, .cpp:384-482): ONE code-predictor
 * position at position_id in [0, 16) over the slot's 16-entry code-predictor cache (reset_kv_cache, :50, has nothing
 * to clear: positions above the current one are never read).  input != NULL: F32 vector of input_size floats
 * (hidden_size when needs_proj and the model has mtp_proj, else cp_hidden_size); input == NULL: row token_id of
 * embedding table embedding_idx (0 = talker codec_embd, 1..14 = code_pred.codec_embd[0..13]).
 * lm_head_idx >= 0: output = logits [cp_vocab]; < 0: output = normalised hidden [cp_hidden]. */
int qtts_cp_step(qtts_engine * e, int slot, const float * input, int32_t input_size, int32_t token_id,
                 int32_t embedding_idx, int32_t position_id, int32_t lm_head_idx, int needs_proj,
                 float * output, int32_t output_size);

/* frame callback: frame_callback_t (src/tts_transformer.h:164-170): codes for n_frames frames,
 * row-major [n_frames][n_codebooks]; return 0 to stop generation. */
typedef int (*qtts_frame_cb)(const int32_t * codes, int32_t n_frames, int32_t n_codebooks,
                             void * user);

/* generate (src/tts_transformer.cpp:2342-2574).  codes_out must hold max_len*n_codebooks
 * values; *n_frames_out receives the number of frames produced.  seed drives the sampler
 * when temperature > 0 (the reference seeds from std::random_device,
 * src/tts_transformer.h:298, so its sampled streams are not reproducible). */
int qtts_generate(qtts_engine * e, int slot, const int32_t * text_tokens, int32_t n_tokens,
                  const float * speaker_embd, int32_t max_len, int32_t language_id,
                  float repetition_penalty, float temperature, int32_t top_k, uint64_t seed,
                  int32_t * codes_out, int32_t * n_frames_out,
                  qtts_frame_cb on_frames, int32_t callback_interval, void * user);

/* ---- concurrent-utterance engine (the axis the reference lacks: SURVEY.md §8e) ----------- */

/* Start an utterance in `slot`: build_prefill + forward_prefill; the stream is then advanced
 * by qtts_step_frames together with every other active stream. */
int qtts_stream_begin(qtts_engine * e, int slot, const int32_t * text_tokens, int32_t n_tokens,
                      const float * speaker_embd, int32_t max_len, int32_t language_id,
                      float repetition_penalty, float temperature, int32_t top_k, uint64_t seed);
/* The same in two phases, so that the prefill of MANY utterances is one batched pass:
 * qtts_stream_prepare = build_prefill_graph (src/tts_transformer.cpp:1093-1231) + per-stream state, the stream is
 * "pending"; qtts_prefill_pending = forward_prefill (:1829-1920) of every pending stream, per lock-step group all
 * together on the batched kernel (one weight pass per prefill position for the group; FAST attention numerics),
 * after which they are active.  qtts_step_frames runs it implicitly for streams still pending.
 * qtts_last_loop_ms afterwards = device time of the batched prefill. */
int qtts_stream_prepare(qtts_engine * e, int slot, const int32_t * text_tokens, int32_t n_tokens,
                        const float * speaker_embd, int32_t max_len, int32_t language_id,
                        float repetition_penalty, float temperature, int32_t top_k, uint64_t seed);
int qtts_prefill_pending(qtts_engine * e);
/* Advance every active stream by up to n_frames frames.  codes_out is
 * [max_streams][n_frames][n_codebooks]; produced[slot] = frames written for that slot;
 * active[slot] = 1 while the stream has neither hit EOS nor max_len. */
int qtts_step_frames(qtts_engine * e, int32_t n_frames, int32_t * codes_out,
                     int32_t * produced, int32_t * active);

/* Parity hooks for a stream begun / prepared but not yet stepped (all arrays are host memory, copied):
 *   forced_codes [n_forced][n_codebooks]: teacher forcing - in frames [0, n_forced) the codes fed back (code-predictor
 *     inputs, next Talker input, repetition-penalty set) are these, while the codes the engine emits stay its own choice;
 *     an own EOS inside the forced range does not end the stream.
 *   uniforms [n_uniform][n_codebooks]: the uniform in [0,1) each sampled code consumes ([f][0] codebook 0, [f][c] code
 *     c) instead of the engine's counter-based generator - TRTCodePredictor::run_sampling_loop's rand_vals
 *     (src/trt_code_predictor.cpp:602-726, src/tts_transformer.cpp:2663-2665) extended to codebook 0.
 * Either pointer may be NULL.  n_* <= max_ctx. */
int qtts_stream_force(qtts_engine * e, int slot, const int32_t * forced_codes, int32_t n_forced,
                      const float * uniforms, int32_t n_uniform);
/* Codebook-0 post-processing + sampling of ONE frame on caller-supplied raw Talker logits [V]
 * (src/tts_transformer.cpp:2417-2499, the block generate() runs inline): seen_tokens = the distinct
 * codebook-0 tokens emitted so far, u = the uniform draw (ignored for temperature <= 0).
 * token_out = chosen id (codec_eos_id = stop); processed_out (may be NULL) [V] = logits after
 * suppress / repetition penalty / EOS ramp.  Uses the engine's numerics mode for the top-k threshold. */
int qtts_sample_cb0(qtts_engine * e, int slot, const float * raw_logits, int32_t frame, int32_t n_text_tokens,
                    const int32_t * seen_tokens, int32_t n_seen, float repetition_penalty, float temperature,
                    int32_t top_k, float u, int32_t * token_out, float * processed_out);
/* debug switches: bit 0 = the batch-1 frame kernel also keeps the 15 code-predictor logit vectors of the last frame
 * (qtts_stream_logits; the batched kernel always does). */
int qtts_set_debug(qtts_engine * e, int flags);
/* KV capacity per stream without touching the weights (init_kv_cache, src/tts_transformer.cpp:794-842, which the
 * reference calls from generate() whenever prefill_len + max_len + 8 outgrows the cache, :2383-2388).  max_text =
 * longest utterance in text tokens (trailing-row capacity; <= 0 keeps the current one).  Streams in flight are
 * dropped.  n_ctx is capped at 4608 KV slots per stream (the attention scratch; the reference has no cap). */
int qtts_resize_kv(qtts_engine * e, int32_t n_ctx, int32_t max_text);

/* Vocoder hand-off: frames [first_frame, first_frame + n_frames) of `slot`, read from the
 * stream's device-resident code ring, in the input layout of the reference's streamed vocoder: int64
 * [1][n_codebooks][fixed_frames], frames beyond n_frames zero (TRTVocoderDecoder::decode, src/trt_vocoder.cpp:98-170,
 * called with the 40-frame chunks of generate()'s callback, src/tts_transformer.cpp:2517-2523, 2563-2570).
 * out_is_device = 1: `out` is a DEVICE pointer (e.g. the TensorRT input binding: the chunk never visits the host);
 * 0: host memory.  n_frames <= fixed_frames; the frames must have been generated. */
int qtts_codes_chunk(qtts_engine * e, int slot, int32_t first_frame, int32_t n_frames, int32_t fixed_frames,
                     int64_t * out, int out_is_device);

/* No counterpart in the reference.  Allocation-free form of
 * qtts_codes_chunk with a HOST destination, for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it): qtts_codes_chunk stages the chunk in a device buffer it allocates on
 * first use and regrows on demand, this entry point stages it in a file-scope `static __device__` array of
 * QTTS_NOALLOC_CHUNK_WORDS int64 words (64 KB; every GPU that loads the library holds it until the process exits) and
 * calls no cudaMalloc / cudaFree / allocating library routine.  Same arguments, result and layout as qtts_codes_chunk
 * with out_is_device = 0 (int64 [1][n_codebooks][fixed_frames], frames beyond n_frames zero).
 * Limit: n_codebooks * fixed_frames <= QTTS_NOALLOC_CHUNK_WORDS (16 codebooks x 512 frames); larger requests fail
 * with a message naming the limit.  All calls on one GPU share the array (it keeps the last call's contents; every
 * word that is read back is rewritten first, so nothing needs clearing): calls that OVERLAP on a GPU - two engines or
 * two host threads - race on it; the call is synchronous, so serialise callers per GPU. */
#define QTTS_NOALLOC_CHUNK_WORDS 8192
int qtts_codes_chunk_noalloc(qtts_engine * e, int slot, int32_t first_frame, int32_t n_frames, int32_t fixed_frames,
                             int64_t * out_host);

/* 1 (default): qtts_step_frames advances all active streams in lock step with the batched tcgen05
 * kernel (groups of <= 256 slots, env QTTS_BGROUP lowers it); 0: the batch-1 kernel, one stream after another. */
int qtts_set_batched(qtts_engine * e, int on);
/* B-operand (weight) element type of the batched kernel's tensor-core GEMMs.  QTTS_F16 (default):
 * the F16 image of the engine's BF16 weights (exact: BF16-exact values inside F16's range), so both
 * kernels multiply the same real numbers; activations are F16 as in the reference's ggml_mul_mat.
 * QTTS_BF16 is rejected: tcgen05.mma kind::f16 traps on mixed F16 x BF16 operands (measured on
 * B200).  Must be called before qtts_finalize. */
int qtts_set_batched_weight_format(qtts_engine * e, int fmt);
/* parity hook: processed codebook-0 logits [V] (after suppress / penalty / EOS ramp) and, batched
 * mode only, the 15 code-predictor logit vectors [15*cp_vocab] of the LAST frame of `slot`. */
int qtts_stream_logits(qtts_engine * e, int slot, float * cb0_out, float * cp_out);

/* ---- instrumentation -------------------------------------------------------------------- */
/* stage profiler: CTA 0 of the frame kernel accumulates SM-clock cycles per op type
 * (0 GEMV, 1 ATTN, 2 SAMPLE_CB0, 3 SAMPLE_CP, 4 FRAME_END, 5 COPY): out64[type*8 + {0 count,
 * 1 stage, 2 barrier wait, 3 GEMV input load, 4 GEMV norm, 5 GEMV rows}].  out64 (may be NULL)
 * receives the 64 + 4*256 counters accumulated so far (after the first 64: per-CTA [cta][gemv stage, gemv barrier, attn stage, attn barrier]); they are then cleared.  on = 0: off; on = 1 + c:
 * CTA c records.  [6] = cycles thread 0 waited for ring chunks. */
int qtts_profile(qtts_engine * e, int on, uint64_t * out64);
/* stage timeline of the batch-1 frame kernel: thread 0 of CTA `cta` stores SM-clock stamps (low 32 bits) for the
 * first 8192 stages of every launch, 8 words per stage: [type | op index << 8, start, own input poll done, input
 * staged, rows done, stage end (after the grid barrier if any), frame, 0].  Store-only, so the timeline is not
 * distorted by read-modify-write round trips.  out (may be NULL) receives the previous recording, which is cleared. */
int qtts_trace(qtts_engine * e, int on, int cta, uint32_t * out, int32_t max_stages);
/* instrumentation: microseconds per grid-wide barrier of `variant` (see csrc/qtts_microbench.cu) */
double qtts_debug_barrier_us(int device, int variant, int iters);
/* batched kernel stage timers (CTA 0, SM clock cycles): out64[(2*type + 8*f16_input)*4 + {0 count,
 * 1 stage cycles, 2 barrier cycles}], type 0 GEMM, 1 code-predictor attention, 2 Talker attention,
 * 3 sampling; cleared after the read. */
int qtts_batched_profile(qtts_engine * e, int on, uint64_t * out64);
/* unit test of ONE projection stage of the batched kernel on caller-supplied operands: W [rows][K],
 * X [n_tok*n_streams][K] (F32 on the host; W rounded to BF16, X to F16 when in_f16).  norm_w, res,
 * bias, ssq_out may be NULL.  epi: 0 store, 1 +res, 2 SwiGLU (out [rows_act][rows/2]), 3 +bias. */
int qtts_debug_bgemm(int device, int weight_fmt, int n_streams, int n_tok, int rows, int K, int epi, int in_f16,
                     const float * W, const float * X, const float * norm_w, const float * res, const float * bias,
                     float eps, float * out, float * ssq_out);
/* number of kernel launches issued by this engine so far (bench.py's gpu_launches) */
int64_t qtts_launch_count(const qtts_engine * e);
/* device time (ms, CUDA events on the engine's stream) of the last qtts_generate /
 * qtts_step_frames frame loop, prefill excluded */
double qtts_last_loop_ms(const qtts_engine * e);
/* the engine's CUDA stream as a void* (cudaStream_t), for callers that time with events */
void * qtts_stream_handle(const qtts_engine * e);

#ifdef __cplusplus
}
#endif
#endif /* QTTS_B200_H */
