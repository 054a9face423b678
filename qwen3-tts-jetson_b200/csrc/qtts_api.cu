// qtts_api.cu — the extern "C" surface declared in include/qtts_b200.h.
#include "qtts_engine.h"
#include <cstring>
#include <cstdio>
#include <algorithm>
#include <mutex>

using namespace qtts;

namespace qtts {
extern "C" __global__ void qtts_run_program(const Op * ops, int n_ops, StreamState * st, GridBarrier * bar,
                                            int n_frames, RunCfg cfg);
extern "C" __global__ void qtts_run_batched(const __grid_constant__ BRun run_param);
extern "C" __global__ void qtts_run_batched_ext(const __grid_constant__ BRun run_param);
}

struct qtts_engine { Engine e; };

static std::string g_create_err;

static Slot * get_slot(Engine * e, int slot) {
    if (!e->finalized) { e->err = "engine not finalized (call qtts_finalize after the uploads)"; return nullptr; }
    if (slot < 0 || slot >= (int) e->slots.size()) { e->err = "slot out of range"; return nullptr; }
    return &e->slots[slot];
}

extern "C" {

const char * qtts_create_error(void) { return g_create_err.c_str(); }

int qtts_create(const qtts_config * cfg, int device, int max_streams, int max_ctx, qtts_engine ** out) {
    *out = nullptr;
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0) {
        cudaGetLastError();
        g_create_err = "no CUDA device: the B200 engine has no CPU fallback";
        return 0;
    }
    if (device < 0 || device >= n_dev) { g_create_err = "device index out of range"; return 0; }
    if (cudaSetDevice(device) != cudaSuccess) { g_create_err = "cudaSetDevice failed"; return 0; }
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, device);
    if (prop.major < 10) { g_create_err = "device is not sm_100-class (this library only carries sm_100a code)"; return 0; }
    if (max_streams < 1 || max_ctx < 32) { g_create_err = "max_streams must be >= 1 and max_ctx >= 32"; return 0; }
    qtts_engine * h = new qtts_engine();
    Engine * e = &h->e;
    e->cfg = *cfg;
    e->device = device;
    e->max_streams = max_streams;
    e->max_ctx = max_ctx;
    e->max_text = std::max(max_ctx, 64);
    e->n_sms = prop.multiProcessorCount;
    // flag-carrying activations need every CTA to own rows of every stage of a chain (a CTA that
    // only reads could be overtaken by the next writer of the buffer)
    if (const char * v = getenv("QTTS_LL")) e->ll = atoi(v) ? 1 : 0;
    if (std::min(cfg->hidden_size, cfg->cp_hidden_size) / 2 < e->n_sms) e->ll = 0;
    int k_max = std::max({cfg->hidden_size, cfg->intermediate_size, cfg->cp_hidden_size, cfg->cp_intermediate_size,
                          cfg->text_embd_dim, cfg->n_heads * cfg->head_dim, cfg->cp_n_heads * cfg->cp_head_dim});
    // widest input of a two-token (merged code-predictor) stage
    const int k2_max = std::max({cfg->cp_hidden_size, cfg->cp_intermediate_size, cfg->cp_n_heads * cfg->cp_head_dim,
                                 cfg->cp_hidden_size != cfg->hidden_size ? cfg->hidden_size : 0});
    if (max_ctx > kMaxScores) { g_create_err = "max_ctx exceeds the engine limit (kMaxScores)"; delete h; return 0; }
    e->n_slots = n_slots_for(k_max, k2_max, (size_t) prop.sharedMemPerBlockOptin);
    if (e->n_slots < 2) { g_create_err = "model too wide for the shared-memory weight ring"; delete h; return 0; }
    e->slot_bytes = (int) slot_bytes_for(k_max);
    e->xs_elems = (int) xs_elems_for(k_max, k2_max);
    e->smem_bytes = smem_required(k_max, k2_max, e->n_slots);
    if (cudaFuncSetAttribute(qtts_run_program, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) e->smem_bytes) != cudaSuccess) {
        g_create_err = std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(cudaGetLastError()); delete h; return 0;
    }
    int occ = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, qtts_run_program, kThreads, e->smem_bytes);
    if (occ < 1 || !prop.cooperativeLaunch) { g_create_err = "cooperative launch of the frame kernel is not possible on this device"; delete h; return 0; }
    cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking);
    cudaEventCreate(&e->ev0);
    cudaEventCreate(&e->ev1);
    *out = h;
    return 1;
}

void qtts_destroy(qtts_engine * h) {
    if (!h) return;
    Engine * e = &h->e;
    cudaSetDevice(e->device);
    cudaStreamSynchronize(e->stream);
    for (void * p : e->allocs) cudaFree(p);
    for (void * p : e->stream_allocs) cudaFree(p);
    if (e->step_host) cudaFreeHost(e->step_host);
    for (auto & kv : e->tensors) if (kv.second.raw) cudaFree(kv.second.raw);
    if (e->ev0) cudaEventDestroy(e->ev0);
    if (e->ev1) cudaEventDestroy(e->ev1);
    if (e->stream) cudaStreamDestroy(e->stream);
    delete h;
}

const char * qtts_last_error(const qtts_engine * h) { return h ? h->e.err.c_str() : "null engine"; }

int qtts_upload_tensor(qtts_engine * h, const char * name, int dtype, const int64_t * dims, int n_dims, const void * host) {
    cudaSetDevice(h->e.device);
    return engine_upload(&h->e, name, dtype, dims, n_dims, host) ? 1 : 0;
}

int qtts_load_gguf(qtts_engine * h, const char * path) {
    cudaSetDevice(h->e.device);
    return engine_load_gguf(&h->e, path) ? 1 : 0;
}

int qtts_load_stats(const qtts_engine * h, double * load_ms, int64_t * bytes) {
    if (load_ms) *load_ms = h->e.load_ms;
    if (bytes) *bytes = h->e.load_bytes;
    return 1;
}

int qtts_finalize(qtts_engine * h) {
    cudaSetDevice(h->e.device);
    return engine_finalize(&h->e) ? 1 : 0;
}

int qtts_weight_stats(const qtts_engine * h, int64_t out[4]) {
    out[0] = h->e.n_weights; out[1] = h->e.n_inexact; out[2] = h->e.weight_bytes; out[3] = h->e.kv_bytes;
    return 1;
}

static bool rebuild_frame_programs(Engine * e) {
    for (auto & s : e->slots) {
        s.frame_host.clear();
        emit_frame(e, s, s.frame_host);
        if (!e->h2d(s.frame_prog, s.frame_host.data(), s.frame_host.size() * sizeof(Op))) return false;
        const std::vector<WSched> S = build_sched(e, s.frame_host);
        if (!e->h2d(s.frame_sched, S.data(), S.size() * sizeof(WSched))) return false;
    }
    return true;
}

int qtts_set_numerics(qtts_engine * h, int mode, int reference_threads) {
    Engine * e = &h->e;
    if (mode != QTTS_NUMERICS_REFERENCE && mode != QTTS_NUMERICS_FAST) { e->err = "unknown numerics mode"; return 0; }
    e->numerics = mode;
    if (reference_threads > 0) e->ref_threads = std::min(reference_threads, 64);
    if (e->finalized) { cudaSetDevice(e->device); cudaStreamSynchronize(e->stream); if (!rebuild_frame_programs(e)) return 0; }
    return 1;
}

int qtts_set_persistent(qtts_engine * h, int persistent) { h->e.persistent = persistent ? 1 : 0; return 1; }

int qtts_set_ignore_eos(qtts_engine * h, int on) { h->e.ignore_eos = on ? 1 : 0; return 1; }

int qtts_clear_kv_cache(qtts_engine * h, int slot) {
    Engine * e = &h->e;
    Slot * s = get_slot(e, slot);
    if (!s) return 0;
    cudaSetDevice(e->device);
    StreamState z{};
    if (cudaMemcpyAsync(s->st, &z, sizeof(z), cudaMemcpyHostToDevice, e->stream) != cudaSuccess) { e->err = "state reset failed"; return 0; }
    cudaStreamSynchronize(e->stream);
    return 1;
}

static bool do_build_prefill(Engine * e, Slot & s, const int32_t * tokens, int32_t n, const float * speaker, int32_t language_id) {
    if (!tokens) return e->fail("text_tokens is null");
    if (speaker) QTTS_CUDA(cudaMemcpyAsync(s.speaker, speaker, e->cfg.hidden_size * sizeof(float), cudaMemcpyHostToDevice, e->stream));
    std::vector<Op> P;
    if (!emit_build_prefill(e, s, P, tokens, n, speaker != nullptr, language_id)) return false;
    return run_program(e, s, P, 1, false);
}

int qtts_build_prefill(qtts_engine * h, int slot, const int32_t * tokens, int32_t n, const float * speaker,
                       int32_t language_id, int32_t * prefill_len, int32_t * trailing_len,
                       float * prefill_out, float * trailing_out, float * pad_out) {
    Engine * e = &h->e;
    Slot * s = get_slot(e, slot);
    if (!s) return 0;
    cudaSetDevice(e->device);
    if (!do_build_prefill(e, *s, tokens, n, speaker, language_id)) return 0;
    const size_t H = e->cfg.hidden_size;
    cudaError_t r = cudaStreamSynchronize(e->stream);
    if (r != cudaSuccess) { e->err = std::string("build_prefill: ") + cudaGetErrorString(r); return 0; }
    if (prefill_len) *prefill_len = s->prefill_len;
    if (trailing_len) *trailing_len = s->trailing_len;
    if (prefill_out) cudaMemcpy(prefill_out, s->prefill, s->prefill_len * H * sizeof(float), cudaMemcpyDeviceToHost);
    if (trailing_out) cudaMemcpy(trailing_out, s->trailing, s->trailing_len * H * sizeof(float), cudaMemcpyDeviceToHost);
    if (pad_out) cudaMemcpy(pad_out, s->pad, H * sizeof(float), cudaMemcpyDeviceToHost);
    return 1;
}

// prefill over rows already resident in s.prefill (token by token; identical arithmetic to the
// reference's causal prefill graph because every K/V it attends to is the F16 cache copy)
static bool do_forward_prefill(Engine * e, Slot & s, int n_tokens, int n_past, bool all_hidden = false) {
    if (n_tokens <= 0) return e->fail("n_tokens must be > 0");                       // src/tts_transformer.cpp:1840-1843
    if (n_tokens > 64) return e->fail("forward_prefill: at most 64 tokens per call");
    if (n_past < 0) return e->fail("forward_prefill: n_past must be >= 0");
    if (n_past + n_tokens > e->max_ctx) return e->fail("Context length exceeded");   // :1852-1855
    std::vector<Op> P;
    const size_t H = e->cfg.hidden_size;
    if (all_hidden && !s.prefill_hidden) {
        s.prefill_hidden = e->dalloc<float>((size_t) 64 * H);
        if (!s.prefill_hidden) return e->fail("out of device memory for the prefill hidden rows");
    }
    for (int t = 0; t < n_tokens; ++t) {
        // the reference returns the normalised hidden of EVERY row (:1893): with all_hidden the head stage (whose
        // prologue is the output norm) runs for every token and leaves its row in prefill_hidden
        const bool head = all_hidden || t == n_tokens - 1;
        emit_talker_step(e, s, P, IN_VEC, s.prefill + t * H, false, n_past + t, ATTN_SOFTMAX_P16, head, false);
        if (all_hidden && t < n_tokens - 1) P.back().hid_out = s.prefill_hidden + t * H;
    }
    if (!run_program(e, s, P, 1, false)) return false;
    const int np = n_past + n_tokens;
    QTTS_CUDA(cudaMemcpyAsync(&s.st->n_past, &np, sizeof(int), cudaMemcpyHostToDevice, e->stream));
    QTTS_CUDA(cudaStreamSynchronize(e->stream));
    return true;
}

int qtts_forward_prefill(qtts_engine * h, int slot, const float * embd, int32_t n_tokens, int32_t n_past,
                         float * hidden_out, float * logits_out) {
    Engine * e = &h->e;
    Slot * s = get_slot(e, slot);
    if (!s) return 0;
    cudaSetDevice(e->device);
    if (!embd) { e->err = "prefill_embd is null"; return 0; }
    if (n_tokens <= 0) { e->err = "n_tokens must be > 0"; return 0; }
    if (n_tokens > 64) { e->err = "forward_prefill: at most 64 tokens per call"; return 0; }
    const size_t H = e->cfg.hidden_size;
    if (cudaMemcpyAsync(s->prefill, embd, n_tokens * H * sizeof(float), cudaMemcpyHostToDevice, e->stream) != cudaSuccess) { e->err = "H2D failed"; return 0; }
    if (!do_forward_prefill(e, *s, n_tokens, n_past, hidden_out != nullptr && n_tokens > 1)) return 0;
    if (hidden_out) {
        if (n_tokens > 1) cudaMemcpy(hidden_out, s->prefill_hidden, (size_t) (n_tokens - 1) * H * sizeof(float), cudaMemcpyDeviceToHost);
        cudaMemcpy(hidden_out + (size_t) (n_tokens - 1) * H, s->hidden, H * sizeof(float), cudaMemcpyDeviceToHost);
    }
    if (logits_out) cudaMemcpy(logits_out, s->logits, e->cfg.codec_vocab_size * sizeof(float), cudaMemcpyDeviceToHost);
    cudaError_t r = cudaGetLastError();
    if (r != cudaSuccess) { e->err = std::string("forward_prefill: ") + cudaGetErrorString(r); return 0; }
    return 1;
}

int qtts_forward_step(qtts_engine * h, int slot, const float * step_embd, int32_t n_past, float * logits_out, float * hidden_out) {
    Engine * e = &h->e;
    Slot * s = get_slot(e, slot);
    if (!s) return 0;
    cudaSetDevice(e->device);
    if (!step_embd) { e->err = "step_embd is null"; return 0; }
    if (n_past < 0) { e->err = "forward_step: n_past must be >= 0"; return 0; }
    if (n_past + 1 > e->max_ctx) { e->err = "Context length exceeded"; return 0; }       // src/tts_transformer.cpp:1972-1975
    const size_t H = e->cfg.hidden_size;
    if (cudaMemcpyAsync(s->in_embd, step_embd, H * sizeof(float), cudaMemcpyHostToDevice, e->stream) != cudaSuccess) { e->err = "H2D failed"; return 0; }
    std::vector<Op> P;
    const int mode = (e->numerics == QTTS_NUMERICS_REFERENCE) ? ATTN_FLASH_F16SEQ : ATTN_FAST;
    emit_talker_step(e, *s, P, IN_VEC, s->in_embd, false, n_past, mode, true, false);
    if (!run_program(e, *s, P, 1, false)) return 0;
    const int np = n_past + 1;
    cudaMemcpyAsync(&s->st->n_past, &np, sizeof(int), cudaMemcpyHostToDevice, e->stream);
    cudaError_t r = cudaStreamSynchronize(e->stream);
    if (r != cudaSuccess) { e->err = std::string("forward_step: ") + cudaGetErrorString(r); return 0; }
    if (logits_out) cudaMemcpy(logits_out, s->logits, e->cfg.codec_vocab_size * sizeof(float), cudaMemcpyDeviceToHost);
    if (hidden_out) cudaMemcpy(hidden_out, s->hidden, H * sizeof(float), cudaMemcpyDeviceToHost);
    return 1;
}

// forward_text (src/tts_transformer.cpp:1922-1950): rows of the folded text-projection table (project_text_tokens,
// :1026-1091) + the speaker row, then the prefill pass
int qtts_forward_text(qtts_engine * h, int slot, const int32_t * tokens, int32_t n_tokens, const float * speaker,
                      int32_t n_past, float * hidden_out) {
    Engine * e = &h->e;
    Slot * s = get_slot(e, slot);
    if (!s) return 0;
    cudaSetDevice(e->device);
    if (!tokens) { e->err = "text_tokens is null"; return 0; }
    if (n_tokens <= 0) { e->err = "n_tokens must be > 0"; return 0; }
    if (n_tokens > 64) { e->err = "forward_text: at most 64 tokens per call"; return 0; }
    const qtts_config & c = e->cfg;
    const size_t H = c.hidden_size;
    for (int i = 0; i < n_tokens; ++i)
        if (tokens[i] < 0 || tokens[i] >= c.text_vocab_size) { e->err = "text token id out of range"; return 0; }
    if (cudaMemcpyAsync(s->tokens, tokens, n_tokens * sizeof(int32_t), cudaMemcpyHostToDevice, e->stream) != cudaSuccess) { e->err = "H2D failed"; return 0; }
    if (speaker) cudaMemcpyAsync(s->speaker, speaker, H * sizeof(float), cudaMemcpyHostToDevice, e->stream);
    cudaStreamSynchronize(e->stream);
    std::vector<Op> P;
    for (int t = 0; t < n_tokens; ++t) {
        float * dst = s->prefill + (size_t) t * H;
        if (e->text_tab) {
            // row of the F32 table selected on the host (ids were validated above) + the speaker row
            Op o;
            memset(&o, 0, sizeof(o));
            o.type = OP_COPY; o.tok_slot = -1; o.n_tok = 1; o.temperature_override = -1e30f;
            o.out = dst; o.n = (int) H;
            o.a = e->text_tab + (size_t) tokens[t] * H; o.b = speaker ? s->speaker : nullptr;
            P.push_back(o);
        } else {
            e->err = "forward_text needs the folded text table (QTTS_FOLD_TEXT=0 disables it)"; return 0;
        }
    }
    if (!run_program(e, *s, P, 1, false)) return 0;
    if (!do_forward_prefill(e, *s, n_tokens, n_past, hidden_out != nullptr && n_tokens > 1)) return 0;
    if (hidden_out) {
        if (n_tokens > 1) cudaMemcpy(hidden_out, s->prefill_hidden, (size_t) (n_tokens - 1) * H * sizeof(float), cudaMemcpyDeviceToHost);
        cudaMemcpy(hidden_out + (size_t) (n_tokens - 1) * H, s->hidden, H * sizeof(float), cudaMemcpyDeviceToHost);
    }
    cudaError_t r = cudaGetLastError();
    if (r != cudaSuccess) { e->err = std::string("forward_text: ") + cudaGetErrorString(r); return 0; }
    return 1;
}

int qtts_get_hidden_states(qtts_engine * h, int slot, float * hidden_out) {
    Engine * e = &h->e;
    Slot * s = get_slot(e, slot);
    if (!s || !hidden_out) return 0;
    cudaSetDevice(e->device);
    cudaStreamSynchronize(e->stream);
    if (cudaMemcpy(hidden_out, s->hidden, e->cfg.hidden_size * sizeof(float), cudaMemcpyDeviceToHost) != cudaSuccess) { e->err = "get_hidden_states: copy failed"; return 0; }
    return 1;
}

// forward_codec (src/tts_transformer.cpp:2030-2040)
int qtts_forward_codec(qtts_engine * h, int slot, int32_t codec_token, int32_t n_past, float * logits_out) {
    Engine * e = &h->e;
    Slot * s = get_slot(e, slot);
    if (!s) return 0;
    cudaSetDevice(e->device);
    const qtts_config & c = e->cfg;
    if (codec_token < 0 || codec_token >= c.codec_vocab_size) { e->err = "forward_codec: token out of range"; return 0; }
    if (n_past < 0) { e->err = "forward_codec: n_past must be >= 0"; return 0; }
    if (n_past + 1 > e->max_ctx) { e->err = "Context length exceeded"; return 0; }
    cudaMemcpyAsync(&s->tokens[0], &codec_token, sizeof(int32_t), cudaMemcpyHostToDevice, e->stream);
    cudaStreamSynchronize(e->stream);
    std::vector<Op> P;
    const int mode = (e->numerics == QTTS_NUMERICS_REFERENCE) ? ATTN_FLASH_F16SEQ : ATTN_FAST;
    emit_talker_step(e, *s, P, IN_EMBED_ROW, nullptr, false, n_past, mode, true, false);
    P[0].table = e->codec_embd; P[0].tok_ptr = s->tokens; P[0].tok_slot = -1;
    if (!run_program(e, *s, P, 1, false)) return 0;
    const int np = n_past + 1;
    cudaMemcpyAsync(&s->st->n_past, &np, sizeof(int), cudaMemcpyHostToDevice, e->stream);
    cudaError_t r = cudaStreamSynchronize(e->stream);
    if (r != cudaSuccess) { e->err = std::string("forward_codec: ") + cudaGetErrorString(r); return 0; }
    if (logits_out) cudaMemcpy(logits_out, s->logits, c.codec_vocab_size * sizeof(float), cudaMemcpyDeviceToHost);
    return 1;
}

int qtts_predict_codes(qtts_engine * h, int slot, const float * hidden, int32_t cb0, float temperature, int32_t top_k,
                       const float * uniforms15, const int32_t * forced15, int32_t * out15, float * logits_out) {
    Engine * e = &h->e;
    Slot * s = get_slot(e, slot);
    if (!s) return 0;
    cudaSetDevice(e->device);
    const qtts_config & c = e->cfg;
    if (cb0 < 0 || cb0 >= c.codec_vocab_size) { e->err = "predict_codes: codebook-0 token out of range"; return 0; }
    const int nb = c.n_codebooks - 1;
    cudaMemcpyAsync(s->cp_hidden_in, hidden, c.hidden_size * sizeof(float), cudaMemcpyHostToDevice, e->stream);
    cudaMemcpyAsync(&s->st->feed_codes[0], &cb0, sizeof(int), cudaMemcpyHostToDevice, e->stream);
    cudaMemcpyAsync(&s->st->cur_codes[0], &cb0, sizeof(int), cudaMemcpyHostToDevice, e->stream);
    if (uniforms15) cudaMemcpyAsync(s->uniforms, uniforms15, nb * sizeof(float), cudaMemcpyHostToDevice, e->stream);
    if (forced15) cudaMemcpyAsync(s->forced, forced15, nb * sizeof(int32_t), cudaMemcpyHostToDevice, e->stream);
    std::vector<Op> P;
    emit_code_predictor(e, *s, P, s->cp_hidden_in, logits_out ? s->cp_dump : nullptr, forced15 ? s->forced : nullptr,
                        uniforms15 ? s->uniforms : nullptr, true, temperature, top_k);
    if (!run_program(e, *s, P, 1, false)) return 0;
    cudaError_t r = cudaStreamSynchronize(e->stream);
    if (r != cudaSuccess) { e->err = std::string("predict_codes: ") + cudaGetErrorString(r); return 0; }
    StreamState hs;
    cudaMemcpy(&hs, s->st, sizeof(hs), cudaMemcpyDeviceToHost);
    for (int i = 0; i < nb; ++i) out15[i] = hs.cur_codes[i + 1];
    if (logits_out) cudaMemcpy(logits_out, s->cp_dump, (size_t) nb * c.cp_vocab_size * sizeof(float), cudaMemcpyDeviceToHost);
    return 1;
}

// TRTCodePredictor::step / step_with_token (src/trt_code_predictor.cpp:384-482): ONE code-predictor position.
// input != NULL: an F32 vector of input_size floats (hidden_size when needs_proj and the model has mtp_proj, else
// cp_hidden_size); input == NULL: row token_id of embedding table embedding_idx (0 = talker codec_embd,
// 1..14 = code_pred.codec_embd[0..13], :46 of the header), always projected when the model has mtp_proj.
// lm_head_idx in [0, 15): output = logits [cp_vocab]; < 0: output = normalised hidden [cp_hidden] (:342-382).
int qtts_cp_step(qtts_engine * h, int slot, const float * input, int32_t input_size, int32_t token_id, int32_t embedding_idx,
                 int32_t position_id, int32_t lm_head_idx, int needs_proj, float * output, int32_t output_size) {
    Engine * e = &h->e;
    Slot * s = get_slot(e, slot);
    if (!s) return 0;
    cudaSetDevice(e->device);
    const qtts_config & c = e->cfg;
    if (position_id < 0 || position_id >= 16) { e->err = "cp_step: position_id must be in [0, 16)"; return 0; }
    if (lm_head_idx >= c.n_codebooks - 1) { e->err = "cp_step: lm_head_idx out of range"; return 0; }
    const int want_out = lm_head_idx >= 0 ? c.cp_vocab_size : c.cp_hidden_size;
    if (!output || output_size < want_out) { e->err = "cp_step: output buffer too small"; return 0; }
    const __half * table = nullptr;
    bool proj;
    if (input) {
        proj = needs_proj != 0 && e->mtp != nullptr;
        const int want_in = proj ? c.hidden_size : c.cp_hidden_size;
        if (input_size != want_in) { e->err = "cp_step: input_size does not match the model"; return 0; }
        cudaMemcpyAsync(s->cp_hidden_in, input, (size_t) want_in * sizeof(float), cudaMemcpyHostToDevice, e->stream);
    } else {
        if (embedding_idx < 0 || embedding_idx >= c.n_codebooks - 1) { e->err = "Invalid embedding index: " + std::to_string(embedding_idx); return 0; }
        const int vocab = embedding_idx == 0 ? c.codec_vocab_size : c.cp_vocab_size;
        if (token_id < 0 || token_id >= vocab) { e->err = "Token ID out of range: " + std::to_string(token_id); return 0; }
        table = embedding_idx == 0 ? e->codec_embd : e->cp_embd[embedding_idx - 1];
        proj = e->mtp != nullptr;
        if (!proj && c.cp_hidden_size != c.hidden_size) { e->err = "cp_step: embedding rows need mtp_proj"; return 0; }
        cudaMemcpyAsync(&s->tokens[0], &token_id, sizeof(int32_t), cudaMemcpyHostToDevice, e->stream);
    }
    cudaStreamSynchronize(e->stream);
    std::vector<Op> P;
    emit_cp_step(e, *s, P, s->cp_hidden_in, table, s->tokens, position_id, lm_head_idx, proj, s->misc);
    if (!run_program(e, *s, P, 1, false)) return 0;
    cudaError_t r = cudaStreamSynchronize(e->stream);
    if (r != cudaSuccess) { e->err = std::string("cp_step: ") + cudaGetErrorString(r); return 0; }
    cudaMemcpy(output, lm_head_idx >= 0 ? s->cp_logits : s->misc, (size_t) want_out * sizeof(float), cudaMemcpyDeviceToHost);
    return 1;
}

// embeddings (A12) + per-stream state; the Talker prefill (A3) is left to the caller
static bool stream_prepare(Engine * e, Slot & s, const int32_t * tokens, int32_t n, const float * speaker, int32_t max_len,
                           int32_t language_id, float rep_pen, float temperature, int32_t top_k, uint64_t seed) {
    if (!tokens) return e->fail("text_tokens is null");
    if (n < 4) return e->fail("Need at least 4 text tokens for generation");
    s.active = false; s.pending = false;
    if (max_len <= 0) {
        // generate() returns an empty output for max_len <= 0 (src/tts_transformer.cpp:2364-2367): the stream is born done
        StreamState st{};
        st.done = 1;
        QTTS_CUDA(cudaMemcpyAsync(s.st, &st, sizeof(st), cudaMemcpyHostToDevice, e->stream));
        QTTS_CUDA(cudaStreamSynchronize(e->stream));
        s.prefill_len = 0;
        return true;
    }
    if (!do_build_prefill(e, s, tokens, n, speaker, language_id)) return false;
    // required context as in the reference (src/tts_transformer.cpp:2383)
    if (s.prefill_len + max_len + 8 > e->max_ctx) return e->fail("max_len does not fit the engine's max_ctx");
    StreamState st{};
    st.max_len = max_len; st.n_text_tokens = n; st.trailing_len = s.trailing_len;
    st.top_k = top_k; st.rep_pen = rep_pen; st.temperature = temperature; st.seed = seed;
    st.ignore_eos = e->ignore_eos;
    st.sample_mode = (e->numerics == QTTS_NUMERICS_REFERENCE) ? 1 : 0;
    QTTS_CUDA(cudaMemcpyAsync(s.st, &st, sizeof(st), cudaMemcpyHostToDevice, e->stream));
    QTTS_CUDA(cudaStreamSynchronize(e->stream));
    return true;
}

static bool stream_begin(Engine * e, Slot & s, const int32_t * tokens, int32_t n, const float * speaker, int32_t max_len,
                         int32_t language_id, float rep_pen, float temperature, int32_t top_k, uint64_t seed) {
    if (!stream_prepare(e, s, tokens, n, speaker, max_len, language_id, rep_pen, temperature, top_k, seed)) return false;
    if (max_len <= 0) return true;
    if (!do_forward_prefill(e, s, s.prefill_len, 0)) return false;
    s.active = true;
    return true;
}

// Talker prefill of every prepared stream: per lock-step group, all of them together on the batched kernel (one
// weight pass per prefill position for the whole group); the batch-1 kernel, stream after stream, when the batched
// path is switched off.  The streams become active.
static bool prefill_pending(Engine * e) {
    bool any = false;
    for (const Slot & s : e->slots) any |= s.pending;
    if (!any) return true;
    if (e->batched && e->batched_ready) {
        for (BGroup & G : e->groups) {
            std::vector<const float *> rows(G.n, nullptr);
            std::vector<int32_t> plen(G.n, 0);
            bool some = false;
            for (int i = 0; i < G.n; ++i) {
                const Slot & s = e->slots[G.s0 + i];
                if (s.pending) { rows[i] = s.prefill; plen[i] = s.prefill_len; some = true; }
            }
            if (some && !run_batched_prefill(e, G, rows, plen)) return false;
        }
        QTTS_CUDA(cudaStreamSynchronize(e->stream));
    } else {
        for (Slot & s : e->slots)
            if (s.pending && !do_forward_prefill(e, s, s.prefill_len, 0)) return false;
    }
    for (Slot & s : e->slots) if (s.pending) { s.pending = false; s.active = true; }
    return true;
}

int qtts_generate(qtts_engine * h, int slot, const int32_t * tokens, int32_t n, const float * speaker, int32_t max_len,
                  int32_t language_id, float rep_pen, float temperature, int32_t top_k, uint64_t seed,
                  int32_t * codes_out, int32_t * n_frames_out, qtts_frame_cb on_frames, int32_t cb_interval, void * user) {
    Engine * e = &h->e;
    Slot * s = get_slot(e, slot);
    if (!s) return 0;
    cudaSetDevice(e->device);
    *n_frames_out = 0;
    if (max_len <= 0) return 1;                                   // src/tts_transformer.cpp:2364-2367
    if (!stream_begin(e, *s, tokens, n, speaker, max_len, language_id, rep_pen, temperature, top_k, seed)) return 0;
    const int nb = e->cfg.n_codebooks;
    const int chunk = cb_interval > 0 ? cb_interval : 40;
    int produced = 0;
    bool stop = false;
    cudaEventRecord(e->ev0, e->stream);
    while (!stop) {
        if (!run_device_program(e, *s, s->frame_prog, &s->frame_host, s->frame_prog_n, s->frame_sched, s->frame_sched_n, chunk, true)) return 0;
        StreamState hs;
        if (cudaMemcpyAsync(&hs, s->st, sizeof(hs), cudaMemcpyDeviceToHost, e->stream) != cudaSuccess ||
            cudaStreamSynchronize(e->stream) != cudaSuccess) {
            e->err = std::string("generate: ") + cudaGetErrorString(cudaGetLastError());
            return 0;
        }
        const int newf = hs.frame - produced;
        if (newf > 0) {
            cudaMemcpy(codes_out + (size_t) produced * nb, s->codes + (size_t) produced * nb,
                       (size_t) newf * nb * sizeof(int32_t), cudaMemcpyDeviceToHost);
            // callback contract (src/tts_transformer.cpp:2517-2523, 2563-2570): full chunks of
            // cb_interval frames while generating, the remainder at the end
            if (on_frames && cb_interval > 0) {
                if (newf == cb_interval) {
                    if (!on_frames(codes_out + (size_t) produced * nb, newf, nb, user)) stop = true;
                } else if (hs.done) {
                    on_frames(codes_out + (size_t) produced * nb, newf, nb, user);
                }
            }
            produced = hs.frame;
        }
        if (hs.done) stop = true;
    }
    cudaEventRecord(e->ev1, e->stream);
    cudaEventSynchronize(e->ev1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e->ev0, e->ev1);
    e->last_loop_ms = ms;
    *n_frames_out = produced;
    s->active = false;
    return 1;
}

int qtts_stream_begin(qtts_engine * h, int slot, const int32_t * tokens, int32_t n, const float * speaker, int32_t max_len,
                      int32_t language_id, float rep_pen, float temperature, int32_t top_k, uint64_t seed) {
    Engine * e = &h->e;
    Slot * s = get_slot(e, slot);
    if (!s) return 0;
    cudaSetDevice(e->device);
    return stream_begin(e, *s, tokens, n, speaker, max_len, language_id, rep_pen, temperature, top_k, seed) ? 1 : 0;
}

int qtts_stream_prepare(qtts_engine * h, int slot, const int32_t * tokens, int32_t n, const float * speaker, int32_t max_len,
                        int32_t language_id, float rep_pen, float temperature, int32_t top_k, uint64_t seed) {
    Engine * e = &h->e;
    Slot * s = get_slot(e, slot);
    if (!s) return 0;
    cudaSetDevice(e->device);
    if (!stream_prepare(e, *s, tokens, n, speaker, max_len, language_id, rep_pen, temperature, top_k, seed)) return 0;
    s->pending = max_len > 0;
    return 1;
}

int qtts_prefill_pending(qtts_engine * h) {
    Engine * e = &h->e;
    if (!e->finalized) { e->err = "engine not finalized"; return 0; }
    cudaSetDevice(e->device);
    cudaEventRecord(e->ev0, e->stream);
    if (!prefill_pending(e)) return 0;
    cudaEventRecord(e->ev1, e->stream);
    if (cudaEventSynchronize(e->ev1) != cudaSuccess) { e->err = std::string("prefill_pending: ") + cudaGetErrorString(cudaGetLastError()); return 0; }
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e->ev0, e->ev1);
    e->last_loop_ms = ms;
    return 1;
}

// frames emitted so far by every stream, before the launch (device to device: no host round trip)
__global__ void k_snapshot_frames(const StreamState * __restrict__ st, int n, int32_t * __restrict__ frame0) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) frame0[i] = st[i].frame;
}

// After the launch: the frames every stream produced, out of its device-resident code ring into ONE buffer laid out
// like the caller's [slot][n_frames][n_codebooks] (unproduced rows zero), plus {produced, done} per slot.  One
// gather + one copy to the host replaces one cudaMemcpyAsync per slot (512 of them at BASELINE configs[3]).
__global__ void k_gather_codes(const StreamState * __restrict__ st, const int32_t * __restrict__ frame0,
                               const int32_t * __restrict__ ring, size_t ring_stride, int nb, int n_frames,
                               int32_t * __restrict__ out, int32_t * __restrict__ meta) {
    const int i = blockIdx.x;
    const int f0 = frame0[i], newf = min(max(st[i].frame - f0, 0), n_frames);
    if (threadIdx.x == 0) { meta[2 * i] = newf; meta[2 * i + 1] = st[i].done; }
    const int32_t * src = ring + (size_t) i * ring_stride + (size_t) f0 * nb;
    int32_t * dst = out + (size_t) i * n_frames * nb;
    for (int k = threadIdx.x; k < n_frames * nb; k += blockDim.x) dst[k] = (k < newf * nb) ? src[k] : 0;
}

// int32 [frames][n_codebooks] ring rows -> the TRT vocoder's input layout, int64 [1][n_codebooks][fixed_frames] with
// the frames beyond n zero (src/trt_vocoder.cpp:110-121)
__global__ void k_codes_chunk(const int32_t * __restrict__ ring_rows, int nb, int n, int fixed, long long * __restrict__ out) {
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < nb * fixed; k += gridDim.x * blockDim.x) {
        const int c = k / fixed, f = k - c * fixed;
        out[k] = (f < n) ? (long long) ring_rows[(size_t) f * nb + c] : 0ll;
    }
}

static bool ensure_step_buffers(Engine * e, int n_frames) {
    const size_t ns = e->slots.size();
    const size_t need = ns * (size_t) n_frames * e->cfg.n_codebooks + 2 * ns;      // int32 words: codes + meta
    if (need <= e->step_words && e->step_dev) return true;
    if (e->step_host) { cudaFreeHost(e->step_host); e->step_host = nullptr; }
    // (an older, smaller device buffer stays in stream_allocs until the next resize / destroy)
    e->step_dev = e->dalloc<int32_t>(need + ns, false);
    if (!e->step_dev || cudaMallocHost((void **) &e->step_host, need * sizeof(int32_t)) != cudaSuccess) {
        cudaGetLastError();
        e->step_dev = nullptr; e->step_host = nullptr; e->step_words = 0;
        return e->fail("out of memory for the step_frames staging buffers");
    }
    e->step_words = need;
    return true;
}

// Batched mode (default): all active streams of a lock-step group (<= 256 slots) advance inside ONE launch of the
// tcgen05 frame kernel (qtts_batched.cu), weights streamed once per frame for the whole group.  Sequential mode
// (qtts_set_batched(e, 0)): the batch-1 kernel, one stream after another (kept as the A/B reference of the parity tests).
int qtts_step_frames(qtts_engine * h, int32_t n_frames, int32_t * codes_out, int32_t * produced, int32_t * active) {
    Engine * e = &h->e;
    if (!e->finalized) { e->err = "engine not finalized"; return 0; }
    if (n_frames <= 0 || n_frames > e->max_ctx) { e->err = "step_frames: n_frames out of range"; return 0; }
    cudaSetDevice(e->device);
    if (!ensure_step_buffers(e, n_frames)) return 0;
    // the timed region starts before the batched prefill of streams prepared since the last call (they join here):
    // qtts_last_loop_ms of such a call includes it
    cudaEventRecord(e->ev0, e->stream);
    if (!prefill_pending(e)) return 0;
    const int nb = e->cfg.n_codebooks;
    const size_t ns = e->slots.size();
    int32_t * d_codes = e->step_dev, * d_meta = d_codes + ns * (size_t) n_frames * nb, * d_frame0 = d_meta + 2 * ns;
    k_snapshot_frames<<<(unsigned) ((ns + 255) / 256), 256, 0, e->stream>>>(e->bb.st, (int) ns, d_frame0);
    if (e->batched && e->batched_ready) {
        for (auto & G : e->groups) if (!run_batched_group(e, G, n_frames)) return 0;
    } else {
        for (size_t i = 0; i < ns; ++i) {
            Slot & s = e->slots[i];
            if (!s.active) continue;
            if (!run_device_program(e, s, s.frame_prog, &s.frame_host, s.frame_prog_n, s.frame_sched, s.frame_sched_n, n_frames, true)) return 0;
        }
    }
    cudaEventRecord(e->ev1, e->stream);
    k_gather_codes<<<(unsigned) ns, 128, 0, e->stream>>>(e->bb.st, d_frame0, e->bb.codes, e->bb.codes_stride, nb, n_frames, d_codes, d_meta);
    e->launches += 2;
    const size_t words = ns * (size_t) n_frames * nb + 2 * ns;
    if (cudaMemcpyAsync(e->step_host, d_codes, words * sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream) != cudaSuccess ||
        cudaStreamSynchronize(e->stream) != cudaSuccess) {
        e->err = std::string("step_frames: ") + cudaGetErrorString(cudaGetLastError());
        return 0;
    }
    const int32_t * meta = e->step_host + ns * (size_t) n_frames * nb;
    memcpy(codes_out, e->step_host, ns * (size_t) n_frames * nb * sizeof(int32_t));
    for (size_t i = 0; i < ns; ++i) {
        Slot & s = e->slots[i];
        if (!s.active) {
            produced[i] = 0; active[i] = 0;
            memset(codes_out + i * (size_t) n_frames * nb, 0, (size_t) n_frames * nb * sizeof(int32_t));
            continue;
        }
        produced[i] = meta[2 * i];
        if (meta[2 * i + 1]) s.active = false;
        active[i] = s.active ? 1 : 0;
    }
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e->ev0, e->ev1);
    e->last_loop_ms = ms;
    return 1;
}

// N4 (vocoder hand-off): frames [first_frame, first_frame + n_frames) of `slot` straight out of the device-resident
// code ring in the layout the reference's streamed vocoder consumes - int64 [1][n_codebooks][fixed_frames], frames
// beyond n_frames zero (TRTVocoderDecoder::decode, src/trt_vocoder.cpp:98-170, fed by the 40-frame callback of
// generate(), src/tts_transformer.cpp:2517-2523, src/qwen3_tts.cpp:437-451).  out_is_device = 1: `out` is a device
// pointer (a TensorRT binding: the chunk never visits the host); 0: host memory.
int qtts_codes_chunk(qtts_engine * h, int slot, int32_t first_frame, int32_t n_frames, int32_t fixed_frames,
                     int64_t * out, int out_is_device) {
    Engine * e = &h->e;
    Slot * s = get_slot(e, slot);
    if (!s) return 0;
    cudaSetDevice(e->device);
    const int nb = e->cfg.n_codebooks;
    if (!out || fixed_frames <= 0 || n_frames < 0 || n_frames > fixed_frames || first_frame < 0 || first_frame + n_frames > e->max_ctx) {
        e->err = "codes_chunk: bad frame range"; return 0;
    }
    int32_t have = 0;
    cudaMemcpyAsync(&have, &s->st->frame, sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream);
    cudaStreamSynchronize(e->stream);
    if (first_frame + n_frames > have) { e->err = "codes_chunk: frames not generated yet"; return 0; }
    const size_t words = (size_t) nb * fixed_frames;
    long long * dst = (long long *) out;
    if (!out_is_device) {
        if (words > e->chunk_words) {
            e->chunk_dev = e->dalloc<long long>(words, false);
            if (!e->chunk_dev) { e->chunk_words = 0; e->err = "out of device memory for the code chunk"; return 0; }
            e->chunk_words = words;
        }
        dst = e->chunk_dev;
    }
    k_codes_chunk<<<(unsigned) ((words + 255) / 256), 256, 0, e->stream>>>(s->codes + (size_t) first_frame * nb, nb, n_frames, fixed_frames, dst);
    e->launches += 1;
    if (!out_is_device) cudaMemcpyAsync(out, dst, words * sizeof(long long), cudaMemcpyDeviceToHost, e->stream);
    if (cudaStreamSynchronize(e->stream) != cudaSuccess) { e->err = std::string("codes_chunk: ") + cudaGetErrorString(cudaGetLastError()); return 0; }
    return 1;
}

// ------------------------------------------------------------------------------------------
// qtts_codes_chunk_noalloc: the host-destination form of qtts_codes_chunk for callers that must not allocate (code
// under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// qtts_codes_chunk stages the chunk in a device buffer it cudaMalloc's
// on first use and regrows on demand; here the staging buffer is the file-scope array below: QTTS_NOALLOC_CHUNK_WORDS
// int64 words (64 KB) that every GPU which loads the library holds until the process exits.
//   * In host code `g_chunk_scratch` is a host-side stand-in, NOT a device address; the kernel gets the address
//     cudaGetSymbolAddress returns, looked up once per device with that device current (chunk_scratch_for).
//   * Every call on a GPU - whatever engine or thread it comes from - shares the array, and the array keeps the last
//     call's contents.  Nothing needs clearing: k_codes_chunk writes all n_codebooks x fixed_frames words the copy then
//     reads (frames beyond n_frames as zeros).  Calls that overlap on one GPU race on it; the call itself is
//     synchronous (it returns after its copy has landed), so one thread per GPU, or a lock around the call, is safe.
//   * No cudaMalloc / cudaFree / allocating library call on any path; sizes beyond the array are rejected.
// ------------------------------------------------------------------------------------------
static __device__ __align__(256) long long g_chunk_scratch[QTTS_NOALLOC_CHUNK_WORDS];

constexpr int kScratchMaxDevices = 64;
static long long * g_chunk_scratch_ptr[kScratchMaxDevices];      // device address of g_chunk_scratch on device d (or null)
static std::mutex g_chunk_scratch_mu;

// device must be current
static long long * chunk_scratch_for(Engine * e) {
    if (e->device < 0 || e->device >= kScratchMaxDevices) { e->err = "codes_chunk_noalloc: device index out of range"; return nullptr; }
    std::lock_guard<std::mutex> lk(g_chunk_scratch_mu);
    if (!g_chunk_scratch_ptr[e->device]) {
        long long * p = nullptr;
        const cudaError_t rc = cudaGetSymbolAddress((void **) &p, g_chunk_scratch);
        if (rc != cudaSuccess || !p) {
            e->err = std::string("codes_chunk_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(rc);
            cudaGetLastError();
            return nullptr;
        }
        g_chunk_scratch_ptr[e->device] = p;
    }
    return g_chunk_scratch_ptr[e->device];
}

int qtts_codes_chunk_noalloc(qtts_engine * h, int slot, int32_t first_frame, int32_t n_frames, int32_t fixed_frames, int64_t * out_host) {
    Engine * e = &h->e;
    Slot * s = get_slot(e, slot);
    if (!s) return 0;
    if (cudaSetDevice(e->device) != cudaSuccess) { e->err = "codes_chunk_noalloc: cudaSetDevice failed"; cudaGetLastError(); return 0; }
    const int nb = e->cfg.n_codebooks;
    if (!out_host || fixed_frames <= 0 || n_frames < 0 || n_frames > fixed_frames || first_frame < 0 || first_frame + n_frames > e->max_ctx) {
        e->err = "codes_chunk_noalloc: bad frame range"; return 0;
    }
    const size_t words = (size_t) nb * (size_t) fixed_frames;
    if (words > (size_t) QTTS_NOALLOC_CHUNK_WORDS) {
        e->err = "codes_chunk_noalloc: n_codebooks * fixed_frames = " + std::to_string(words) + " int64 words exceeds the static scratch of " +
                 std::to_string((int) QTTS_NOALLOC_CHUNK_WORDS) + " (QTTS_NOALLOC_CHUNK_WORDS); use qtts_codes_chunk or smaller chunks";
        return 0;
    }
    long long * scratch = chunk_scratch_for(e);
    if (!scratch) return 0;
    int32_t have = 0;
    cudaMemcpyAsync(&have, &s->st->frame, sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream);
    if (cudaStreamSynchronize(e->stream) != cudaSuccess) { e->err = std::string("codes_chunk_noalloc: ") + cudaGetErrorString(cudaGetLastError()); return 0; }
    if (first_frame + n_frames > have) { e->err = "codes_chunk_noalloc: frames not generated yet"; return 0; }
    k_codes_chunk<<<(unsigned) ((words + 255) / 256), 256, 0, e->stream>>>(s->codes + (size_t) first_frame * nb, nb, n_frames, fixed_frames, scratch);
    e->launches += 1;
    cudaMemcpyAsync(out_host, scratch, words * sizeof(long long), cudaMemcpyDeviceToHost, e->stream);
    if (cudaStreamSynchronize(e->stream) != cudaSuccess) { e->err = std::string("codes_chunk_noalloc: ") + cudaGetErrorString(cudaGetLastError()); return 0; }
    return 1;
}

int qtts_stream_force(qtts_engine * h, int slot, const int32_t * forced, int32_t n_forced, const float * uniforms, int32_t n_uniform) {
    Engine * e = &h->e;
    Slot * s = get_slot(e, slot);
    if (!s) return 0;
    cudaSetDevice(e->device);
    if (!s->active && !s->pending) { e->err = "stream_force: the stream has not been begun / prepared"; return 0; }
    const int nb = e->cfg.n_codebooks;
    if (n_forced < 0 || n_uniform < 0 || n_forced > e->max_ctx || n_uniform > e->max_ctx) { e->err = "stream_force: frame count out of range"; return 0; }
    if (forced && n_forced > 0) {
        for (int i = 0; i < n_forced * nb; ++i) {
            const int lim = (i % nb == 0) ? e->cfg.codec_vocab_size : e->cfg.cp_vocab_size;
            if (forced[i] < 0 || forced[i] >= lim) { e->err = "stream_force: code out of range"; return 0; }
        }
        if (!s->force_codes) s->force_codes = e->dalloc<int32_t>((size_t) e->max_ctx * nb);
        if (!s->force_codes) { e->err = "out of device memory for the forced codes"; return 0; }
        cudaMemcpyAsync(s->force_codes, forced, (size_t) n_forced * nb * sizeof(int32_t), cudaMemcpyHostToDevice, e->stream);
    }
    if (uniforms && n_uniform > 0) {
        if (!s->force_u) s->force_u = e->dalloc<float>((size_t) e->max_ctx * nb);
        if (!s->force_u) { e->err = "out of device memory for the uniforms"; return 0; }
        cudaMemcpyAsync(s->force_u, uniforms, (size_t) n_uniform * nb * sizeof(float), cudaMemcpyHostToDevice, e->stream);
    }
    StreamState hs;
    cudaMemcpyAsync(&hs, s->st, sizeof(hs), cudaMemcpyDeviceToHost, e->stream);
    cudaStreamSynchronize(e->stream);
    hs.forced = (forced && n_forced > 0) ? s->force_codes : nullptr; hs.forced_frames = hs.forced ? n_forced : 0;
    hs.uniforms = (uniforms && n_uniform > 0) ? s->force_u : nullptr; hs.uniform_frames = hs.uniforms ? n_uniform : 0;
    cudaMemcpyAsync(s->st, &hs, sizeof(hs), cudaMemcpyHostToDevice, e->stream);
    if (cudaStreamSynchronize(e->stream) != cudaSuccess) { e->err = std::string("stream_force: ") + cudaGetErrorString(cudaGetLastError()); return 0; }
    return 1;
}

int qtts_sample_cb0(qtts_engine * h, int slot, const float * raw, int32_t frame, int32_t n_text_tokens, const int32_t * seen,
                    int32_t n_seen, float rep_pen, float temperature, int32_t top_k, float u, int32_t * token_out,
                    float * processed_out) {
    Engine * e = &h->e;
    Slot * s = get_slot(e, slot);
    if (!s) return 0;
    cudaSetDevice(e->device);
    const qtts_config & c = e->cfg;
    if (!raw || !token_out) { e->err = "sample_cb0: null argument"; return 0; }
    if (frame < 0 || frame >= e->max_ctx) { e->err = "sample_cb0: frame out of range"; return 0; }
    StreamState st{};
    st.frame = frame; st.max_len = frame + 2; st.n_text_tokens = n_text_tokens; st.top_k = top_k; st.rep_pen = rep_pen;
    st.temperature = temperature; st.sample_mode = (e->numerics == QTTS_NUMERICS_REFERENCE) ? 1 : 0;
    for (int i = 0; i < n_seen; ++i) {
        if (seen[i] < 0 || seen[i] >= c.codec_vocab_size) { e->err = "sample_cb0: seen token out of range"; return 0; }
        st.seen[seen[i] >> 5] |= 1u << (seen[i] & 31);
    }
    s->active = false; s->pending = false;
    cudaMemcpyAsync(s->st, &st, sizeof(st), cudaMemcpyHostToDevice, e->stream);
    cudaMemcpyAsync(s->logits, raw, c.codec_vocab_size * sizeof(float), cudaMemcpyHostToDevice, e->stream);
    cudaMemcpyAsync(s->uniforms + frame, &u, sizeof(float), cudaMemcpyHostToDevice, e->stream);
    cudaStreamSynchronize(e->stream);
    Op o;
    memset(&o, 0, sizeof(o));
    o.type = OP_SAMPLE_CB0; o.n_tok = 1; o.tok_slot = -1; o.temperature_override = -1e30f;
    o.logits = s->logits; o.n_vocab = c.codec_vocab_size; o.eos_id = c.codec_eos_id; o.suppress_start = c.codec_vocab_size - 1024;
    o.dump = s->cb0_dump; o.uniforms = s->uniforms; o.n_books = c.n_codebooks;
    std::vector<Op> P{o};
    if (!run_program(e, *s, P, 1, false)) return 0;
    StreamState hs;
    cudaMemcpyAsync(&hs, s->st, sizeof(hs), cudaMemcpyDeviceToHost, e->stream);
    if (cudaStreamSynchronize(e->stream) != cudaSuccess) { e->err = std::string("sample_cb0: ") + cudaGetErrorString(cudaGetLastError()); return 0; }
    *token_out = hs.eos ? c.codec_eos_id : hs.cur_codes[0];
    if (processed_out) cudaMemcpy(processed_out, s->cb0_dump, c.codec_vocab_size * sizeof(float), cudaMemcpyDeviceToHost);
    return 1;
}

int qtts_set_debug(qtts_engine * h, int flags) {
    Engine * e = &h->e;
    e->debug_flags = flags;
    if (e->finalized) { cudaSetDevice(e->device); cudaStreamSynchronize(e->stream); if (!rebuild_frame_programs(e)) return 0; }
    return 1;
}

int qtts_resize_kv(qtts_engine * h, int32_t n_ctx, int32_t max_text) {
    Engine * e = &h->e;
    cudaSetDevice(e->device);
    return engine_resize(e, n_ctx, max_text) ? 1 : 0;
}

int qtts_set_batched(qtts_engine * h, int on) { h->e.batched = on ? 1 : 0; return 1; }

int qtts_set_batched_weight_format(qtts_engine * h, int fmt) {
    Engine * e = &h->e;
    if (e->finalized) { e->err = "weight format must be chosen before qtts_finalize"; return 0; }
    if (fmt == QTTS_BF16) {
        // measured on B200: tcgen05.mma kind::f16 with a_format = F16 and b_format = BF16 traps with
        // "illegal instruction"; the activations must stay F16 (the reference's rounding point)
        e->err = "batched kernel: BF16 weights cannot be mixed with F16 activations in one tcgen05.mma (kind::f16); use QTTS_F16";
        return 0;
    }
    if (fmt != QTTS_F16) { e->err = "unknown weight format"; return 0; }
    e->bfmt = 0;
    return 1;
}

int qtts_stream_logits(qtts_engine * h, int slot, float * cb0_out, float * cp_out) {
    Engine * e = &h->e;
    Slot * s = get_slot(e, slot);
    if (!s) return 0;
    cudaSetDevice(e->device);
    cudaStreamSynchronize(e->stream);
    const qtts_config & c = e->cfg;
    if (e->batched && e->batched_ready) {
        BGroup & G = e->groups[slot / e->bgroup];
        const int b = slot - G.s0;
        if (cb0_out) cudaMemcpy(cb0_out, G.cb0_dump + (size_t) b * c.codec_vocab_size, c.codec_vocab_size * sizeof(float), cudaMemcpyDeviceToHost);
        if (cp_out)
            for (int p = 0; p < c.n_codebooks - 1; ++p)
                cudaMemcpy(cp_out + (size_t) p * c.cp_vocab_size, G.cp_dump + ((size_t) p * G.n + b) * c.cp_vocab_size,
                           c.cp_vocab_size * sizeof(float), cudaMemcpyDeviceToHost);
    } else {
        if (cb0_out) cudaMemcpy(cb0_out, s->cb0_dump, c.codec_vocab_size * sizeof(float), cudaMemcpyDeviceToHost);
        if (cp_out) {
            if (!(e->debug_flags & 1)) { e->err = "code-predictor logits of the last frame: batch-1 kernel keeps them only with qtts_set_debug(e, 1)"; return 0; }
            cudaMemcpy(cp_out, s->cp_dump, (size_t) (c.n_codebooks - 1) * c.cp_vocab_size * sizeof(float), cudaMemcpyDeviceToHost);
        }
    }
    if (cudaGetLastError() != cudaSuccess) { e->err = "stream_logits: copy failed"; return 0; }
    return 1;
}

int qtts_batched_profile(qtts_engine * h, int on, uint64_t * out64) {
    Engine * e = &h->e;
    if (!e->finalized) { e->err = "engine not finalized"; return 0; }
    cudaSetDevice(e->device);
    cudaStreamSynchronize(e->stream);
    if (out64) cudaMemcpy(out64, e->prof + 2048, 64 * sizeof(uint64_t), cudaMemcpyDeviceToHost);
    cudaMemset(e->prof + 2048, 0, 64 * sizeof(uint64_t));
    cudaStreamSynchronize(0);
    e->prof_on = on ? 1 : 0;
    return 1;
}

// Runs ONE projection stage of the batched kernel on caller-supplied operands (unit test of the
// tcgen05 path: descriptors, operand layouts, epilogues).  W [rows][K] and X [rows_act][K] are
// F32 on the host (W is rounded to BF16, X to F16 when in_f16); norm_w / res / bias may be NULL.
// out receives [rows_act][rows] floats (EPI_SWIGLU: [rows_act][rows/2]); ssq_out (may be NULL)
// [rows_act][rows/16].
int qtts_debug_bgemm(int device, int weight_fmt, int n_streams, int n_tok, int rows, int K, int epi, int in_f16,
                     const float * W, const float * X, const float * norm_w, const float * res, const float * bias,
                     float eps, float * out, float * ssq_out) {
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || device >= n_dev) { g_create_err = "no CUDA device"; return 0; }
    cudaSetDevice(device);
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, device);
    const int rows_act = n_streams * n_tok;
    if (rows_act > kBMaxRows || n_tok < 1 || n_tok > 2 || rows % 16 || K % kKc) { g_create_err = "bgemm: bad shape"; return 0; }
    Engine eng;
    Engine * e = &eng;
    e->device = device; e->n_sms = prop.multiProcessorCount; e->bfmt = (weight_fmt == QTTS_BF16) ? 1 : 0;
    cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking);
    auto cleanup = [&](int rc) {
        cudaStreamSynchronize(e->stream);
        for (void * p : e->allocs) cudaFree(p);
        for (void * p : e->stream_allocs) cudaFree(p);
        cudaStreamDestroy(e->stream);
        if (!rc) g_create_err = e->err;
        return rc;
    };
    if (cudaFuncSetAttribute(qtts_run_batched, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) kBSmemBytes) != cudaSuccess ||
        cudaFuncSetAttribute(qtts_run_batched_ext, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) kBSmemBytes) != cudaSuccess) {
        e->err = "bgemm: cudaFuncSetAttribute failed"; return cleanup(0);
    }
    e->bar = e->dalloc<GridBarrier>(1);
    const size_t nW = (size_t) rows * K, nX = (size_t) rows_act * K;
    std::vector<wt_t> wb(nW);       // W rounded to BF16 (the synthetic fixtures' precision), held as its exact F16 image
    for (size_t i = 0; i < nW; ++i) wb[i] = __float2half_rn(__bfloat162float(__float2bfloat16_rn(W[i])));
    wt_t * dW = e->dalloc<wt_t>(nW, false);
    cudaMemcpy(dW, wb.data(), nW * 2, cudaMemcpyHostToDevice);
    TiledW T;
    if (!tile_matrix(e, dW, rows, K, T, in_f16 != 0 && epi == EPI_ADD)) return cleanup(0);
    void * dX = nullptr;
    if (in_f16) {
        // F16 activations live in global memory as UMMA-canonical 128-column chunks (qtts_batched.cu: canon_off)
        const size_t cb = (size_t) ((rows_act + 7) / 8) * 2048, nXc = (size_t) (K / 128) * cb / 2;
        std::vector<__half> xh(nXc);
        for (int r = 0; r < rows_act; ++r)
            for (int k = 0; k < K; ++k)
                xh[((size_t) (k >> 7) * cb + (size_t) (r >> 3) * 2048 + (size_t) ((k & 127) >> 3) * 128 + (r & 7) * 16 + (k & 7) * 2) / 2] =
                    __float2half_rn(X[(size_t) r * K + k]);
        dX = e->dalloc<__half>(nXc, false);
        cudaMemcpy(dX, xh.data(), nXc * 2, cudaMemcpyHostToDevice);
    } else {
        dX = e->dalloc<float>(nX, false);
        cudaMemcpy(dX, X, nX * 4, cudaMemcpyHostToDevice);
    }
    const int outw = (epi == EPI_SWIGLU) ? rows / 2 : rows;
    float * d_out = e->dalloc<float>((size_t) rows_act * rows);
    __half * d_out16 = e->dalloc<__half>((size_t) (rows_act + 8) * rows);
    float * d_res = nullptr, * d_bias = nullptr, * d_norm = nullptr, * d_ssq_in = nullptr;
    float * d_ssq = e->dalloc<float>((size_t) rows_act * (rows / 16));
    if (res) { d_res = e->dalloc<float>((size_t) rows_act * rows, false); cudaMemcpy(d_res, res, (size_t) rows_act * rows * 4, cudaMemcpyHostToDevice); }
    if (bias) { d_bias = e->dalloc<float>(rows, false); cudaMemcpy(d_bias, bias, rows * 4, cudaMemcpyHostToDevice); }
    if (norm_w) {
        d_norm = e->dalloc<float>(K, false); cudaMemcpy(d_norm, norm_w, K * 4, cudaMemcpyHostToDevice);
        // sums of squares split over 4 partials to exercise the reduction
        std::vector<float> ss((size_t) rows_act * 4, 0.f);
        for (int r = 0; r < rows_act; ++r)
            for (int k = 0; k < K; ++k) ss[(size_t) r * 4 + (k * 4) / K] += X[(size_t) r * K + k] * X[(size_t) r * K + k];
        d_ssq_in = e->dalloc<float>(ss.size(), false); cudaMemcpy(d_ssq_in, ss.data(), ss.size() * 4, cudaMemcpyHostToDevice);
    }
    BOp o;
    memset(&o, 0, sizeof(o));
    o.type = BOP_GEMM; o.n_streams = n_streams; o.rows_act = rows_act; o.n_tok = n_tok;
    o.flags = ssq_out ? BF_SSQ_OUT : 0;
    o.Wt = T.p; o.rows = rows; o.K = T.K; o.n_tile = T.n_tile; o.n_tiles = T.n_tiles; o.cta_off = 0; o.kpp = T.kpp; o.n_pieces = T.n_pieces;
    o.ksplit = T.ksplit;
    if (T.ksplit > 1) o.part_ll = e->dalloc<unsigned long long>((size_t) rows * kBMaxRows);
    o.in_kind = in_f16 ? BIN_F16 : BIN_F32; o.in = dX; o.in_row_stride = K; o.in_tok_stride = n_streams * K;
    o.norm_w = d_norm; o.ssq_in = d_ssq_in; o.n_parts_in = 4; o.eps = eps;
    o.epi = epi; o.out32 = d_out; o.out16 = d_out16; o.res = d_res; o.bias = d_bias;
    o.out_row_stride = outw; o.out_tok_stride = n_streams * outw; o.ssq_out = d_ssq;
    BSched sc{T.p, T.n_tile * T.K * 2, T.n_tile * T.kpp * 2, T.n_pieces, T.n_tiles, 0, 0};
    BOp * d_ops = e->dalloc<BOp>(1, false);
    BSched * d_sc = e->dalloc<BSched>(1, false);
    StreamState * d_st = e->dalloc<StreamState>(n_streams);
    int32_t * d_act = e->dalloc<int32_t>(n_streams);
    cudaMemcpy(d_ops, &o, sizeof(o), cudaMemcpyHostToDevice);
    cudaMemcpy(d_sc, &sc, sizeof(sc), cudaMemcpyHostToDevice);
    // two "frames": the second pass exercises ring slot / barrier phase recycling
    int iters = 2;
    if (const char * it = getenv("QTTS_BGEMM_ITERS")) iters = std::max(1, atoi(it));
    cudaEvent_t t0, t1;
    cudaEventCreate(&t0); cudaEventCreate(&t1);
    cudaEventRecord(t0, e->stream);
    if (iters > 2) { e->prof = e->dalloc<unsigned long long>(4096); e->prof_on = 1; }
    if (!launch_batched(e, d_ops, 1, d_sc, 1, d_st, d_act, n_streams, iters, false)) return cleanup(0);
    cudaEventRecord(t1, e->stream);
    cudaError_t r = cudaStreamSynchronize(e->stream);
    if (iters > 2 && r == cudaSuccess) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, t0, t1);
        fprintf(stderr, "[bgemm] rows_act=%d rows=%d K=%d n_tile=%d tiles=%d pieces=%d in_f16=%d epi=%d: %.2f us per stage (incl. grid barrier)\n",
                rows_act, rows, K, T.n_tile, T.n_tiles, T.n_pieces, in_f16, epi, ms * 1e3 / iters);
        unsigned long long pr[64];
        cudaMemcpy(pr, e->prof + 2048, sizeof(pr), cudaMemcpyDeviceToHost);
        if (pr[56]) fprintf(stderr, "[bgemm]   issuer 0 per call (cycles): fence + wait weights %llu, MMA issue %llu, commits %llu (calls %llu)\n",
                            pr[57] / pr[56], pr[58] / pr[56], pr[59] / pr[56], pr[56]);
        if (pr[48]) fprintf(stderr, "[bgemm]   F32 stage, thread 0 of tile 0's CTA (cycles per tile): prologue %llu, wait A buffer %llu, convert+store (incl. load latency) %llu, "
                            "fence+sync %llu, issue %llu, epilogue+sync %llu, total %llu\n", pr[49] / pr[48], pr[50] / pr[48], pr[51] / pr[48],
                            pr[52] / pr[48], pr[53] / pr[48], pr[54] / pr[48], pr[55] / pr[48]);
        if (pr[40]) fprintf(stderr, "[bgemm]   F16 stage, issuer 0 of tile 0 (cycles per tile): first A chunk after %llu, waiting for A %llu, issue+commit+weights %llu, "
                            "loop %llu; warp 1: epilogue (incl. wait for the MMAs) %llu, closing sync %llu\n",
                            pr[41] / pr[40], pr[42] / pr[40], pr[43] / pr[40], pr[44] / pr[40], pr[45] / pr[40], pr[46] / pr[40]);
    }
    cudaEventDestroy(t0); cudaEventDestroy(t1);
    if (r != cudaSuccess) { e->err = std::string("bgemm: ") + cudaGetErrorString(r); return cleanup(0); }
    if (epi == EPI_SWIGLU) {
        const size_t cb = (size_t) ((rows_act + 7) / 8) * 2048;
        std::vector<__half> oh((size_t) (outw / 128 + 1) * cb / 2);
        cudaMemcpy(oh.data(), d_out16, std::min(oh.size(), (size_t) (rows_act + 8) * rows) * 2, cudaMemcpyDeviceToHost);
        for (int r = 0; r < rows_act; ++r)
            for (int k = 0; k < outw; ++k)
                out[(size_t) r * outw + k] = __half2float(oh[((size_t) (k >> 7) * cb + (size_t) (r >> 3) * 2048 +
                                                               (size_t) ((k & 127) >> 3) * 128 + (r & 7) * 16 + (k & 7) * 2) / 2]);
    } else {
        cudaMemcpy(out, d_out, (size_t) rows_act * outw * 4, cudaMemcpyDeviceToHost);
    }
    if (ssq_out) cudaMemcpy(ssq_out, d_ssq, (size_t) rows_act * (rows / 16) * 4, cudaMemcpyDeviceToHost);
    return cleanup(1);
}

int qtts_profile(qtts_engine * h, int on, uint64_t * out64) {
    Engine * e = &h->e;
    if (!e->finalized) { e->err = "engine not finalized"; return 0; }
    cudaSetDevice(e->device);
    cudaStreamSynchronize(e->stream);
    if (out64) cudaMemcpy(out64, e->prof, (64 + 4 * 256) * sizeof(uint64_t), cudaMemcpyDeviceToHost);
    if (on != e->prof_on || on) { cudaMemset(e->prof, 0, (64 + 4 * 256) * sizeof(uint64_t)); cudaStreamSynchronize(0); }
    e->prof_on = on > 0 ? on : 0;      // on = 1 + index of the CTA that records
    return 1;
}

int qtts_trace(qtts_engine * h, int on, int cta, uint32_t * out, int32_t max_stages) {
    Engine * e = &h->e;
    if (!e->finalized) { e->err = "engine not finalized"; return 0; }
    cudaSetDevice(e->device);
    cudaStreamSynchronize(e->stream);
    const int cap = 8192;
    if (!e->trace) {
        e->stream_phase = false;
        e->trace = e->dalloc<uint32_t>((size_t) cap * 8);
        e->stream_phase = true;
        if (!e->trace) { e->err = "out of device memory for the trace"; return 0; }
    }
    if (out && max_stages > 0)
        cudaMemcpy(out, e->trace, (size_t) std::min(max_stages, cap) * 8 * sizeof(uint32_t), cudaMemcpyDeviceToHost);
    cudaMemset(e->trace, 0, (size_t) cap * 8 * sizeof(uint32_t));
    cudaStreamSynchronize(0);
    e->trace_on = on ? 1 : 0; e->trace_cta = cta; e->trace_cap = cap;
    return 1;
}

int64_t qtts_launch_count(const qtts_engine * h) { return h->e.launches; }
double qtts_last_loop_ms(const qtts_engine * h) { return h->e.last_loop_ms; }
void * qtts_stream_handle(const qtts_engine * h) { return (void *) h->e.stream; }

}  // extern "C"
