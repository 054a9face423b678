// qtts_api.cu — the extern "C" surface declared in include/qtts_b200.h.
#include "qtts_engine.h"
#include <cstring>
#include <algorithm>

using namespace qtts;

namespace qtts {
extern "C" __global__ void qtts_run_program(const Op * ops, int n_ops, StreamState * st, GridBarrier * bar,
                                            int n_frames, RunCfg cfg);
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
        QTTS_CUDA(cudaMemcpy(s.frame_prog, s.frame_host.data(), s.frame_host.size() * sizeof(Op), cudaMemcpyHostToDevice));
        const std::vector<WSched> S = build_sched(e, s.frame_host);
        QTTS_CUDA(cudaMemcpy(s.frame_sched, S.data(), S.size() * sizeof(WSched), cudaMemcpyHostToDevice));
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
static bool do_forward_prefill(Engine * e, Slot & s, int n_tokens, int n_past) {
    if (n_tokens <= 0 || n_tokens > 64) return e->fail("forward_prefill: n_tokens must be in 1..64");
    if (n_past + n_tokens > e->max_ctx) return e->fail("forward_prefill: KV cache too small");
    std::vector<Op> P;
    const size_t H = e->cfg.hidden_size;
    for (int t = 0; t < n_tokens; ++t)
        emit_talker_step(e, s, P, IN_VEC, s.prefill + t * H, false, n_past + t, ATTN_SOFTMAX_P16, t == n_tokens - 1, false);
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
    if (n_tokens <= 0 || n_tokens > 64) { e->err = "forward_prefill: n_tokens must be in 1..64"; return 0; }
    const size_t H = e->cfg.hidden_size;
    if (cudaMemcpyAsync(s->prefill, embd, n_tokens * H * sizeof(float), cudaMemcpyHostToDevice, e->stream) != cudaSuccess) { e->err = "H2D failed"; return 0; }
    if (!do_forward_prefill(e, *s, n_tokens, n_past)) return 0;
    if (hidden_out) cudaMemcpy(hidden_out, s->hidden, H * sizeof(float), cudaMemcpyDeviceToHost);
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
    if (n_past + 1 > e->max_ctx) { e->err = "forward_step: KV cache too small"; return 0; }
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

static bool stream_begin(Engine * e, Slot & s, const int32_t * tokens, int32_t n, const float * speaker, int32_t max_len,
                         int32_t language_id, float rep_pen, float temperature, int32_t top_k, uint64_t seed) {
    if (!tokens) return e->fail("text_tokens is null");
    if (n < 4) return e->fail("Need at least 4 text tokens for generation");
    if (!do_build_prefill(e, s, tokens, n, speaker, language_id)) return false;
    // required context as in the reference (src/tts_transformer.cpp:2383)
    if (s.prefill_len + max_len + 8 > e->max_ctx) return e->fail("max_len does not fit the engine's max_ctx");
    StreamState st{};
    st.max_len = max_len; st.n_text_tokens = n; st.trailing_len = s.trailing_len;
    st.top_k = top_k; st.rep_pen = rep_pen; st.temperature = temperature; st.seed = seed;
    st.ignore_eos = e->ignore_eos;
    QTTS_CUDA(cudaMemcpyAsync(s.st, &st, sizeof(st), cudaMemcpyHostToDevice, e->stream));
    QTTS_CUDA(cudaStreamSynchronize(e->stream));
    if (!do_forward_prefill(e, s, s.prefill_len, 0)) return false;
    s.active = true;
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

// v1: streams are advanced one after another with the batch-1 frame kernel (weights re-streamed
// per stream); the batched tcgen05 path replaces this loop.
int qtts_step_frames(qtts_engine * h, int32_t n_frames, int32_t * codes_out, int32_t * produced, int32_t * active) {
    Engine * e = &h->e;
    if (!e->finalized) { e->err = "engine not finalized"; return 0; }
    cudaSetDevice(e->device);
    const int nb = e->cfg.n_codebooks;
    std::vector<int> before(e->slots.size(), 0);
    cudaEventRecord(e->ev0, e->stream);
    for (size_t i = 0; i < e->slots.size(); ++i) {
        Slot & s = e->slots[i];
        produced[i] = 0;
        if (!s.active) { active[i] = 0; continue; }
        StreamState hs;
        cudaMemcpyAsync(&hs, s.st, sizeof(hs), cudaMemcpyDeviceToHost, e->stream);
        cudaStreamSynchronize(e->stream);
        before[i] = hs.frame;
        if (!run_device_program(e, s, s.frame_prog, &s.frame_host, s.frame_prog_n, s.frame_sched, s.frame_sched_n, n_frames, true)) return 0;
    }
    cudaEventRecord(e->ev1, e->stream);
    for (size_t i = 0; i < e->slots.size(); ++i) {
        Slot & s = e->slots[i];
        if (!s.active) continue;
        StreamState hs;
        cudaMemcpyAsync(&hs, s.st, sizeof(hs), cudaMemcpyDeviceToHost, e->stream);
        if (cudaStreamSynchronize(e->stream) != cudaSuccess) { e->err = std::string("step_frames: ") + cudaGetErrorString(cudaGetLastError()); return 0; }
        const int newf = hs.frame - before[i];
        produced[i] = newf;
        if (newf > 0)
            cudaMemcpy(codes_out + ((size_t) i * n_frames) * nb, s.codes + (size_t) before[i] * nb,
                       (size_t) newf * nb * sizeof(int32_t), cudaMemcpyDeviceToHost);
        if (hs.done) s.active = false;
        active[i] = s.active ? 1 : 0;
    }
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e->ev0, e->ev1);
    e->last_loop_ms = ms;
    return 1;
}

int qtts_profile(qtts_engine * h, int on, uint64_t * out64) {
    Engine * e = &h->e;
    if (!e->finalized) { e->err = "engine not finalized"; return 0; }
    cudaSetDevice(e->device);
    cudaStreamSynchronize(e->stream);
    if (out64) cudaMemcpy(out64, e->prof, (64 + 4 * 256) * sizeof(uint64_t), cudaMemcpyDeviceToHost);
    if (on != e->prof_on || on) cudaMemset(e->prof, 0, (64 + 4 * 256) * sizeof(uint64_t));
    e->prof_on = on > 0 ? on : 0;      // on = 1 + index of the CTA that records
    return 1;
}

int64_t qtts_launch_count(const qtts_engine * h) { return h->e.launches; }
double qtts_last_loop_ms(const qtts_engine * h) { return h->e.last_loop_ms; }
void * qtts_stream_handle(const qtts_engine * h) { return (void *) h->e.stream; }

}  // extern "C"
