// TEST INFRASTRUCTURE — C-ABI harness around the UNMODIFIED reference class
// qwen3_tts::TTSTransformer (reference: src/tts_transformer.h:164-245) running on
// the ggml CPU backend.  Built by oracle/Makefile.ref into oracle/_ref/libqtts_ref.so
// together with the reference's own translation units (compiled where they lie under
// /root/reference; no reference source is copied into this repo).
//
// Used only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs, through ctypes.  The product never loads it.
//
// The harness code below is ours: it calls the reference's public methods, plus a
// few private ones (graph builders, embedding assembly) that are reached by
// re-declaring `private` for this one translation unit so that per-step code-predictor
// logits and the host-side prefill embeddings can be dumped for parity fixtures.

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <map>
#include <random>
#include <string>
#include <unordered_set>
#include <vector>

#include "ggml.h"
#include "ggml-backend.h"
#include "ggml-cpu.h"
#include "gguf.h"

#define private public
#include "tts_transformer.h"
#undef private
#include "text_tokenizer.h"

using qwen3_tts::TTSTransformer;

namespace {
struct RefHandle {
    TTSTransformer tts;
    qwen3_tts::TextTokenizer tok;
    bool tok_loaded = false;
    std::string err;
};
double now_ms() {
    using namespace std::chrono;
    return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}
}  // namespace

extern "C" {

void * ref_create() { return new RefHandle(); }

void ref_destroy(void * h) { delete static_cast<RefHandle *>(h); }

const char * ref_error(void * h) {
    auto * r = static_cast<RefHandle *>(h);
    if (!r->err.empty()) return r->err.c_str();
    return r->tts.get_error().c_str();
}

// load_model (reference: src/tts_transformer.cpp:51)
int ref_load(void * h, const char * gguf_path) {
    auto * r = static_cast<RefHandle *>(h);
    if (!r->tts.load_model(gguf_path)) return 0;
    r->tok_loaded = false;
    return 1;
}

// The reference ignores -j (src/main.cpp:212-214); the ggml CPU backend default is 4
// threads (ggml/include/ggml.h:232).  n <= 0 keeps that default.
void ref_set_threads(void * h, int n) {
    auto * r = static_cast<RefHandle *>(h);
    if (n > 0 && r->tts.state_.backend) {
        ggml_backend_cpu_set_n_threads(r->tts.state_.backend, n);
    }
}

// cfg_out[0..15]: hidden, n_layers, n_heads, n_kv_heads, ffn, head_dim, codec_vocab, n_codebooks,
// cp_layers, cp_vocab, cp_hidden, cp_ffn, cp_heads, cp_kv_heads, cp_head_dim, text_embd_dim
void ref_get_config(void * h, int32_t * cfg_out, float * eps, float * theta) {
    auto * r = static_cast<RefHandle *>(h);
    const auto & c = r->tts.get_config();
    int32_t v[16] = {c.hidden_size, c.n_layers, c.n_attention_heads, c.n_key_value_heads,
                     c.intermediate_size, c.head_dim, c.codec_vocab_size, c.n_codebooks,
                     c.code_pred_layers, c.code_pred_vocab_size, c.code_pred_hidden_size,
                     c.code_pred_intermediate_size, c.code_pred_n_attention_heads,
                     c.code_pred_n_kv_heads, c.code_pred_head_dim, c.text_embd_dim};
    memcpy(cfg_out, v, sizeof(v));
    *eps = c.rms_norm_eps;
    *theta = c.rope_theta;
}

// TextTokenizer::encode_for_tts (reference: src/text_tokenizer.cpp:293-330)
int ref_tokenize(void * h, const char * gguf_path, const char * text, int32_t * out, int max_out) {
    auto * r = static_cast<RefHandle *>(h);
    if (!r->tok_loaded) {
        struct ggml_context * meta = nullptr;
        struct gguf_init_params p = {true, &meta};
        struct gguf_context * g = gguf_init_from_file(gguf_path, p);
        if (!g) { r->err = "tokenizer: cannot open gguf"; return -1; }
        bool ok = r->tok.load_from_gguf(g);
        gguf_free(g);
        if (meta) ggml_free(meta);
        if (!ok) { r->err = "tokenizer: " + r->tok.get_error(); return -1; }
        r->tok_loaded = true;
    }
    std::vector<int32_t> ids = r->tok.encode_for_tts(text);
    if ((int) ids.size() > max_out) return -2;
    memcpy(out, ids.data(), ids.size() * sizeof(int32_t));
    return (int) ids.size();
}

// build_prefill_graph (reference: src/tts_transformer.cpp:1093-1231): host-side embedding assembly.
// Returns prefill_len; trailing_len via *trailing_len_out.  Buffers must be large enough:
// prefill_out[(3+7+1)*H], trailing_out[(n_tokens)*H], pad_out[H].
int ref_build_prefill(void * h, const int32_t * tokens, int n_tokens, const float * speaker,
                      int language_id, float * prefill_out, float * trailing_out, float * pad_out,
                      int * trailing_len_out) {
    auto * r = static_cast<RefHandle *>(h);
    std::vector<float> pre, trail, pad;
    if (!r->tts.build_prefill_graph(tokens, n_tokens, speaker, language_id, pre, trail, pad)) return -1;
    const int H = r->tts.get_config().hidden_size;
    memcpy(prefill_out, pre.data(), pre.size() * sizeof(float));
    memcpy(trailing_out, trail.data(), trail.size() * sizeof(float));
    memcpy(pad_out, pad.data(), pad.size() * sizeof(float));
    *trailing_len_out = (int) (trail.size() / H);
    return (int) (pre.size() / H);
}

int ref_init_kv(void * h, int n_ctx) {
    auto * r = static_cast<RefHandle *>(h);
    if (!r->tts.init_kv_cache(n_ctx)) return 0;
    r->tts.clear_kv_cache();
    return 1;
}

// forward_prefill (reference: src/tts_transformer.cpp:1829-1920)
int ref_forward_prefill(void * h, const float * embd, int n_tokens, int n_past,
                        float * hidden_last_out, float * logits_out) {
    auto * r = static_cast<RefHandle *>(h);
    std::vector<float> hid, logits;
    if (!r->tts.forward_prefill(embd, n_tokens, n_past, hid, &logits)) return 0;
    const int H = r->tts.get_config().hidden_size;
    if (hidden_last_out) memcpy(hidden_last_out, hid.data() + (size_t) (n_tokens - 1) * H, H * sizeof(float));
    if (logits_out) memcpy(logits_out, logits.data(), logits.size() * sizeof(float));
    return 1;
}

// forward_step (reference: src/tts_transformer.cpp:1952-2028)
int ref_forward_step(void * h, const float * step_embd, int n_past, float * logits_out,
                     float * hidden_out) {
    auto * r = static_cast<RefHandle *>(h);
    std::vector<float> logits, hid;
    if (!r->tts.forward_step(step_embd, n_past, logits, &hid)) return 0;
    if (logits_out) memcpy(logits_out, logits.data(), logits.size() * sizeof(float));
    if (hidden_out) memcpy(hidden_out, hid.data(), hid.size() * sizeof(float));
    return 1;
}

// predict_codes_autoregressive through the PUBLIC method (reference: :2153-2340)
int ref_predict_codes(void * h, const float * hidden, int cb0, float temperature, int top_k,
                      int32_t * out15) {
    auto * r = static_cast<RefHandle *>(h);
    std::vector<int32_t> out;
    if (!r->tts.predict_codes_autoregressive(hidden, cb0, out, temperature, top_k)) return 0;
    memcpy(out15, out.data(), 15 * sizeof(int32_t));
    return 1;
}

// Same sequence of reference graphs as predict_codes_autoregressive's GGML branch
// (reference: src/tts_transformer.cpp:2185-2337), greedy, but also returning the 15 logit
// vectors.  If forced15 != nullptr the fed-back code of step s is forced15[s] (teacher forcing)
// while out15 still holds the argmax of each step.
int ref_predict_codes_logits(void * h, const float * hidden, int cb0, const int32_t * forced15,
                             int32_t * out15, float * logits_out /* [15 * cp_vocab] */) {
    auto * r = static_cast<RefHandle *>(h);
    auto & t = r->tts;
    const auto & cfg = t.get_config();
    const int V = cfg.code_pred_vocab_size;
    if (t.state_.code_pred_cache.n_ctx < 16) {
        if (!t.init_code_pred_kv_cache(16)) return 0;
    }
    t.clear_code_pred_kv_cache();
    std::vector<float> cb0_embd(cfg.hidden_size);
    if (!t.lookup_single_embedding_row(t.model_.codec_embd, cb0, cb0_embd.data())) return 0;
    auto amax = [&](const float * d) {
        int best = 0;
        for (int i = 1; i < V; ++i) if (d[i] > d[best]) best = i;
        return best;
    };
    {
        struct ggml_cgraph * gf = t.build_code_pred_prefill_graph();
        if (!ggml_backend_sched_alloc_graph(t.state_.sched, gf)) return 0;
        ggml_backend_tensor_set(ggml_graph_get_tensor(gf, "inp_hidden"), hidden, 0, cfg.hidden_size * sizeof(float));
        ggml_backend_tensor_set(ggml_graph_get_tensor(gf, "inp_cb0_embd"), cb0_embd.data(), 0, cfg.hidden_size * sizeof(float));
        int32_t positions[2] = {0, 1};
        ggml_backend_tensor_set(ggml_graph_get_tensor(gf, "inp_pos"), positions, 0, sizeof(positions));
        if (ggml_backend_sched_graph_compute(t.state_.sched, gf) != GGML_STATUS_SUCCESS) return 0;
        ggml_backend_tensor_get(ggml_graph_get_tensor(gf, "logits"), logits_out, 0, V * sizeof(float));
        out15[0] = amax(logits_out);
        ggml_backend_sched_reset(t.state_.sched);
    }
    for (int step = 1; step < 15; ++step) {
        const int32_t n_past = step + 1;
        struct ggml_cgraph * gf = t.build_code_pred_step_graph(n_past, step);
        if (!ggml_backend_sched_alloc_graph(t.state_.sched, gf)) return 0;
        // inp_hidden only feeds the graph when generation_step == 0 (reference :1700-1707)
        if (struct ggml_tensor * ih = ggml_graph_get_tensor(gf, "inp_hidden"))
            ggml_backend_tensor_set(ih, hidden, 0, cfg.hidden_size * sizeof(float));
        int32_t prev = forced15 ? forced15[step - 1] : out15[step - 1];
        ggml_backend_tensor_set(ggml_graph_get_tensor(gf, "inp_code"), &prev, 0, sizeof(int32_t));
        int32_t pos = n_past;
        ggml_backend_tensor_set(ggml_graph_get_tensor(gf, "inp_pos"), &pos, 0, sizeof(int32_t));
        if (ggml_backend_sched_graph_compute(t.state_.sched, gf) != GGML_STATUS_SUCCESS) return 0;
        float * dst = logits_out + (size_t) step * V;
        ggml_backend_tensor_get(ggml_graph_get_tensor(gf, "logits"), dst, 0, V * sizeof(float));
        out15[step] = amax(dst);
        ggml_backend_sched_reset(t.state_.sched);
    }
    return 1;
}

// generate (reference: src/tts_transformer.cpp:2342-2574).  Returns frames produced (or -1).
// ms_out[0] = wall ms of the whole call (prefill included).
int ref_generate(void * h, const int32_t * tokens, int n_tokens, const float * speaker, int max_len,
                 int language_id, float rep_pen, float temperature, int top_k,
                 int32_t * codes_out /* [max_len*16] */, double * ms_out) {
    auto * r = static_cast<RefHandle *>(h);
    std::vector<int32_t> out;
    const double t0 = now_ms();
    if (!r->tts.generate(tokens, n_tokens, speaker, max_len, out, language_id, rep_pen,
                         temperature, top_k, nullptr, 40)) return -1;
    const double t1 = now_ms();
    if (ms_out) ms_out[0] = t1 - t0;
    memcpy(codes_out, out.data(), out.size() * sizeof(int32_t));
    return (int) (out.size() / r->tts.get_config().n_codebooks);
}

}  // extern "C"
