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

// the product's restatement of libstdc++'s introselect (header-only, host + device), checked below
// against std::nth_element itself
#include "../qwen3-tts-jetson_b200/csrc/qtts_nth.cuh"

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

// ---- sampled runs made reproducible -------------------------------------------------------
// The reference seeds its sampler from std::random_device (src/tts_transformer.h:298), so its
// sampled streams differ from run to run.  The harness re-seeds that very member (private,
// reached through the re-declaration above), which changes nothing else of the reference.
void ref_seed_rng(void * h, uint32_t seed) {
    static_cast<RefHandle *>(h)->tts.rng_.seed(seed);
}

// Codebook-0 sampling on already post-processed logits: a line-by-line restatement of
// src/tts_transformer.cpp:2451-2494 with the SAME libstdc++ calls (std::nth_element on an index
// vector with the value comparator, std::max_element, expf, double sum, float CDF scan), so the
// threshold it picks is the one the reference picks for these logits: scored_idx[top_k - 1] after
// nth_element(begin, begin + top_k, end) is SOME member of the top-k set, not necessarily the
// k-th largest.  `u` stands for the reference's uniform_real_distribution<float> draw.
// kept_out (may be null) receives the number of candidates that survived the mask.
int ref_sample_cb0(const float * processed, int vocab, int eos_id, float temperature, int top_k,
                   float u, int * kept_out) {
    std::vector<float> logits(processed, processed + vocab), probs(vocab), scored_vals(vocab);
    std::vector<int32_t> scored_idx(vocab);
    for (int32_t i = 0; i < vocab; ++i) logits[i] /= temperature;
    float eos_logit = logits[eos_id];
    if (top_k > 0 && top_k < vocab) {
        for (int32_t i = 0; i < vocab; ++i) { scored_vals[i] = logits[i]; scored_idx[i] = i; }
        std::nth_element(scored_idx.begin(), scored_idx.begin() + top_k, scored_idx.end(),
            [&](int32_t a, int32_t b) { return scored_vals[a] > scored_vals[b]; });
        float threshold = scored_vals[scored_idx[top_k - 1]];
        for (int32_t i = 0; i < vocab; ++i) if (logits[i] < threshold) logits[i] = -INFINITY;
    }
    logits[eos_id] = eos_logit;
    if (kept_out) { int k = 0; for (int32_t i = 0; i < vocab; ++i) k += std::isfinite(logits[i]) ? 1 : 0; *kept_out = k; }
    float max_logit = *std::max_element(logits.data(), logits.data() + vocab);
    double sum = 0.0;
    for (int32_t i = 0; i < vocab; ++i) { probs[i] = expf(logits[i] - max_logit); sum += probs[i]; }
    for (int32_t i = 0; i < vocab; ++i) probs[i] = (float) (probs[i] / sum);
    float cumsum = 0.0f;
    int32_t next_token = vocab - 1;
    for (int32_t i = 0; i < vocab; ++i) { cumsum += probs[i]; if (cumsum >= u) { next_token = i; break; } }
    return next_token;
}

// std::nth_element as the reference calls it (index vector, value-descending comparator,
// src/tts_transformer.cpp:2458-2465) next to the product's restatement (csrc/qtts_nth.cuh) on the
// same values: both index arrays are returned so the test can require them to be IDENTICAL.
void ref_nth_check(const float * vals, int n, int k, int32_t * idx_std, int32_t * idx_ours) {
    std::vector<int32_t> a(n);
    for (int i = 0; i < n; ++i) { a[i] = i; idx_ours[i] = i; }
    std::nth_element(a.begin(), a.begin() + k, a.end(), [&](int32_t x, int32_t y) { return vals[x] > vals[y]; });
    memcpy(idx_std, a.data(), n * sizeof(int32_t));
    qtts::nth::nth_element_desc(idx_ours, k, n, vals);
}

// generate() replayed frame by frame through the reference's OWN methods (forward_prefill,
// predict_codes_autoregressive, lookup_single_embedding_row, forward_step, rng_) with the inline
// codebook-0 block of src/tts_transformer.cpp:2417-2499 restated between them, so that the
// per-frame intermediates the real generate() never exposes can be dumped.  The caller checks the
// codes against ref_generate() on the same inputs (and seed): equal codes = faithful replay.
//   raw_out       [rows][V]   Talker logits as forward_prefill / forward_step returned them
//   proc_out      [rows][V]   after suppress / repetition penalty / EOS ramp (before temperature)
//   hidden_out    [rows][H]   last_hidden_ handed to the code predictor
//   step_embd_out [rows][H]   next Talker input assembled after the frame (:2529-2553)
//   u_cb0_out     [rows]      the uniform_real_distribution<float> draw of the frame (temperature > 0)
//   u_cp_out      [rows][15]  the generate_canonical<double,53> draws std::discrete_distribution
//                             consumes for the frame's 15 codes (taken from a COPY of rng_ before the call)
//   codes_out     [rows][16]
// rows = frames produced (+1 when the loop ended by EOS: that frame's raw / proc / u_cb0 are
// filled too, *eos_frame_out = its index, else -1).  Any output pointer may be null.  Returns the
// number of frames produced, or -1.
int ref_generate_trace(void * h, const int32_t * tokens, int n_tokens, const float * speaker, int max_len,
                       int language_id, float rep_pen, float temperature, int top_k,
                       float * raw_out, float * proc_out, float * hidden_out, float * step_embd_out,
                       float * u_cb0_out, double * u_cp_out, int32_t * codes_out, int * eos_frame_out) {
    auto * r = static_cast<RefHandle *>(h);
    auto & t = r->tts;
    const auto & cfg = t.get_config();
    const int H = cfg.hidden_size, V = cfg.codec_vocab_size, NB = cfg.n_codebooks;
    if (eos_frame_out) *eos_frame_out = -1;
    if (n_tokens < 4 || max_len <= 0) return -1;
    std::vector<float> prefill, trailing, pad;
    if (!t.build_prefill_graph(tokens, n_tokens, speaker, language_id, prefill, trailing, pad)) return -1;
    const int prefill_len = (int) (prefill.size() / H), trailing_len = (int) (trailing.size() / H);
    const int required_ctx = prefill_len + max_len + 8;
    if (t.state_.cache.n_ctx < required_ctx || t.state_.cache.n_ctx > std::max<int32_t>(required_ctx * 2, 512))
        if (!t.init_kv_cache(required_ctx)) return -1;
    t.clear_kv_cache();
    std::vector<float> hidden_all, logits;
    if (!t.forward_prefill(prefill.data(), prefill_len, 0, hidden_all, &logits)) return -1;
    int n_past = prefill_len, produced = 0;
    std::unordered_set<int32_t> seen;
    std::vector<float> step_embd(H), row(H);
    std::vector<int32_t> codes;
    for (int frame = 0; frame < max_len; ++frame) {
        if (raw_out) memcpy(raw_out + (size_t) frame * V, logits.data(), V * sizeof(float));
        for (int32_t i = V - 1024; i < V; ++i) if (i != cfg.codec_eos_id) logits[i] = -INFINITY;
        if (rep_pen != 1.0f)
            for (int32_t tok : seen) { if (logits[tok] > 0.0f) logits[tok] /= rep_pen; else logits[tok] *= rep_pen; }
        const int32_t expected = std::max(20, n_tokens * 4);
        if (frame >= expected) {
            float ramp = std::min(1.0f, (float) (frame - expected) / (float) expected);
            float max_lg = *std::max_element(logits.data(), logits.data() + V);
            float target = max_lg + 5.0f;
            logits[cfg.codec_eos_id] += ramp * (target - logits[cfg.codec_eos_id]);
        }
        if (proc_out) memcpy(proc_out + (size_t) frame * V, logits.data(), V * sizeof(float));
        int32_t next;
        if (temperature <= 0.0f) {
            next = 0;
            for (int32_t i = 1; i < V; ++i) if (logits[i] > logits[next]) next = i;
        } else {
            std::uniform_real_distribution<float> udist(0.0f, 1.0f);
            const float u = udist(t.rng_);
            if (u_cb0_out) u_cb0_out[frame] = u;
            next = ref_sample_cb0(logits.data(), V, cfg.codec_eos_id, temperature, top_k, u, nullptr);
        }
        if (next == cfg.codec_eos_id) { if (eos_frame_out) *eos_frame_out = frame; break; }
        seen.insert(next);
        if (hidden_out) memcpy(hidden_out + (size_t) frame * H, t.last_hidden_.data(), H * sizeof(float));
        if (u_cp_out && temperature > 0.0f) {
            std::mt19937 peek = t.rng_;
            for (int s = 0; s < 15; ++s) u_cp_out[(size_t) frame * 15 + s] = std::generate_canonical<double, 53>(peek);
        }
        if (!t.predict_codes_autoregressive(t.last_hidden_.data(), next, codes, temperature, top_k)) return -1;
        if (codes_out) {
            codes_out[(size_t) frame * NB] = next;
            for (int cb = 1; cb < NB; ++cb) codes_out[(size_t) frame * NB + cb] = codes[cb - 1];
        }
        produced = frame + 1;
        if (frame + 1 >= max_len) break;
        if (!t.lookup_single_embedding_row(t.model_.codec_embd, next, row.data())) return -1;
        for (int i = 0; i < H; ++i) step_embd[i] = row[i];
        for (int cb = 1; cb < NB; ++cb) {
            if (!t.lookup_single_embedding_row(t.model_.code_pred_embd[cb - 1], codes[cb - 1], row.data())) return -1;
            for (int i = 0; i < H; ++i) step_embd[i] += row[i];
        }
        const float * tr = (frame < trailing_len) ? trailing.data() + (size_t) frame * H : pad.data();
        for (int i = 0; i < H; ++i) step_embd[i] += tr[i];
        if (step_embd_out) memcpy(step_embd_out + (size_t) frame * H, step_embd.data(), H * sizeof(float));
        if (!t.forward_step(step_embd.data(), n_past, logits)) return -1;
        n_past++;
    }
    return produced;
}

}  // extern "C"
