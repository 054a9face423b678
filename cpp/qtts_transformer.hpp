// qtts_transformer.hpp — C++ host adapter over the C ABI of libqtts_b200.so (include/qtts_b200.h).
//
// Same class surface, argument meaning and error behaviour (bool + get_error()) as the reference's
// qwen3_tts::TTSTransformer (src/tts_transformer.h:164-245) and the loop entry points of
// qwen3_tts::TRTCodePredictor (src/trt_code_predictor.h:25-87), so that the reference's own callers
// (Qwen3TTS::synthesize_internal, src/qwen3_tts.cpp:459-463) compile against it unchanged.  The
// class only marshals pointers; every computation happens in the CUDA library, and every method
// fails (false + message) when that library reports an error — there is no CPU fallback.
#pragma once
#include <algorithm>
#include <cstdint>
#include <functional>
#include <string>
#include <vector>

#include "../include/qtts_b200.h"
#include "qtts_gguf.hpp"

namespace qtts_b200 {

// reference: tts_transformer_config (src/tts_transformer.h:21-69); the C ABI struct carries the same fields
using tts_transformer_config = qtts_config;

namespace detail {
struct KeyU32 { int32_t qtts_config::* field; std::vector<const char *> keys; int64_t def; int32_t qtts_config::* fallback; };

// parse_config (src/tts_transformer.cpp:288-442): same alternative keys, same defaults
inline void parse_config(const GgufFile & g, qtts_config & c) {
    const std::vector<KeyU32> tab = {
        {&qtts_config::text_vocab_size, {"qwen3-tts.text.vocab_size", "qwen3-tts.text_vocab_size"}, 151936, nullptr},
        {&qtts_config::text_embd_dim, {"qwen3-tts.text.embedding_dim", "qwen3-tts.text_hidden_size"}, 2048, nullptr},
        {&qtts_config::hidden_size, {"qwen3-tts.talker.embedding_length", "qwen3-tts.embedding_length"}, 1024, nullptr},
        {&qtts_config::n_layers, {"qwen3-tts.talker.block_count", "qwen3-tts.block_count"}, 28, nullptr},
        {&qtts_config::n_heads, {"qwen3-tts.talker.attention.head_count", "qwen3-tts.attention.head_count"}, 16, nullptr},
        {&qtts_config::n_kv_heads, {"qwen3-tts.talker.attention.head_count_kv", "qwen3-tts.attention.head_count_kv"}, 8, nullptr},
        {&qtts_config::intermediate_size, {"qwen3-tts.talker.feed_forward_length", "qwen3-tts.feed_forward_length"}, 3072, nullptr},
        {&qtts_config::head_dim, {"qwen3-tts.talker.attention.key_length", "qwen3-tts.attention.key_length"}, 128, nullptr},
        {&qtts_config::codec_vocab_size, {"qwen3-tts.talker.codec_vocab_size", "qwen3-tts.vocab_size"}, 3072, nullptr},
        {&qtts_config::n_codebooks, {"qwen3-tts.talker.num_codebooks", "qwen3-tts.num_code_groups"}, 16, nullptr},
        {&qtts_config::cp_layers, {"qwen3-tts.code_pred.layer_count", "qwen3-tts.code_predictor.layer_count"}, 5, nullptr},
        {&qtts_config::cp_vocab_size, {"qwen3-tts.code_pred.vocab_size", "qwen3-tts.code_predictor.vocab_size"}, 2048, nullptr},
        {&qtts_config::cp_hidden_size, {"qwen3-tts.code_predictor.embedding_length"}, -1, &qtts_config::hidden_size},
        {&qtts_config::cp_intermediate_size, {"qwen3-tts.code_predictor.feed_forward_length"}, -1, &qtts_config::intermediate_size},
        {&qtts_config::cp_n_heads, {"qwen3-tts.code_predictor.attention.head_count"}, -1, &qtts_config::n_heads},
        {&qtts_config::cp_n_kv_heads, {"qwen3-tts.code_predictor.attention.head_count_kv"}, -1, &qtts_config::n_kv_heads},
        {&qtts_config::cp_head_dim, {"qwen3-tts.code_predictor.attention.key_length"}, -1, &qtts_config::head_dim},
        {&qtts_config::codec_pad_id, {"qwen3-tts.codec.pad_id"}, 2148, nullptr},
        {&qtts_config::codec_bos_id, {"qwen3-tts.codec.bos_id"}, 2149, nullptr},
        {&qtts_config::codec_eos_id, {"qwen3-tts.codec.eos_id", "qwen3-tts.codec.eos_token_id"}, 2150, nullptr},
        {&qtts_config::tts_bos_id, {"qwen3-tts.tts_bos_token_id", "qwen3-tts.tts.bos_token_id", "qwen3-tts.tts.bos_id"}, 151672, nullptr},
        {&qtts_config::tts_eos_id, {"qwen3-tts.tts_eos_token_id", "qwen3-tts.tts.eos_token_id", "qwen3-tts.tts.eos_id"}, 151673, nullptr},
        {&qtts_config::tts_pad_id, {"qwen3-tts.tts_pad_token_id", "qwen3-tts.tts.pad_token_id", "qwen3-tts.tts.pad_id"}, 151671, nullptr},
        {&qtts_config::codec_think_id, {"qwen3-tts.codec.think_id", "qwen3-tts.codec_think_id"}, 2154, nullptr},
        {&qtts_config::codec_nothink_id, {"qwen3-tts.codec.nothink_id", "qwen3-tts.codec_nothink_id"}, 2155, nullptr},
        {&qtts_config::codec_think_bos_id, {"qwen3-tts.codec.think_bos_id", "qwen3-tts.codec_think_bos_id"}, 2156, nullptr},
        {&qtts_config::codec_think_eos_id, {"qwen3-tts.codec.think_eos_id", "qwen3-tts.codec_think_eos_id"}, 2157, nullptr},
    };
    for (const auto & e : tab) {
        int64_t v = e.fallback ? (int64_t) (c.*(e.fallback)) : e.def;
        for (const char * k : e.keys) if (g.has(k)) { v = (int64_t) g.num(k); break; }
        c.*(e.field) = (int32_t) v;
    }
    c.rms_norm_eps = 1e-6f;
    for (const char * k : {"qwen3-tts.talker.attention.layer_norm_rms_epsilon", "qwen3-tts.attention.layer_norm_rms_epsilon"})
        if (g.has(k)) { c.rms_norm_eps = (float) g.num(k); break; }
    c.rope_theta = 1000000.0f;
    for (const char * k : {"qwen3-tts.talker.rope.freq_base", "qwen3-tts.rope.freq_base"})
        if (g.has(k)) { c.rope_theta = (float) g.num(k); break; }
}
}  // namespace detail

// Drop-in for qwen3_tts::TTSTransformer (src/tts_transformer.h:164-245) on the hot path.
class TTSTransformer {
public:
    // frame_callback_t, src/tts_transformer.h:222-223
    using frame_callback_t = std::function<bool(const int32_t *, int32_t, int32_t)>;

    TTSTransformer() = default;
    ~TTSTransformer() { unload_model(); }
    TTSTransformer(const TTSTransformer &) = delete;
    TTSTransformer & operator=(const TTSTransformer &) = delete;

    // engine-creation knobs without a counterpart in the reference (it owns one stream on one device)
    void set_device(int device) { device_ = device; }
    void set_max_streams(int n) { max_streams_ = n; }
    void set_seed(uint64_t seed) { seed_ = seed; }

    // load_model (src/tts_transformer.cpp:155-286): parse config, create tensors, copy the data
    bool load_model(const std::string & model_path) {
        unload_model();
        GgufFile g;
        if (!g.open(model_path, error_msg_)) return false;
        cfg_ = qtts_config{};
        detail::parse_config(g, cfg_);
        model_path_ = model_path;
        // the engine allocates its KV cache at creation; init_kv_cache may re-create it with another n_ctx
        return create_engine(n_ctx_);
    }

    // unload_model (src/tts_transformer.cpp:2576-2640)
    void unload_model() {
        if (eng_) qtts_destroy(eng_);
        eng_ = nullptr;
    }

    // init_kv_cache (src/tts_transformer.cpp:794-870): the reference frees and re-allocates
    bool init_kv_cache(int32_t n_ctx) {
        if (!eng_) { error_msg_ = "Model not loaded"; return false; }
        if (n_ctx <= 0) { error_msg_ = "init_kv_cache: n_ctx must be positive"; return false; }
        if (n_ctx != n_ctx_) {
            unload_model();
            if (!create_engine(n_ctx)) return false;
        }
        return check(qtts_clear_kv_cache(eng_, 0));
    }
    void clear_kv_cache() { if (eng_) qtts_clear_kv_cache(eng_, 0); }
    // the code predictor cache is part of the engine's per-stream state (5 layers x 16 positions)
    bool init_code_pred_kv_cache(int32_t n_ctx) {
        if (!eng_) { error_msg_ = "Model not loaded"; return false; }
        if (n_ctx > 16) { error_msg_ = "code predictor context is at most 16"; return false; }
        return true;
    }
    void clear_code_pred_kv_cache() {}

    // forward_prefill (src/tts_transformer.cpp:1829-1920): output = normalised hidden of the last token
    bool forward_prefill(const float * prefill_embd, int32_t n_tokens, int32_t n_past,
                         std::vector<float> & output, std::vector<float> * logits_out = nullptr) {
        if (!eng_) { error_msg_ = "Model not loaded"; return false; }
        output.resize(cfg_.hidden_size);
        std::vector<float> logits(cfg_.codec_vocab_size);
        if (!check(qtts_forward_prefill(eng_, 0, prefill_embd, n_tokens, n_past, output.data(), logits.data()))) return false;
        last_hidden_ = output;
        if (logits_out) *logits_out = std::move(logits);
        return true;
    }

    // forward_step (src/tts_transformer.cpp:1952-2028): output = codec logits
    bool forward_step(const float * step_embd, int32_t n_past, std::vector<float> & output,
                      std::vector<float> * hidden_out = nullptr) {
        if (!eng_) { error_msg_ = "Model not loaded"; return false; }
        output.resize(cfg_.codec_vocab_size);
        last_hidden_.resize(cfg_.hidden_size);
        if (!check(qtts_forward_step(eng_, 0, step_embd, n_past, output.data(), last_hidden_.data()))) return false;
        if (hidden_out) *hidden_out = last_hidden_;
        return true;
    }

    // get_hidden_states (src/tts_transformer.cpp:2030-2040)
    bool get_hidden_states(std::vector<float> & hidden) const {
        if (last_hidden_.empty()) return false;
        hidden = last_hidden_;
        return true;
    }

    // build_prefill_graph + project_text_tokens (src/tts_transformer.cpp:1026-1231): private in the
    // reference (generate() calls it); public here so a caller can drive prefill/step by hand
    bool build_prefill(const int32_t * text_tokens, int32_t n_tokens, const float * speaker_embd,
                       int32_t language_id, std::vector<float> & prefill_embd,
                       std::vector<float> & trailing_text_hidden, std::vector<float> & tts_pad_embed) {
        if (!eng_) { error_msg_ = "Model not loaded"; return false; }
        int32_t pl = 0, tl = 0;
        const int H = cfg_.hidden_size;
        // first call sizes the outputs, second fills them
        if (!check(qtts_build_prefill(eng_, 0, text_tokens, n_tokens, speaker_embd, language_id, &pl, &tl,
                                      nullptr, nullptr, nullptr))) return false;
        prefill_embd.resize((size_t) pl * H);
        trailing_text_hidden.resize((size_t) tl * H);
        tts_pad_embed.resize(H);
        return check(qtts_build_prefill(eng_, 0, text_tokens, n_tokens, speaker_embd, language_id, &pl, &tl,
                                        prefill_embd.data(), trailing_text_hidden.data(), tts_pad_embed.data()));
    }

    // predict_codes_autoregressive (src/tts_transformer.cpp:2153-2340)
    bool predict_codes_autoregressive(const float * hidden, int32_t codebook_0_token,
                                      std::vector<int32_t> & output, float temperature = 0.9f,
                                      int32_t top_k = 50) {
        if (!eng_) { error_msg_ = "Model not loaded"; return false; }
        output.resize(cfg_.n_codebooks - 1);
        return check(qtts_predict_codes(eng_, 0, hidden, codebook_0_token, temperature, top_k,
                                        nullptr, nullptr, output.data(), nullptr));
    }

    // generate (src/tts_transformer.cpp:2342-2574)
    bool generate(const int32_t * text_tokens, int32_t n_tokens, const float * speaker_embd,
                  int32_t max_len, std::vector<int32_t> & output, int32_t language_id = 2050,
                  float repetition_penalty = 1.05f, float temperature = 0.9f, int32_t top_k = 50,
                  frame_callback_t on_frames = nullptr, int32_t callback_interval = 40) {
        if (!eng_) { error_msg_ = "Model not loaded"; return false; }
        if (max_len + 64 + n_tokens > n_ctx_) {
            // the reference sizes the cache from the request (:2378-2386); do the same
            unload_model();
            if (!create_engine(max_len + n_tokens + 64)) return false;
        }
        output.assign((size_t) max_len * cfg_.n_codebooks, 0);
        int32_t n_frames = 0;
        const int rc = qtts_generate(eng_, 0, text_tokens, n_tokens, speaker_embd, max_len, language_id,
                                     repetition_penalty, temperature, top_k, seed_, output.data(), &n_frames,
                                     on_frames ? &TTSTransformer::trampoline : nullptr, callback_interval,
                                     on_frames ? (void *) &on_frames : nullptr);
        if (!check(rc)) { output.clear(); return false; }
        output.resize((size_t) n_frames * cfg_.n_codebooks);
        return true;
    }

    const tts_transformer_config & get_config() const { return cfg_; }
    const std::string & get_error() const { return error_msg_; }
    qtts_engine * engine() const { return eng_; }

private:
    static int trampoline(const int32_t * codes, int32_t n_frames, int32_t n_codebooks, void * user) {
        return (*static_cast<frame_callback_t *>(user))(codes, n_frames, n_codebooks) ? 1 : 0;
    }
    bool check(int rc) {
        if (rc != 0) return true;   // C ABI: 1 = success, 0 = failure
        error_msg_ = eng_ ? qtts_last_error(eng_) : qtts_create_error();
        return false;
    }
    bool create_engine(int32_t n_ctx) {
        GgufFile g;
        if (!g.open(model_path_, error_msg_)) return false;
        if (qtts_create(&cfg_, device_, max_streams_, n_ctx, &eng_) == 0) {
            error_msg_ = qtts_create_error();
            eng_ = nullptr;
            return false;
        }
        n_ctx_ = n_ctx;
        std::vector<uint8_t> buf;
        for (const auto & t : g.tensors()) {
            // vocoder / speaker-encoder tensors live in other GGUF files; unknown names are skipped by the engine
            if (!g.read_tensor(t, buf, error_msg_)) { unload_model(); return false; }
            const int dtype = t.ggml_type == 0 ? QTTS_F32 : QTTS_F16;
            if (qtts_upload_tensor(eng_, t.name.c_str(), dtype, t.dims.data(), (int) t.dims.size(), buf.data()) == 0) {
                error_msg_ = qtts_last_error(eng_);
                unload_model();
                return false;
            }
        }
        if (qtts_finalize(eng_) == 0) {
            error_msg_ = qtts_last_error(eng_);
            unload_model();
            return false;
        }
        return true;
    }

    qtts_engine * eng_ = nullptr;
    qtts_config cfg_{};
    std::string model_path_, error_msg_;
    std::vector<float> last_hidden_;
    int device_ = 0, max_streams_ = 1;
    int32_t n_ctx_ = 4096;
    uint64_t seed_ = 0;
};

// Loop entry points of qwen3_tts::TRTCodePredictor (src/trt_code_predictor.h:61-79) over the same
// engine: the reference keeps a second, TensorRT-built code predictor next to the transformer
// (tts_transformer.h:304-311); here both views share one set of weights.
class CodePredictor {
public:
    explicit CodePredictor(TTSTransformer & owner) : tr_(owner) {}

    bool is_ready() const { return tr_.engine() != nullptr; }

    // run_greedy_loop (src/trt_code_predictor.cpp:484-600).  needs_proj: the 1.7B Talker->code
    // predictor projection (mtp_proj) is part of the engine's program whenever the model has one,
    // so the flag is only checked for consistency with the reference call.
    bool run_greedy_loop(const float * hidden, int32_t hidden_size, int32_t cb0_token, bool needs_proj,
                         int32_t * output_tokens, int32_t n_outputs) {
        return run(hidden, hidden_size, cb0_token, needs_proj, 0.0f, 0, nullptr, output_tokens, n_outputs);
    }

    // run_sampling_loop (src/trt_code_predictor.cpp:602-726): rand_vals = one uniform in [0,1) per code
    bool run_sampling_loop(const float * hidden, int32_t hidden_size, int32_t cb0_token, bool needs_proj,
                           float temperature, int32_t top_k, const float * rand_vals,
                           int32_t * output_tokens, int32_t n_outputs) {
        return run(hidden, hidden_size, cb0_token, needs_proj, temperature, top_k, rand_vals, output_tokens, n_outputs);
    }

    const std::string & get_error() const { return err_.empty() ? tr_.get_error() : err_; }

private:
    bool run(const float * hidden, int32_t hidden_size, int32_t cb0, bool needs_proj, float temperature, int32_t top_k,
             const float * rand_vals, int32_t * out, int32_t n_outputs) {
        err_.clear();
        if (!is_ready()) { err_ = "Model not loaded"; return false; }
        const auto & c = tr_.get_config();
        if (hidden_size != c.hidden_size) { err_ = "hidden_size does not match the model"; return false; }
        if (needs_proj != (c.cp_hidden_size != c.hidden_size)) { err_ = "needs_proj does not match the model"; return false; }
        if (n_outputs != c.n_codebooks - 1) { err_ = "n_outputs must be n_codebooks - 1"; return false; }
        if (qtts_predict_codes(tr_.engine(), 0, hidden, cb0, temperature, top_k, rand_vals, nullptr, out, nullptr) == 0) {
            err_ = qtts_last_error(tr_.engine());
            return false;
        }
        return true;
    }
    TTSTransformer & tr_;
    std::string err_;
};

// Concurrent utterances on one GPU (the reference serves one utterance at a time):
// begin() fills a slot, step() advances every live slot in lock step with the batched kernel.
class StreamServer {
public:
    explicit StreamServer(TTSTransformer & owner, int max_streams) : tr_(owner), n_(max_streams) {}

    bool begin(int slot, const int32_t * text_tokens, int32_t n_tokens, const float * speaker_embd, int32_t max_len,
               int32_t language_id = 2050, float repetition_penalty = 1.05f, float temperature = 0.9f,
               int32_t top_k = 50, uint64_t seed = 0) {
        if (!tr_.engine()) { err_ = "Model not loaded"; return false; }
        if (qtts_stream_begin(tr_.engine(), slot, text_tokens, n_tokens, speaker_embd, max_len, language_id,
                              repetition_penalty, temperature, top_k, seed) == 0) {
            err_ = qtts_last_error(tr_.engine());
            return false;
        }
        return true;
    }

    // begin() in two phases: prepare() many slots, then prefill() once - the Talker prefill of all of them is one
    // batched pass (step() also runs it for slots still pending)
    bool prepare(int slot, const int32_t * text_tokens, int32_t n_tokens, const float * speaker_embd, int32_t max_len,
                 int32_t language_id = 2050, float repetition_penalty = 1.05f, float temperature = 0.9f,
                 int32_t top_k = 50, uint64_t seed = 0) {
        if (!tr_.engine()) { err_ = "Model not loaded"; return false; }
        if (qtts_stream_prepare(tr_.engine(), slot, text_tokens, n_tokens, speaker_embd, max_len, language_id,
                                repetition_penalty, temperature, top_k, seed) == 0) {
            err_ = qtts_last_error(tr_.engine());
            return false;
        }
        return true;
    }
    bool prefill() {
        if (!tr_.engine()) { err_ = "Model not loaded"; return false; }
        if (qtts_prefill_pending(tr_.engine()) == 0) { err_ = qtts_last_error(tr_.engine()); return false; }
        return true;
    }

    // codes: [max_streams][n_frames][n_codebooks]; produced / active: [max_streams]
    bool step(int32_t n_frames, std::vector<int32_t> & codes, std::vector<int32_t> & produced,
              std::vector<int32_t> & active) {
        if (!tr_.engine()) { err_ = "Model not loaded"; return false; }
        codes.assign((size_t) n_ * n_frames * tr_.get_config().n_codebooks, 0);
        produced.assign(n_, 0);
        active.assign(n_, 0);
        if (qtts_step_frames(tr_.engine(), n_frames, codes.data(), produced.data(), active.data()) == 0) {
            err_ = qtts_last_error(tr_.engine());
            return false;
        }
        return true;
    }
    const std::string & get_error() const { return err_; }

private:
    TTSTransformer & tr_;
    int n_;
    std::string err_;
};

}  // namespace qtts_b200
