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
#include <memory>
#include <random>
#include <string>
#include <vector>

#include "../include/qtts_b200.h"
#include "qtts_gguf.hpp"

namespace qtts_b200 {

// tts_transformer_config (src/tts_transformer.h:21-69): same field names and defaults, so that callers reading
// get_config().n_attention_heads / code_pred_hidden_size / tts_bos_token_id ... compile unchanged.  The C ABI's
// qtts_config carries the same values under shorter names (to_abi / from_abi below).
struct tts_transformer_config {
    int32_t text_vocab_size = 151936;
    int32_t text_embd_dim = 2048;
    int32_t hidden_size = 1024;
    int32_t n_layers = 28;
    int32_t n_attention_heads = 16;
    int32_t n_key_value_heads = 8;
    int32_t intermediate_size = 3072;
    int32_t head_dim = 128;
    float rms_norm_eps = 1e-6f;
    float rope_theta = 1000000.0f;
    int32_t mrope_section[3] = {24, 20, 20};      // parsed by the reference, never read (src/tts_transformer.h:37)
    int32_t codec_vocab_size = 3072;
    int32_t n_codebooks = 16;
    int32_t code_pred_layers = 5;
    int32_t code_pred_vocab_size = 2048;
    int32_t code_pred_hidden_size = 0;
    int32_t code_pred_intermediate_size = 0;
    int32_t code_pred_n_attention_heads = 0;
    int32_t code_pred_n_kv_heads = 0;
    int32_t code_pred_head_dim = 0;
    int32_t codec_pad_id = 2148;
    int32_t codec_bos_id = 2149;
    int32_t codec_eos_id = 2150;
    int32_t tts_bos_token_id = 151672;
    int32_t tts_eos_token_id = 151673;
    int32_t tts_pad_token_id = 151671;
    int32_t codec_think_id = 2154;
    int32_t codec_nothink_id = 2155;
    int32_t codec_think_bos_id = 2156;
    int32_t codec_think_eos_id = 2157;
    int32_t english_language_id = 2050;
};

inline tts_transformer_config from_abi(const qtts_config & c) {
    tts_transformer_config r;
    r.text_vocab_size = c.text_vocab_size; r.text_embd_dim = c.text_embd_dim; r.hidden_size = c.hidden_size;
    r.n_layers = c.n_layers; r.n_attention_heads = c.n_heads; r.n_key_value_heads = c.n_kv_heads;
    r.intermediate_size = c.intermediate_size; r.head_dim = c.head_dim; r.rms_norm_eps = c.rms_norm_eps;
    r.rope_theta = c.rope_theta; r.codec_vocab_size = c.codec_vocab_size; r.n_codebooks = c.n_codebooks;
    r.code_pred_layers = c.cp_layers; r.code_pred_vocab_size = c.cp_vocab_size; r.code_pred_hidden_size = c.cp_hidden_size;
    r.code_pred_intermediate_size = c.cp_intermediate_size; r.code_pred_n_attention_heads = c.cp_n_heads;
    r.code_pred_n_kv_heads = c.cp_n_kv_heads; r.code_pred_head_dim = c.cp_head_dim;
    r.codec_pad_id = c.codec_pad_id; r.codec_bos_id = c.codec_bos_id; r.codec_eos_id = c.codec_eos_id;
    r.tts_bos_token_id = c.tts_bos_id; r.tts_eos_token_id = c.tts_eos_id; r.tts_pad_token_id = c.tts_pad_id;
    r.codec_think_id = c.codec_think_id; r.codec_nothink_id = c.codec_nothink_id;
    r.codec_think_bos_id = c.codec_think_bos_id; r.codec_think_eos_id = c.codec_think_eos_id;
    return r;
}

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

    TTSTransformer() : seed_(std::random_device{}()) {}      // like rng_{std::random_device{}()}, src/tts_transformer.h:298
    ~TTSTransformer() { unload_model(); }
    TTSTransformer(const TTSTransformer &) = delete;
    TTSTransformer & operator=(const TTSTransformer &) = delete;

    // engine-creation knobs without a counterpart in the reference (it owns one stream on one device)
    void set_device(int device) { device_ = device; }
    void set_max_streams(int n) { max_streams_ = n; }
    // Sampling is seeded from std::random_device like the reference's rng_; set_seed makes a run reproducible.
    // Every generate() call advances the seed, so repeated calls do not replay one uniform sequence.
    void set_seed(uint64_t seed) { seed_ = seed; }

    // load_model (src/tts_transformer.cpp:51-286): parse config, create tensors, copy the data
    bool load_model(const std::string & model_path) {
        unload_model();
        GgufFile g;
        if (!g.open(model_path, error_msg_)) return false;
        abi_cfg_ = qtts_config{};
        detail::parse_config(g, abi_cfg_);
        cfg_ = from_abi(abi_cfg_);
        model_path_ = model_path;
        return create_engine(g);
    }

    // unload_model (src/tts_transformer.cpp:2576-2640)
    void unload_model() {
        if (eng_) qtts_destroy(eng_);
        eng_ = nullptr;
        last_hidden_.clear();
    }

    // init_kv_cache (src/tts_transformer.cpp:794-842): the reference frees and re-allocates the cache tensors; here
    // qtts_resize_kv re-sizes the per-stream slabs and leaves the weights where they are
    bool init_kv_cache(int32_t n_ctx) {
        if (!eng_) { error_msg_ = "Model not loaded"; return false; }
        if (n_ctx <= 0) { error_msg_ = "init_kv_cache: n_ctx must be positive"; return false; }
        if (n_ctx != n_ctx_) {
            if (!check(qtts_resize_kv(eng_, std::max(n_ctx, 32), max_text_))) return false;     // on failure the old cache stays
            n_ctx_ = std::max(n_ctx, 32);
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
    void clear_code_pred_kv_cache() {}      // positions above the current one are never read

    // forward_text (src/tts_transformer.cpp:1922-1950): output = hidden states [n_tokens, hidden_size]
    bool forward_text(const int32_t * text_tokens, int32_t n_tokens, const float * speaker_embd, int32_t n_past,
                      std::vector<float> & output) {
        if (!eng_) { error_msg_ = "Model not loaded"; return false; }
        if (!text_tokens) { error_msg_ = "text_tokens is null"; return false; }
        if (n_tokens <= 0) { error_msg_ = "n_tokens must be > 0"; return false; }
        output.resize((size_t) n_tokens * cfg_.hidden_size);
        if (!check(qtts_forward_text(eng_, 0, text_tokens, n_tokens, speaker_embd, n_past, output.data()))) return false;
        last_hidden_.assign(output.end() - cfg_.hidden_size, output.end());
        return true;
    }

    // forward_prefill (src/tts_transformer.cpp:1829-1920): output = normalised hidden of EVERY token
    // [n_tokens, hidden_size] (:1893), logits_out = logits of the last token
    bool forward_prefill(const float * prefill_embd, int32_t n_tokens, int32_t n_past,
                         std::vector<float> & output, std::vector<float> * logits_out = nullptr) {
        if (!eng_) { error_msg_ = "Model not loaded"; return false; }
        if (!prefill_embd) { error_msg_ = "prefill_embd is null"; return false; }
        if (n_tokens <= 0) { error_msg_ = "n_tokens must be > 0"; return false; }
        output.resize((size_t) n_tokens * cfg_.hidden_size);
        std::vector<float> logits(cfg_.codec_vocab_size);
        if (!check(qtts_forward_prefill(eng_, 0, prefill_embd, n_tokens, n_past, output.data(), logits.data()))) return false;
        last_hidden_.assign(output.end() - cfg_.hidden_size, output.end());
        if (logits_out) *logits_out = std::move(logits);
        return true;
    }

    // forward_codec (src/tts_transformer.cpp:2030-2040): output = codec logits
    bool forward_codec(int32_t codec_token, int32_t n_past, std::vector<float> & output) {
        if (!eng_) { error_msg_ = "Model not loaded"; return false; }
        output.resize(cfg_.codec_vocab_size);
        if (!check(qtts_forward_codec(eng_, 0, codec_token, n_past, output.data()))) return false;
        last_hidden_.resize(cfg_.hidden_size);
        return check(qtts_get_hidden_states(eng_, 0, last_hidden_.data()));
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

    // get_hidden_states (src/tts_transformer.cpp:2042-2048)
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
        if (!fit(n_tokens, 0)) return false;
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
        output.resize(15);                  // always 15, whatever n_codebooks says (:2192)
        return check(qtts_predict_codes(eng_, 0, hidden, codebook_0_token, temperature, top_k,
                                        nullptr, nullptr, output.data(), nullptr));
    }

    // generate (src/tts_transformer.cpp:2342-2574)
    bool generate(const int32_t * text_tokens, int32_t n_tokens, const float * speaker_embd,
                  int32_t max_len, std::vector<int32_t> & output, int32_t language_id = 2050,
                  float repetition_penalty = 1.05f, float temperature = 0.9f, int32_t top_k = 50,
                  frame_callback_t on_frames = nullptr, int32_t callback_interval = 40) {
        if (!eng_) { error_msg_ = "Model not loaded"; return false; }
        if (!text_tokens) { error_msg_ = "text_tokens is null"; return false; }
        if (n_tokens < 4) { error_msg_ = "Need at least 4 text tokens for generation"; return false; }
        if (max_len <= 0) { output.clear(); return true; }                   // :2364-2367
        if (!fit(n_tokens, max_len)) return false;
        output.assign((size_t) max_len * cfg_.n_codebooks, 0);
        int32_t n_frames = 0;
        const int rc = qtts_generate(eng_, 0, text_tokens, n_tokens, speaker_embd, max_len, language_id,
                                     repetition_penalty, temperature, top_k, seed_, output.data(), &n_frames,
                                     on_frames ? &TTSTransformer::trampoline : nullptr, callback_interval,
                                     on_frames ? (void *) &on_frames : nullptr);
        seed_ = seed_ * 6364136223846793005ull + 1442695040888963407ull;       // the next call draws a fresh sequence
        if (!check(rc)) { output.clear(); return false; }
        output.resize((size_t) n_frames * cfg_.n_codebooks);
        last_hidden_.resize(cfg_.hidden_size);
        qtts_get_hidden_states(eng_, 0, last_hidden_.data());
        return true;
    }

    const tts_transformer_config & get_config() const { return cfg_; }
    const std::string & get_error() const { return error_msg_; }
    qtts_engine * engine() const { return eng_; }
    const qtts_config & abi_config() const { return abi_cfg_; }

private:
    static int trampoline(const int32_t * codes, int32_t n_frames, int32_t n_codebooks, void * user) {
        return (*static_cast<frame_callback_t *>(user))(codes, n_frames, n_codebooks) ? 1 : 0;
    }
    bool check(int rc) {
        if (rc != 0) return true;   // C ABI: 1 = success, 0 = failure
        error_msg_ = eng_ ? qtts_last_error(eng_) : qtts_create_error();
        return false;
    }
    // The reference re-sizes its KV cache from the request: required = prefill_len + max_len + 8, re-allocated when
    // the cache is too small or more than max(2 * required, 512) (src/tts_transformer.cpp:2383-2388); prefill_len is
    // at most 11.  Same policy, through qtts_resize_kv (weights untouched); a request beyond the engine's cap fails
    // BEFORE anything is freed, the loaded engine stays usable.
    bool fit(int32_t n_tokens, int32_t max_len) {
        const int32_t required = 11 + max_len + 8;
        int32_t n_ctx = n_ctx_, max_text = max_text_;
        if (n_ctx < required || n_ctx > std::max<int32_t>(required * 2, 512)) n_ctx = std::max(required, 32);
        if (n_tokens > max_text) max_text = std::max(n_tokens, 2 * max_text);
        if (n_ctx == n_ctx_ && max_text == max_text_) return true;
        if (!check(qtts_resize_kv(eng_, n_ctx, max_text))) return false;
        n_ctx_ = n_ctx; max_text_ = max_text;
        return true;
    }
    bool create_engine(GgufFile & g) {
        if (qtts_create(&abi_cfg_, device_, max_streams_, n_ctx_, &eng_) == 0) {
            error_msg_ = qtts_create_error();
            eng_ = nullptr;
            return false;
        }
        // tensors: the library's own loader (mapped file -> pinned staging ring -> HBM; quantised types rejected by name)
        (void) g;
        if (qtts_load_gguf(eng_, model_path_.c_str()) == 0) {
            error_msg_ = qtts_last_error(eng_);
            unload_model();
            return false;
        }
        if (qtts_finalize(eng_) == 0) {
            error_msg_ = qtts_last_error(eng_);
            unload_model();
            return false;
        }
        return true;
    }

    qtts_engine * eng_ = nullptr;
    qtts_config abi_cfg_{};
    tts_transformer_config cfg_{};
    std::string model_path_, error_msg_;
    std::vector<float> last_hidden_;
    int device_ = 0, max_streams_ = 1;
    int32_t n_ctx_ = 512, max_text_ = 512;
    uint64_t seed_;
};

// Drop-in for qwen3_tts::TRTCodePredictor (src/trt_code_predictor.h:25-87): the whole method set.  The reference
// keeps this second, TensorRT-built copy of the code predictor next to the transformer (src/tts_transformer.h:304-311)
// and feeds it from load_model (src/tts_transformer.cpp:155-230: load_engine, upload_lm_heads, upload_mtp_proj,
// upload_codec_embeddings, warmup_cublas).  Here the code predictor lives in the same engine as the Talker:
//   * attach(transformer): share the transformer's engine (the INTEGRATION.md section 2 hook);
//   * load_engine(path): stand-alone - `path` is the model GGUF (there is no serialized .trt engine to load);
//   * upload_*: the engine already holds these tensors from the GGUF, so the calls VALIDATE the shapes the
//     reference would upload (vocab x hidden heads, cp_hidden x talker_hidden projection, 15 embedding tables of
//     talker width) and keep nothing.
class TRTCodePredictor {
public:
    TRTCodePredictor() = default;
    explicit TRTCodePredictor(TTSTransformer & owner) : tr_(&owner) {}
    ~TRTCodePredictor() { unload(); }

    void attach(TTSTransformer & owner) { unload(); tr_ = &owner; }

    bool load_engine(const std::string & engine_path) {
        unload();
        own_.reset(new TTSTransformer());
        if (!own_->load_model(engine_path)) { error_msg_ = "Failed to load code predictor: " + own_->get_error(); own_.reset(); return false; }
        tr_ = own_.get();
        return true;
    }

    bool upload_lm_heads(const std::vector<std::vector<float>> & weights, int32_t vocab_size, int32_t hidden_size) {
        if (!ready()) return false;
        const auto & c = tr_->get_config();
        if ((int) weights.size() != 15 || vocab_size != c.code_pred_vocab_size || hidden_size != c.code_pred_hidden_size)
            return fail("upload_lm_heads: shape does not match the loaded model");
        for (const auto & w : weights) if (w.size() != (size_t) vocab_size * hidden_size) return fail("upload_lm_heads: bad weight size");
        return true;
    }
    bool upload_mtp_proj(const std::vector<float> & weight, const std::vector<float> & bias, int32_t cp_hidden, int32_t talker_hidden) {
        if (!ready()) return false;
        const auto & c = tr_->get_config();
        if (cp_hidden != c.code_pred_hidden_size || talker_hidden != c.hidden_size || weight.size() != (size_t) cp_hidden * talker_hidden ||
            (!bias.empty() && bias.size() != (size_t) cp_hidden))
            return fail("upload_mtp_proj: shape does not match the loaded model");
        if (c.code_pred_hidden_size == c.hidden_size) return fail("upload_mtp_proj: the loaded model has no mtp_proj");
        return true;
    }
    bool upload_codec_embeddings(const std::vector<std::vector<float>> & weights, int32_t embd_dim) {
        if (!ready()) return false;
        const auto & c = tr_->get_config();
        if (weights.empty() || weights.size() > 15 || embd_dim != c.hidden_size) return fail("upload_codec_embeddings: shape does not match the loaded model");
        for (size_t i = 0; i < weights.size(); ++i) {
            const size_t vocab = i == 0 ? c.codec_vocab_size : c.code_pred_vocab_size;
            if (weights[i].size() != vocab * (size_t) embd_dim) return fail("upload_codec_embeddings: bad table size");
        }
        return true;
    }

    void reset_kv_cache() {}                // positions above the current one are never read
    void warmup_cublas() {}                 // nothing to warm: no cuBLAS, no JIT

    // one position with a host input vector (src/trt_code_predictor.cpp:384-424)
    bool step(const float * input, int32_t input_size, int32_t position_id, int32_t lm_head_idx, bool needs_proj,
              float * output, int32_t output_size) {
        if (!ready()) return false;
        if (qtts_cp_step(tr_->engine(), 0, input, input_size, 0, 0, position_id, lm_head_idx, needs_proj ? 1 : 0, output, output_size) == 0)
            return fail(qtts_last_error(tr_->engine()));
        return true;
    }
    // one position from a device-resident embedding row (src/trt_code_predictor.cpp:426-482)
    bool step_with_token(int32_t token_id, int32_t embedding_idx, int32_t position_id, int32_t lm_head_idx,
                         float * output, int32_t output_size) {
        if (!ready()) return false;
        if (qtts_cp_step(tr_->engine(), 0, nullptr, 0, token_id, embedding_idx, position_id, lm_head_idx, 1, output, output_size) == 0)
            return fail(qtts_last_error(tr_->engine()));
        return true;
    }

    // run_greedy_loop (src/trt_code_predictor.cpp:484-600).  needs_proj: the 1.7B Talker->code
    // predictor projection (mtp_proj) is part of the engine's program whenever the model has one,
    // so the flag is only checked for consistency with the reference call.
    bool run_greedy_loop(const float * hidden, int32_t hidden_size, int32_t cb0_token, bool needs_proj,
                         int32_t * output_tokens, int32_t n_outputs) {
        return run(hidden, hidden_size, cb0_token, needs_proj, 0.0f, 0, nullptr, output_tokens, n_outputs);
    }
    // run_sampling_loop (src/trt_code_predictor.cpp:602-726): rand_vals = one uniform in [0,1) per code.
    // Sampling rule: the GGML path's (mask `< k-th largest`, ties kept, softmax with a double sum, CDF scan against
    // the uniform) - NOT k_sample_topk_f32's -1e30 mask and `rand * sum` scan (src/trt_cuda_kernels.cu:124,172);
    // both select from the same candidate set, the id can differ when the uniform falls within float rounding of a step.
    bool run_sampling_loop(const float * hidden, int32_t hidden_size, int32_t cb0_token, bool needs_proj,
                           float temperature, int32_t top_k, const float * rand_vals,
                           int32_t * output_tokens, int32_t n_outputs) {
        return run(hidden, hidden_size, cb0_token, needs_proj, temperature, top_k, rand_vals, output_tokens, n_outputs);
    }

    bool is_loaded() const { return tr_ && tr_->engine() != nullptr; }
    bool is_ready() const { return is_loaded(); }
    const std::string & get_error() const { return error_msg_; }
    void unload() { own_.reset(); tr_ = nullptr; }

private:
    bool fail(const std::string & m) { error_msg_ = m; return false; }
    bool ready() { if (is_loaded()) return true; return fail("TRT code predictor not loaded"); }
    bool run(const float * hidden, int32_t hidden_size, int32_t cb0, bool needs_proj, float temperature, int32_t top_k,
             const float * rand_vals, int32_t * out, int32_t n_outputs) {
        if (!ready()) return false;
        const auto & c = tr_->get_config();
        if (hidden_size != c.hidden_size) return fail("hidden_size does not match the model");
        if (needs_proj != (c.code_pred_hidden_size != c.hidden_size)) return fail("needs_proj does not match the model");
        if (n_outputs != c.n_codebooks - 1) return fail("n_outputs must be n_codebooks - 1");
        if (qtts_predict_codes(tr_->engine(), 0, hidden, cb0, temperature, top_k, rand_vals, nullptr, out, nullptr) == 0)
            return fail(qtts_last_error(tr_->engine()));
        return true;
    }
    TTSTransformer * tr_ = nullptr;
    std::unique_ptr<TTSTransformer> own_;
    std::string error_msg_;
};
using CodePredictor = TRTCodePredictor;     // round-1 name

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
