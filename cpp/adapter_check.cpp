// adapter_check.cpp — drives the C++ host adapter (qtts_transformer.hpp) the way the reference's
// Qwen3TTS::synthesize_internal drives its TTSTransformer (src/qwen3_tts.cpp:459-463), and prints
// what it got as JSON on stdout so tests/test_cpp_adapter.py can compare it with the ctypes path.
//
//   adapter_check <model.gguf> <tokens.i32> <max_len> <temperature> <top_k> <seed>
//
// Build: g++ -std=c++17 -O2 cpp/adapter_check.cpp -o cpp/adapter_check -L<pkg> -lqtts_b200 -Wl,-rpath,<pkg>
#include <cstdio>
#include <cstdlib>

#include "qtts_transformer.hpp"

using qtts_b200::TRTCodePredictor;
using qtts_b200::TTSTransformer;

static void print_ints(const char * key, const std::vector<int32_t> & v, bool comma = true) {
    printf("\"%s\": [", key);
    for (size_t i = 0; i < v.size(); ++i) printf("%s%d", i ? "," : "", v[i]);
    printf("]%s\n", comma ? "," : "");
}

// --config <model.gguf>: the GGUF reader and parse_config alone (no GPU needed)
static int dump_config(const char * path) {
    qtts_b200::GgufFile g;
    std::string err;
    if (!g.open(path, err)) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
    qtts_config c{};
    qtts_b200::detail::parse_config(g, c);
    size_t bytes = 0;
    for (const auto & t : g.tensors()) bytes += t.n_bytes;
    std::vector<uint8_t> first;
    if (!g.tensors().empty() && !g.read_tensor(g.tensors()[0], first, err)) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
    unsigned long sum = 0;
    for (uint8_t b : first) sum += b;
    printf("{\"hidden_size\": %d, \"n_layers\": %d, \"n_heads\": %d, \"n_kv_heads\": %d, \"intermediate_size\": %d, "
           "\"head_dim\": %d, \"codec_vocab_size\": %d, \"n_codebooks\": %d, \"cp_layers\": %d, \"cp_vocab_size\": %d, "
           "\"cp_hidden_size\": %d, \"cp_intermediate_size\": %d, \"cp_n_heads\": %d, \"cp_n_kv_heads\": %d, \"cp_head_dim\": %d, "
           "\"text_vocab_size\": %d, \"text_embd_dim\": %d, \"rms_norm_eps\": %.9g, \"rope_theta\": %.9g, "
           "\"codec_pad_id\": %d, \"codec_bos_id\": %d, \"codec_eos_id\": %d, \"tts_bos_id\": %d, \"tts_eos_id\": %d, \"tts_pad_id\": %d, "
           "\"n_tensors\": %zu, \"tensor_bytes\": %zu, \"first_tensor\": \"%s\", \"first_tensor_bytesum\": %lu}\n",
           c.hidden_size, c.n_layers, c.n_heads, c.n_kv_heads, c.intermediate_size, c.head_dim, c.codec_vocab_size, c.n_codebooks,
           c.cp_layers, c.cp_vocab_size, c.cp_hidden_size, c.cp_intermediate_size, c.cp_n_heads, c.cp_n_kv_heads, c.cp_head_dim,
           c.text_vocab_size, c.text_embd_dim, c.rms_norm_eps, c.rope_theta, c.codec_pad_id, c.codec_bos_id, c.codec_eos_id,
           c.tts_bos_id, c.tts_eos_id, c.tts_pad_id, g.tensors().size(), bytes,
           g.tensors().empty() ? "" : g.tensors()[0].name.c_str(), sum);
    return 0;
}

// --tensors <model.gguf>: walk the tensor table the way the library's loader does (type check per tensor, no GPU)
static int check_tensors(const char * path) {
    qtts_b200::GgufFile g;
    std::string err;
    if (!g.open(path, err) || !g.map(err)) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
    size_t n = 0;
    for (const auto & t : g.tensors()) {
        if (t.ggml_type != 0 && t.ggml_type != 1) {
            fprintf(stderr, "tensor %s: unsupported GGUF type %s (only F32/F16; quantised GGUF files are rejected)\n",
                    t.name.c_str(), qtts_b200::GgufFile::type_name(t.ggml_type));
            return 1;
        }
        if (!g.data(t)) { fprintf(stderr, "tensor %s lies outside the file\n", t.name.c_str()); return 1; }
        ++n;
    }
    printf("{\"n_tensors\": %zu}\n", n);
    return 0;
}

int main(int argc, char ** argv) {
    if (argc == 3 && std::string(argv[1]) == "--config") return dump_config(argv[2]);
    if (argc == 3 && std::string(argv[1]) == "--tensors") return check_tensors(argv[2]);
    if (argc < 7) {
        fprintf(stderr, "usage: %s model.gguf tokens.i32 max_len temperature top_k seed\n", argv[0]);
        return 2;
    }
    const std::string model = argv[1];
    const int max_len = atoi(argv[3]);
    const float temperature = (float) atof(argv[4]);
    const int top_k = atoi(argv[5]);
    const uint64_t seed = strtoull(argv[6], nullptr, 10);

    std::vector<int32_t> tokens;
    if (FILE * f = fopen(argv[2], "rb")) {
        int32_t t;
        while (fread(&t, sizeof(t), 1, f) == 1) tokens.push_back(t);
        fclose(f);
    }

    TTSTransformer tr;
    // error behaviour first: a bad path must fail with the reference's message
    if (tr.load_model("/nonexistent/model.gguf") || tr.get_error().find("Failed to open GGUF file") == std::string::npos) {
        fprintf(stderr, "bad-path load did not fail as expected: %s\n", tr.get_error().c_str());
        return 1;
    }
    tr.set_seed(seed);
    if (!tr.load_model(model)) {
        // no GPU / no sm_100 device: report and exit 3 so the CPU test can assert the loud failure
        printf("{\"load_error\": \"%s\"}\n", tr.get_error().c_str());
        return 3;
    }
    const auto & cfg = tr.get_config();

    // 1. generate() with a frame callback
    std::vector<int32_t> codes;
    int cb_calls = 0, cb_frames = 0;
    auto on_frames = [&](const int32_t *, int32_t n_frames, int32_t n_cb) {
        ++cb_calls;
        cb_frames += n_frames;
        return n_cb == cfg.n_codebooks;
    };
    if (!tr.generate(tokens.data(), (int32_t) tokens.size(), nullptr, max_len, codes, 2050, 1.05f, temperature, top_k,
                     on_frames, 4)) {
        fprintf(stderr, "generate failed: %s\n", tr.get_error().c_str());
        return 1;
    }

    // 2. the same first frame by hand: build_prefill -> forward_prefill -> greedy cb0 -> code predictor loop
    std::vector<float> prefill, trailing, pad, hidden, logits;
    if (!tr.init_kv_cache(256) || !tr.build_prefill(tokens.data(), (int32_t) tokens.size(), nullptr, 2050, prefill, trailing, pad)) {
        fprintf(stderr, "build_prefill failed: %s\n", tr.get_error().c_str());
        return 1;
    }
    const int n_prefill = (int) (prefill.size() / cfg.hidden_size);
    if (!tr.forward_prefill(prefill.data(), n_prefill, 0, hidden, &logits)) {
        fprintf(stderr, "forward_prefill failed: %s\n", tr.get_error().c_str());
        return 1;
    }
    // greedy over the unsuppressed range (generate(): src/tts_transformer.cpp:2432-2440 keeps [0, vocab-1024) + EOS)
    int cb0 = 0;
    for (int i = 1; i < cfg.codec_vocab_size - 1024; ++i) if (logits[i] > logits[cb0]) cb0 = i;
    if ((int) hidden.size() != n_prefill * cfg.hidden_size) { fprintf(stderr, "forward_prefill: output is not [n_tokens x H]\n"); return 1; }
    const float * last_hidden = hidden.data() + (size_t) (n_prefill - 1) * cfg.hidden_size;
    TRTCodePredictor cp(tr);
    const bool needs_proj = cfg.code_pred_hidden_size != cfg.hidden_size;
    std::vector<int32_t> greedy15(cfg.n_codebooks - 1), via_class, stepwise;
    if (!cp.run_greedy_loop(last_hidden, cfg.hidden_size, cb0, needs_proj, greedy15.data(),
                            (int32_t) greedy15.size())) {
        fprintf(stderr, "run_greedy_loop failed: %s\n", cp.get_error().c_str());
        return 1;
    }
    if (!tr.predict_codes_autoregressive(last_hidden, cb0, via_class, 0.0f, 0)) {
        fprintf(stderr, "predict_codes_autoregressive failed: %s\n", tr.get_error().c_str());
        return 1;
    }
    // the same 15 codes position by position through step / step_with_token, the way the reference's first TRT
    // integration drove the engine (src/trt_code_predictor.h:53-66): pos 0 = hidden (no head), pos 1 = codec_embd[cb0]
    // -> head 0, pos s + 1 = code_pred_embd[s - 1][code] -> head s
    {
        std::vector<float> lg(cfg.code_pred_vocab_size), hid(cfg.code_pred_hidden_size);
        cp.reset_kv_cache();
        if (!cp.step(last_hidden, cfg.hidden_size, 0, -1, needs_proj, hid.data(), (int32_t) hid.size())) {
            fprintf(stderr, "step failed: %s\n", cp.get_error().c_str());
            return 1;
        }
        int tok = cb0;
        for (int s = 0; s < cfg.n_codebooks - 1; ++s) {
            if (!cp.step_with_token(tok, s, s + 1, s, lg.data(), (int32_t) lg.size())) {
                fprintf(stderr, "step_with_token failed: %s\n", cp.get_error().c_str());
                return 1;
            }
            int best = 0;
            for (int i = 1; i < cfg.code_pred_vocab_size; ++i) if (lg[i] > lg[best]) best = i;
            stepwise.push_back(best);
            tok = greedy15[s];          // teacher-forced on the loop's codes: every position is comparable
        }
    }
    // argument checking mirrors the reference's (wrong width is an error, not UB)
    int32_t dummy[15];
    const bool bad_ok = cp.run_greedy_loop(last_hidden, cfg.hidden_size + 1, cb0, false, dummy, 15);
    // the upload_* calls validate the shapes the reference's load_model block would pass (src/tts_transformer.cpp:186-230)
    std::vector<std::vector<float>> heads(15, std::vector<float>((size_t) cfg.code_pred_vocab_size * cfg.code_pred_hidden_size));
    const bool up_ok = cp.upload_lm_heads(heads, cfg.code_pred_vocab_size, cfg.code_pred_hidden_size) &&
                       !cp.upload_lm_heads(heads, cfg.code_pred_vocab_size + 1, cfg.code_pred_hidden_size);
    // a second generate() must not replay the first one's uniforms (the seed advances), and a request beyond the
    // engine's context cap must fail without losing the loaded model
    std::vector<int32_t> codes2, codes3;
    tr.generate(tokens.data(), (int32_t) tokens.size(), nullptr, max_len, codes2, 2050, 1.05f, temperature, top_k);
    const bool too_long_fails = !tr.generate(tokens.data(), (int32_t) tokens.size(), nullptr, 100000, codes3);
    const bool still_loaded = tr.generate(tokens.data(), (int32_t) tokens.size(), nullptr, 4, codes3, 2050, 1.05f, 0.0f, 0) && codes3.size() == 4u * cfg.n_codebooks;

    printf("{\n");
    printf("\"n_codebooks\": %d,\n\"n_frames\": %d,\n\"cb_calls\": %d,\n\"cb_frames\": %d,\n\"n_prefill\": %d,\n\"cb0\": %d,\n",
           cfg.n_codebooks, (int) (codes.size() / cfg.n_codebooks), cb_calls, cb_frames, n_prefill, cb0);
    printf("\"bad_width_rejected\": %s,\n", bad_ok ? "false" : "true");
    printf("\"upload_validation\": %s,\n\"too_long_fails\": %s,\n\"still_loaded\": %s,\n", up_ok ? "true" : "false",
           too_long_fails ? "true" : "false", still_loaded ? "true" : "false");
    print_ints("codes_second_call", codes2);
    print_ints("stepwise15", stepwise);
    print_ints("codes", codes);
    print_ints("greedy15", greedy15);
    print_ints("predict15", via_class, false);
    printf("}\n");
    return 0;
}
