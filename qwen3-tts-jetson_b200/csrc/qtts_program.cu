// qtts_program.cu — host-side builders of the per-frame stage programs and the launcher.
//
// One codec frame (reference: the body of the loop at src/tts_transformer.cpp:2416-2560) is
//   SAMPLE_CB0 -> code predictor (16 positions x [5 layers x 5 stages] + 15 x [head, sample])
//   -> FRAME_END -> Talker step (28 layers x 5 stages + head), whose first stage gathers the
//   16 code embeddings + text row (src/tts_transformer.cpp:2529-2553).
#include "qtts_engine.h"
#include <algorithm>
#include <utility>
#include <cstring>
#include <cstdlib>

namespace qtts {

extern "C" __global__ void qtts_run_program(const Op * ops, int n_ops, StreamState * st, GridBarrier * bar,
                                            int n_frames, RunCfg cfg);

static Op blank() {
    Op o;
    memset(&o, 0, sizeof(o));
    o.temperature_override = -1e30f;
    o.n_tok = 1;
    o.tok_slot = -1;
    return o;
}

static Op gemv(const wt_t * W, int rows, int K, const float * in, const float * norm_w, float eps,
               int epi, float * out) {
    Op o = blank();
    o.type = OP_GEMV; o.W = W; o.rows = rows; o.K = K; o.in_kind = IN_VEC; o.in = in;
    o.norm_w = norm_w; o.eps = eps; o.epi = epi; o.out = out;
    return o;
}

struct LayerBuf { float * x, * qkv, * attn, * act; __half * kc, * vc; size_t kv_layer_stride; float * attn_part; unsigned int * attn_cnt;
                  unsigned long long * x_ll, * act_ll, * qkv_ll, * attn_ll; };

// five stages of one decoder layer over n_tok tokens (2 only for the merged code-predictor
// positions 0/1); `first` (if non-null) replaces the QKV stage's input setup
static void emit_layer(Engine * e, std::vector<Op> & P, const LayerW & L, const Dims & D, const LayerBuf & B,
                       int layer, const Op * first, bool use_n_past, int pos_off, int mode, int n_tok = 1) {
    const bool ll = e->ll != 0, ll_in = layer > 0;
    const float eps = e->cfg.rms_norm_eps;
    const int qkv_rows = (D.nh + 2 * D.nkv) * kHeadDim;
    const int aw = D.nh * kHeadDim;
    Op q = gemv(L.wqkv, qkv_rows, D.H, B.x, L.attn_norm, eps, EPI_STORE, B.qkv);
    q.n_tok = n_tok; q.in_stride = D.H; q.out_stride = qkv_rows;
    if (first) {
        q.in_kind = first->in_kind; q.in = first->in; q.table = first->table; q.tok_ptr = first->tok_ptr;
        q.tok_slot = first->tok_slot; q.es = first->es; q.x_out = first->x_out;
    } else if (ll && ll_in) {
        q.in_ll = B.x_ll; q.in_dist = 1;      // previous layer's down projection
    }
    // the attention stage polls the projections; without a barrier in between, the residual the O
    // projection adds must be flag-carrying too (x_out of a first layer, else the previous down projection)
    const bool attn_ll_out = ll && mode != ATTN_FAST;     // ATTN_FAST merges split-KV partials across CTAs: barrier stays
    int o_res_dist = 0;
    if (ll) {
        q.out_ll = B.qkv_ll; q.flags |= F_LL_OUT;
        if (attn_ll_out) {
            if (q.x_out) { q.x_out_ll = B.x_ll; o_res_dist = 2; }
            else if (!first && ll_in) o_res_dist = 3;
        }
    }
    P.push_back(q);
    Op a = blank();
    a.type = OP_ATTN; a.qkv = B.qkv; a.q_norm_w = L.q_norm; a.k_norm_w = L.k_norm;
    a.kcache = B.kc + (size_t) layer * B.kv_layer_stride;
    a.vcache = B.vc + (size_t) layer * B.kv_layer_stride;
    a.rope = e->rope; a.n_heads = D.nh; a.n_kv_heads = D.nkv; a.pos_off = pos_off; a.use_n_past = use_n_past ? 1 : 0;
    a.mode = mode; a.ref_threads = e->ref_threads; a.kq_scale = 1.0f / sqrtf((float) kHeadDim);
    a.eps = eps; a.out = B.attn;
    a.n_tok = n_tok; a.qkv_stride = qkv_rows; a.out_stride = aw;
    a.attn_part = B.attn_part; a.attn_cnt = B.attn_cnt;
    if (ll) { a.in_ll = B.qkv_ll; a.in_dist = 1; }
    // (a first layer fed by a plain vector without x_out - the mtp_proj case - keeps the barrier)
    const bool a_ll_out = attn_ll_out && o_res_dist != 0;
    if (a_ll_out) { a.out_ll = B.attn_ll; a.flags |= F_LL_OUT; }
    P.push_back(a);
    Op o = gemv(L.wo, D.H, aw, B.attn, nullptr, eps, EPI_ADD, B.x);
    o.res = B.x; o.n_tok = n_tok; o.in_stride = aw; o.out_stride = D.H;
    if (ll) { o.out_ll = B.x_ll; o.flags |= F_LL_OUT; }
    if (a_ll_out) { o.in_ll = B.attn_ll; o.in_dist = 1; o.res_ll = B.x_ll; o.res_dist = o_res_dist; }
    P.push_back(o);
    Op gu = gemv(L.wgu, 2 * D.ffn, D.H, B.x, L.ffn_norm, eps, EPI_SWIGLU, B.act);
    gu.n_tok = n_tok; gu.in_stride = D.H; gu.out_stride = D.ffn;
    if (ll) { gu.in_ll = B.x_ll; gu.in_dist = 1; gu.out_ll = B.act_ll; gu.flags |= F_LL_OUT; }
    P.push_back(gu);
    Op d = gemv(L.wdown, D.H, D.ffn, B.act, nullptr, eps, EPI_ADD, B.x);
    d.res = B.x; d.n_tok = n_tok; d.in_stride = D.ffn; d.out_stride = D.H;
    // the stage after a layer's down projection is always a GEMV (next layer's QKV or the head)
    if (ll) { d.in_ll = B.act_ll; d.in_dist = 1; d.res_ll = B.x_ll; d.res_dist = 2; d.out_ll = B.x_ll; d.flags |= F_LL_OUT; }
    P.push_back(d);
}

// One Talker token: forward_step graph (src/tts_transformer.cpp:1376-1512).
void emit_talker_step(Engine * e, Slot & s, std::vector<Op> & P, int in_kind, const float * in_vec,
                      bool use_n_past, int pos_abs, int attn_mode, bool with_head, bool inc_npast) {
    const qtts_config & c = e->cfg;
    const Dims D{c.hidden_size, c.n_heads, c.n_kv_heads, c.intermediate_size};
    LayerBuf B{s.x, s.qkv, s.attn, s.act, s.kc, s.vc, (size_t) e->max_ctx * c.n_kv_heads * kHeadDim, s.attn_part, s.attn_cnt,
               s.x_ll, s.act_ll, s.qkv_ll, s.attn_ll};
    Op first = blank();
    first.in_kind = in_kind; first.in = in_vec; first.es = s.es; first.x_out = s.x;
    for (int l = 0; l < c.n_layers; ++l)
        emit_layer(e, P, e->talker[l], D, B, l, l == 0 ? &first : nullptr, use_n_past, use_n_past ? 0 : pos_abs, attn_mode);
    if (with_head) {
        Op h = gemv(e->codec_head, c.codec_vocab_size, c.hidden_size, s.x, e->output_norm, c.rms_norm_eps, EPI_STORE, s.logits);
        h.hid_out = s.hidden;
        if (e->ll) { h.in_ll = s.x_ll; h.in_dist = 1; }
        if (inc_npast) h.flags |= F_INC_NPAST;
        P.push_back(h);
    } else {
        // no head: nothing polls the last down projection, the stage ends with a grid barrier again
        P.back().flags &= ~F_LL_OUT;
        if (inc_npast) P.back().flags |= F_INC_NPAST;
    }
}

// predict_codes_autoregressive (src/tts_transformer.cpp:2153-2340): positions 0 and 1 as ONE
// two-token pass with the all-F32 attention of the reference's prefill graph (:1514-1664,
// :1613-1626), then 14 single-token passes with the F16-cache / F16-P attention of its step graph
// (:1666-1827).  Sampling stages are not followed by a grid barrier (every CTA samples).
void emit_code_predictor(Engine * e, Slot & s, std::vector<Op> & P, const float * hidden,
                         float * dump, const int32_t * forced, const float * uniforms,
                         bool override_sampling, float temperature, int top_k) {
    const qtts_config & c = e->cfg;
    const Dims D{c.cp_hidden_size, c.cp_n_heads, c.cp_n_kv_heads, c.cp_intermediate_size};
    LayerBuf B{s.cp_x, s.cp_qkv, s.cp_attn, s.cp_act, s.cp_kc, s.cp_vc, (size_t) 16 * c.cp_n_kv_heads * kHeadDim,
               s.attn_part, s.attn_cnt, s.cp_x_ll, s.cp_act_ll, s.cp_qkv_ll, s.cp_attn_ll};
    const int n_pos = c.n_codebooks;   // 16 positions -> 15 sampled codes
    for (int p = 1; p < n_pos; ++p) {
        const int n_tok = (p == 1) ? 2 : 1;
        Op first = blank();
        if (p == 1) { first.in_kind = IN_CP_FIRST; first.in = hidden; first.table = e->codec_embd; }
        else { first.in_kind = IN_EMBED_ROW; first.table = e->cp_embd[p - 2]; first.tok_slot = p - 1; }
        if (e->mtp) {
            // talker space -> code-predictor space (src/tts_transformer.cpp:1709-1715)
            Op m = gemv(e->mtp, c.cp_hidden_size, c.hidden_size, hidden, nullptr, c.rms_norm_eps,
                        e->mtp_b ? EPI_BIAS : EPI_STORE, s.cp_x);
            m.in_kind = first.in_kind; m.table = first.table; m.tok_slot = first.tok_slot; m.bias = e->mtp_b;
            m.n_tok = n_tok; m.out_stride = c.cp_hidden_size;
            P.push_back(m);
            first = blank(); first.in_kind = IN_VEC; first.in = s.cp_x;
        } else {
            first.x_out = s.cp_x;
        }
        const int mode = (p == 1) ? ATTN_F32 : ATTN_SOFTMAX_P16;
        for (int l = 0; l < c.cp_layers; ++l)
            emit_layer(e, P, e->cp[l], D, B, l, l == 0 ? &first : nullptr, false, p == 1 ? 0 : p, mode, n_tok);
        // the head reads the LAST token of the pass
        Op hd = gemv(e->lm_head[p - 1], c.cp_vocab_size, c.cp_hidden_size,
                     s.cp_x + (size_t) (n_tok - 1) * c.cp_hidden_size, e->cp_output_norm,
                     c.rms_norm_eps, EPI_STORE, s.cp_logits);
        if (e->ll) { hd.in_ll = s.cp_x_ll + (size_t) (n_tok - 1) * c.cp_hidden_size; hd.in_dist = 1; }
        P.push_back(hd);
        Op sp = blank();
        sp.type = OP_SAMPLE_CP; sp.logits = s.cp_logits; sp.n_vocab = c.cp_vocab_size; sp.book = p;
        sp.uniforms = uniforms; sp.forced = forced; sp.n_books = c.n_codebooks;
        sp.dump = dump ? dump + (size_t) (p - 1) * c.cp_vocab_size : nullptr;
        if (override_sampling) { sp.temperature_override = temperature; sp.top_k_override = top_k; }
        P.push_back(sp);
    }
}

// One single-token code-predictor position (TRTCodePredictor::step / step_with_token,
// src/trt_code_predictor.cpp:384-482 -> run_trt_and_lm_head :303-382): input = an F32 vector already on the device
// (in_vec) or row tok_ptr[0] of an F16 embedding table, projected by mtp_proj when the model has one and `proj` is
// set; 5 layers at `pos` attending to the slot's code-predictor cache [0, pos]; output norm; lm_head[head] -> cp_logits
// (head < 0: lm_head 0 runs only for its prologue, the normalised hidden lands in hid_out).
void emit_cp_step(Engine * e, Slot & s, std::vector<Op> & P, const float * in_vec, const __half * table,
                  const int32_t * tok_ptr, int pos, int head, bool proj, float * hid_out) {
    const qtts_config & c = e->cfg;
    const Dims D{c.cp_hidden_size, c.cp_n_heads, c.cp_n_kv_heads, c.cp_intermediate_size};
    LayerBuf B{s.cp_x, s.cp_qkv, s.cp_attn, s.cp_act, s.cp_kc, s.cp_vc, (size_t) 16 * c.cp_n_kv_heads * kHeadDim,
               s.attn_part, s.attn_cnt, s.cp_x_ll, s.cp_act_ll, s.cp_qkv_ll, s.cp_attn_ll};
    Op first = blank();
    if (table) { first.in_kind = IN_EMBED_ROW; first.table = table; first.tok_ptr = tok_ptr; first.tok_slot = -1; }
    else { first.in_kind = IN_VEC; first.in = in_vec; }
    if (e->mtp && proj) {
        Op m = gemv(e->mtp, c.cp_hidden_size, c.hidden_size, in_vec, nullptr, c.rms_norm_eps, e->mtp_b ? EPI_BIAS : EPI_STORE, s.cp_x);
        m.in_kind = first.in_kind; m.table = first.table; m.tok_ptr = first.tok_ptr; m.tok_slot = -1; m.bias = e->mtp_b;
        m.out_stride = c.cp_hidden_size;
        P.push_back(m);
        first = blank(); first.in_kind = IN_VEC; first.in = s.cp_x;
    } else {
        first.x_out = s.cp_x;
    }
    for (int l = 0; l < c.cp_layers; ++l)
        emit_layer(e, P, e->cp[l], D, B, l, l == 0 ? &first : nullptr, false, pos, ATTN_SOFTMAX_P16, 1);
    Op hd = gemv(e->lm_head[head >= 0 ? head : 0], c.cp_vocab_size, c.cp_hidden_size, s.cp_x, e->cp_output_norm,
                 c.rms_norm_eps, EPI_STORE, s.cp_logits);
    hd.hid_out = hid_out;
    if (e->ll) { hd.in_ll = s.cp_x_ll; hd.in_dist = 1; }
    P.push_back(hd);
}

void emit_frame(Engine * e, Slot & s, std::vector<Op> & P) {
    const qtts_config & c = e->cfg;
    Op s0 = blank();
    s0.type = OP_SAMPLE_CB0; s0.logits = s.logits; s0.n_vocab = c.codec_vocab_size;
    s0.eos_id = c.codec_eos_id; s0.suppress_start = c.codec_vocab_size - 1024;
    s0.dump = s.cb0_dump; s0.n_books = c.n_codebooks;
    P.push_back(s0);
    emit_code_predictor(e, s, P, s.hidden, (e->debug_flags & 1) ? s.cp_dump : nullptr, nullptr, nullptr, false, 0.f, 0);
    Op fe = blank();
    fe.type = OP_FRAME_END; fe.codes_out = s.codes; fe.n_books = c.n_codebooks;
    P.push_back(fe);
    const int mode = (e->numerics == QTTS_NUMERICS_REFERENCE) ? ATTN_FLASH_F16SEQ : ATTN_FAST;
    emit_talker_step(e, s, P, IN_EMBED_SUM, nullptr, true, 0, mode, true, true);
}

static Op copy_op(float * out, const float * a, const float * b, int n) {
    Op o = blank();
    o.type = OP_COPY; o.out = out; o.a = a; o.b = b; o.n = n;
    return o;
}
static Op row_op(float * out, const __half * table, const int32_t * tok_ptr, const float * b, int n) {
    Op o = blank();
    o.type = OP_COPY; o.out = out; o.table = table; o.tok_ptr = tok_ptr; o.b = b; o.n = n;
    return o;
}

__global__ void k_assemble_prefill(const float * __restrict__ tab, const __half * __restrict__ codec_embd,
                                   const int32_t * __restrict__ ids, int n, int n_codec, int has_speaker,
                                   const float * __restrict__ speaker, int H, int plen, int trailing_len,
                                   float * __restrict__ prefill, float * __restrict__ trailing, float * __restrict__ pad);

// build_prefill_graph (src/tts_transformer.cpp:1093-1231) on the device.  Token ids are staged in
// s.tokens: [0..n) the text tokens, then the 3 special text ids, then the codec ids.
bool emit_build_prefill(Engine * e, Slot & s, std::vector<Op> & P, const int32_t * tokens, int n_tokens,
                        bool has_speaker, int language_id) {
    const qtts_config & c = e->cfg;
    const int H = c.hidden_size, E = c.text_embd_dim;
    if (n_tokens < 4) return e->fail("Need at least 4 text tokens for prefill");
    if (n_tokens > e->max_text) return e->fail("text longer than the engine's max_text");
    std::vector<int32_t> ids(tokens, tokens + n_tokens);
    for (int i = 0; i < n_tokens; ++i)
        if (ids[i] < 0 || ids[i] >= c.text_vocab_size) return e->fail("text token id out of range");
    const int sp0 = n_tokens;
    ids.push_back(c.tts_bos_id); ids.push_back(c.tts_eos_id); ids.push_back(c.tts_pad_id);
    std::vector<int32_t> codec;
    if (language_id < 0) codec = {c.codec_nothink_id, c.codec_think_bos_id, c.codec_think_eos_id};
    else codec = {c.codec_think_id, c.codec_think_bos_id, language_id, c.codec_think_eos_id};
    const int n_codec_prefill = (int) codec.size();
    codec.push_back(c.codec_pad_id); codec.push_back(c.codec_bos_id);
    for (int v : codec) if (v < 0 || v >= c.codec_vocab_size) return e->fail("codec prefill id out of range");
    const int cd0 = (int) ids.size();
    ids.insert(ids.end(), codec.begin(), codec.end());
    QTTS_CUDA(cudaMemcpyAsync(s.tokens, ids.data(), ids.size() * sizeof(int32_t), cudaMemcpyHostToDevice, e->stream));
    QTTS_CUDA(cudaStreamSynchronize(e->stream));   // ids is a stack-lifetime host buffer
    if (e->text_tab) {
        const int plen = 3 + (n_codec_prefill + (has_speaker ? 1 : 0) + 1) + 1;
        const int trailing_tokens = n_tokens - 9 > 0 ? n_tokens - 9 : 0;
        k_assemble_prefill<<<plen + trailing_tokens + 2, 256, 0, e->stream>>>(
            e->text_tab, e->codec_embd, s.tokens, n_tokens, n_codec_prefill, has_speaker ? 1 : 0, s.speaker, H,
            plen, trailing_tokens + 1, s.prefill, s.trailing, s.pad);
        QTTS_CUDA(cudaGetLastError());
        s.prefill_len = plen; s.trailing_len = trailing_tokens + 1; s.n_text_tokens = n_tokens;
        return true;            // P stays empty: nothing left for the stage program
    }

    auto text_proj = [&](int tok_index, float * dst) {   // project_text_tokens (:1026-1091)
        Op f1 = gemv(e->fc1, E, E, nullptr, nullptr, c.rms_norm_eps, EPI_BIAS_SILU, s.tp_tmp);
        f1.in_kind = IN_EMBED_ROW; f1.table = e->text_embd; f1.tok_ptr = s.tokens + tok_index; f1.bias = e->fc1_b;
        P.push_back(f1);
        Op f2 = gemv(e->fc2, H, E, s.tp_tmp, nullptr, c.rms_norm_eps, EPI_BIAS, dst);
        f2.bias = e->fc2_b;
        P.push_back(f2);
    };
    float * bos = s.misc, * eos = s.misc + H, * first_text = s.misc + 2 * (size_t) H;
    text_proj(sp0 + 0, bos);
    text_proj(sp0 + 1, eos);
    text_proj(sp0 + 2, s.pad);
    for (int t = 0; t < 3; ++t) text_proj(t, s.prefill + (size_t) t * H);     // role rows
    // codec rows (+ speaker) with the tts_pad / tts_bos overlay (:1165-1193)
    const int codec_input_len = n_codec_prefill + (has_speaker ? 1 : 0) + 2;
    const int overlay_len = codec_input_len - 1;
    for (int t = 0; t < overlay_len; ++t) {
        const float * overlay = (t == overlay_len - 1) ? bos : s.pad;
        float * dst = s.prefill + (size_t) (3 + t) * H;
        if (has_speaker && t == n_codec_prefill) {
            P.push_back(copy_op(dst, overlay, s.speaker, H));
        } else {
            const int ci = (t < n_codec_prefill) ? t : t - (has_speaker ? 1 : 0);
            P.push_back(row_op(dst, e->codec_embd, s.tokens + cd0 + ci, overlay, H));
        }
    }
    // last prefill row: text_proj(tokens[3]) + codec_embd(codec_bos) (:1195-1204)
    text_proj(3, first_text);
    const int prefill_len = 3 + overlay_len + 1;
    P.push_back(row_op(s.prefill + (size_t) (prefill_len - 1) * H, e->codec_embd,
                       s.tokens + cd0 + n_codec_prefill + 1, first_text, H));
    // trailing text rows + tts_eos (:1214-1228)
    const int trailing_tokens = n_tokens - 9 > 0 ? n_tokens - 9 : 0;
    for (int t = 0; t < trailing_tokens; ++t) text_proj(4 + t, s.trailing + (size_t) t * H);
    P.push_back(copy_op(s.trailing + (size_t) trailing_tokens * H, eos, nullptr, H));
    s.prefill_len = prefill_len;
    s.trailing_len = trailing_tokens + 1;
    s.n_text_tokens = n_tokens;
    return true;
}

// ------------------------------------------------------------------------------------------
// Folded text table: project_text_tokens (src/tts_transformer.cpp:1026-1091) is a pure function of the token id
// (text_embd row -> fc1 + bias, SiLU -> fc2 + bias), so it is evaluated ONCE per vocabulary entry at load time, by
// the same two GEMV stages the per-token path runs (bit-identical rows), into an F32 table [text_vocab][H].
// build_prefill_graph (:1093-1231) then is one gather kernel per utterance instead of two dependent GEMV
// stages per text token.  QTTS_FOLD_TEXT=0 keeps the per-token path.
// ------------------------------------------------------------------------------------------
bool fold_text_table(Engine * e) {
    if (const char * v = getenv("QTTS_FOLD_TEXT")) if (atoi(v) == 0) return true;
    const qtts_config & c = e->cfg;
    if (!e->fc1 || !e->fc2 || !e->text_embd || e->slots.empty()) return true;
    const int H = c.hidden_size, E = c.text_embd_dim, V = c.text_vocab_size;
    e->stream_phase = false;            // the table belongs to the weights: it survives qtts_resize_kv
    e->text_tab = e->dalloc<float>((size_t) V * H, false);
    int32_t * ids = e->dalloc<int32_t>(V, false);
    e->stream_phase = true;
    if (!e->text_tab || !ids) { e->text_tab = nullptr; return e->fail("out of device memory for the folded text table"); }
    std::vector<int32_t> iota(V);
    for (int i = 0; i < V; ++i) iota[i] = i;
    if (!e->h2d(ids, iota.data(), (size_t) V * sizeof(int32_t))) return false;
    Slot & s = e->slots[0];
    const int chunk = 2048;
    std::vector<Op> P;
    for (int t0 = 0; t0 < V; t0 += chunk) {
        P.clear();
        for (int t = t0; t < std::min(V, t0 + chunk); ++t) {
            Op f1 = gemv(e->fc1, E, E, nullptr, nullptr, c.rms_norm_eps, EPI_BIAS_SILU, s.tp_tmp);
            f1.in_kind = IN_EMBED_ROW; f1.table = e->text_embd; f1.tok_ptr = ids + t; f1.bias = e->fc1_b;
            P.push_back(f1);
            Op f2 = gemv(e->fc2, H, E, s.tp_tmp, nullptr, c.rms_norm_eps, EPI_BIAS, e->text_tab + (size_t) t * H);
            f2.bias = e->fc2_b;
            P.push_back(f2);
        }
        if (!run_program(e, s, P, 1, false)) { e->text_tab = nullptr; return false; }
    }
    QTTS_CUDA(cudaStreamSynchronize(e->stream));
    return true;
}

// rows of one utterance out of the folded table.  ids: [0, n) text tokens, then tts_bos / tts_eos / tts_pad, then
// the codec prefill ids (n_codec of them), codec_pad, codec_bos.  Row r of the grid:
//   [0, plen)            prefill rows        [plen, plen + trailing_len)  trailing text rows        last: the pad row
__global__ void k_assemble_prefill(const float * __restrict__ tab, const __half * __restrict__ codec_embd,
                                   const int32_t * __restrict__ ids, int n, int n_codec, int has_speaker,
                                   const float * __restrict__ speaker, int H, int plen, int trailing_len,
                                   float * __restrict__ prefill, float * __restrict__ trailing, float * __restrict__ pad) {
    const int r = blockIdx.x;
    const int sp0 = n, cd0 = n + 3;
    const float * bos = tab + (size_t) ids[sp0] * H, * eos = tab + (size_t) ids[sp0 + 1] * H, * padr = tab + (size_t) ids[sp0 + 2] * H;
    const float * a = nullptr, * b = nullptr;
    const __half * hrow = nullptr;
    float * dst;
    if (r < plen) {
        dst = prefill + (size_t) r * H;
        const int overlay_len = plen - 4;
        if (r < 3) a = tab + (size_t) ids[r] * H;                               // role rows
        else if (r < plen - 1) {                                                // codec rows + tts_pad / tts_bos overlay
            const int t = r - 3;
            b = (t == overlay_len - 1) ? bos : padr;
            if (has_speaker && t == n_codec) a = speaker;
            else hrow = codec_embd + (size_t) ids[cd0 + ((t < n_codec) ? t : t - has_speaker)] * H;
        } else {                                                                // text_proj(tokens[3]) + codec_embd(codec_bos)
            hrow = codec_embd + (size_t) ids[cd0 + n_codec + 1] * H;
            b = tab + (size_t) ids[3] * H;
        }
    } else if (r < plen + trailing_len) {
        const int t = r - plen;
        dst = trailing + (size_t) t * H;
        a = (t == trailing_len - 1) ? eos : tab + (size_t) ids[4 + t] * H;
    } else {
        dst = pad; a = padr;
    }
    for (int i = threadIdx.x; i < H; i += blockDim.x) {
        // operand order of the program path: (codec row | overlay) + (overlay | speaker); F32 addition commutes
        const float x = hrow ? __half2float(hrow[i]) : a[i];
        dst[i] = b ? x + b[i] : x;
    }
}

std::vector<WSched> build_sched(Engine * e, const std::vector<Op> & P) {
    std::vector<WSched> S;
    for (const Op & op : P)
        if (op.type == OP_GEMV) S.push_back(WSched{op.W, op.K, op.rows >> 1, e->slot_bytes / (4 * op.K), 0});
    return S;
}

bool run_device_program(Engine * e, Slot & s, const Op * d_ops, const std::vector<Op> * host_ops, int n_ops,
                        const WSched * d_sched, int n_sched, int n_frames, bool check_done) {
    if (n_ops == 0) return true;
    if (e->persistent) {
        QTTS_CUDA(cudaMemsetAsync(&e->bar->count, 0, sizeof(unsigned int), e->stream));
        // stage tags never repeat: a stale word of an earlier launch can not satisfy a poll
        const uint64_t n_tags = (uint64_t) n_ops * (uint64_t) n_frames + 1;
        if ((uint64_t) e->ll_seq + n_tags > 0xfff00000ull) {
            QTTS_CUDA(cudaStreamSynchronize(e->stream));
            for (Slot & t : e->slots) {
                const qtts_config & c = e->cfg;
                const size_t maxH = (size_t) std::max(c.hidden_size, c.cp_hidden_size);
                const size_t qkvw = (size_t) (c.n_heads + 2 * c.n_kv_heads) * kHeadDim;
                const size_t cqkvw = (size_t) (c.cp_n_heads + 2 * c.cp_n_kv_heads) * kHeadDim;
                const std::pair<unsigned long long *, size_t> bufs[] = {
                    {t.x_ll, maxH}, {t.act_ll, (size_t) c.intermediate_size}, {t.cp_x_ll, 2 * maxH},
                    {t.cp_act_ll, 2 * (size_t) c.cp_intermediate_size}, {t.qkv_ll, qkvw},
                    {t.attn_ll, (size_t) c.n_heads * kHeadDim}, {t.cp_qkv_ll, 2 * cqkvw},
                    {t.cp_attn_ll, 2 * (size_t) c.cp_n_heads * kHeadDim}};
                for (const auto & b : bufs) QTTS_CUDA(cudaMemsetAsync(b.first, 0, b.second * 8, e->stream));
            }
            e->ll_seq = 0;
        }
        RunCfg cfg{e->slot_bytes, e->n_slots, e->xs_elems, 1, check_done ? 1 : 0, e->trace_on ? e->trace_cta : e->prof_on - 1, n_sched, d_sched,
                   e->ll_seq, e->prof_on ? e->prof : nullptr, e->trace_on ? e->trace : nullptr, e->trace_cap};
        e->ll_seq += (uint32_t) n_tags;
        StreamState * st = s.st;
        GridBarrier * bar = e->bar;
        void * args[] = {(void *) &d_ops, (void *) &n_ops, (void *) &st, (void *) &bar, (void *) &n_frames,
                         (void *) &cfg};
        // the attribute is per function, not per engine: another engine (other K -> other ring size) may have changed it
        QTTS_CUDA(cudaFuncSetAttribute(qtts_run_program, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) e->smem_bytes));
        QTTS_CUDA(cudaLaunchCooperativeKernel((const void *) qtts_run_program, dim3(e->n_sms), dim3(kThreads),
                                              args, e->smem_bytes, e->stream));
        e->launches += 1;
        return true;
    }
    // debug path: one launch per stage; needs the host copy of the program to find sampling stages
    if (!host_ops) return e->fail("stage-by-stage mode needs the host copy of the program");
    QTTS_CUDA(cudaFuncSetAttribute(qtts_run_program, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) e->smem_bytes));
    for (int f = 0; f < n_frames; ++f) {
        int si = 0;     // schedule entry of the i-th op when it is a GEMV
        for (int i = 0; i < n_ops; ++i) {
            const bool is_gemv = (*host_ops)[i].type == OP_GEMV;
            RunCfg cfg{e->slot_bytes, e->n_slots, e->xs_elems, 0, 0, 0, is_gemv ? 1 : 0, d_sched + si, 0u, nullptr, nullptr, 0};
            if (is_gemv) ++si;
            qtts_run_program<<<e->n_sms, kThreads, e->smem_bytes, e->stream>>>(d_ops + i, 1, s.st, e->bar, 1, cfg);
            e->launches += 1;
            const int type = host_ops ? (*host_ops)[i].type : -1;
            if (check_done && (type == OP_SAMPLE_CB0 || type == OP_FRAME_END || !host_ops)) {
                int done = 0;
                QTTS_CUDA(cudaMemcpyAsync(&done, &s.st->done, sizeof(int), cudaMemcpyDeviceToHost, e->stream));
                QTTS_CUDA(cudaStreamSynchronize(e->stream));
                if (done) return true;
            }
        }
    }
    QTTS_CUDA(cudaGetLastError());
    return true;
}

bool run_program(Engine * e, Slot & s, const std::vector<Op> & P, int n_frames, bool check_done) {
    if (P.empty()) return true;
    if ((int) P.size() > s.scratch_cap) {
        const int cap = (int) P.size() + 256;
        Op * p = e->dalloc<Op>(cap, false);
        if (!p) return e->fail("out of device memory for program");
        s.scratch_prog = p; s.scratch_cap = cap;   // older buffer stays in allocs until destroy
    }
    const std::vector<WSched> S = build_sched(e, P);
    if ((int) S.size() > s.scratch_sched_cap) {
        const int cap = (int) S.size() + 256;
        WSched * p = e->dalloc<WSched>(cap, false);
        if (!p) return e->fail("out of device memory for program schedule");
        s.scratch_sched = p; s.scratch_sched_cap = cap;
    }
    QTTS_CUDA(cudaMemcpyAsync(s.scratch_prog, P.data(), P.size() * sizeof(Op), cudaMemcpyHostToDevice, e->stream));
    if (!S.empty())
        QTTS_CUDA(cudaMemcpyAsync(s.scratch_sched, S.data(), S.size() * sizeof(WSched), cudaMemcpyHostToDevice, e->stream));
    QTTS_CUDA(cudaStreamSynchronize(e->stream));   // P / S may be freed by the caller right after
    return run_device_program(e, s, s.scratch_prog, &P, (int) P.size(), s.scratch_sched, (int) S.size(), n_frames, check_done);
}

}  // namespace qtts
