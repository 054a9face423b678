// qtts_ops.cuh — op descriptors shared by the host-side program builder and the device stages.
//
// The frame is the unit of GPU work: the host builds, once per stream, a linear "program" of
// stage descriptors (Op) for one codec frame; a persistent cooperative kernel walks it with a
// grid-wide barrier between stages (qtts_kernels.cu).  The same device code runs with one
// launch per stage when persistent mode is switched off (debug path).
#pragma once
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>

namespace qtts {

// Projection weights are kept in the GGUF's own type, F16 (scripts/convert_tts_to_gguf.py:289-297), exactly the
// values the reference's ggml_mul_mat multiplies; the activations are rounded to F16 like ggml does
// (ggml/src/ggml-cpu/ggml-cpu.c:1241-1243), and sm_100's mixed-precision FMA (fma.rn.f32.f16 -> SASS FHFMA: F16 x F16
// + F32 with one rounding) accumulates in FP32 without a conversion instruction per operand.
using wt_t = __half;

constexpr int kThreads   = 512;   // threads per CTA (16 warps)
constexpr int kWarps     = kThreads / 32;
constexpr int kHeadDim   = 128;   // every Qwen3-TTS config uses 128 (src/tts_transformer.h:33,47)
constexpr int kMaxVocab  = 4096;  // codec vocab 3072, code-pred vocab 2048
constexpr int kMaxBooks  = 32;
constexpr int kNumSlots  = 10;    // shared-memory weight ring: slots of slot_bytes each
constexpr int kMaxSplits = 16;    // ATTN_FAST: key ranges per KV group
constexpr int kSplitKeys = 128;   // ATTN_FAST: target keys per split
constexpr int kPartStride = 132;  // m, l, pad, pad, o[128]
constexpr int kMaxScores = 4608;  // attention scores kept in shared memory (>= max_ctx)

enum OpType : int32_t {
    OP_GEMV       = 0,   // skinny projection: y = W . f16round(x)  (+ prologue / epilogue)
    OP_ATTN       = 1,   // QK-norm + NEOX RoPE + KV append + attention over the cache
    OP_SAMPLE_CB0 = 2,   // codebook-0 logit post-processing + argmax / top-k sample
    OP_SAMPLE_CP  = 3,   // code-predictor argmax / top-k sample of one codebook
    OP_FRAME_END  = 4,   // append the 16 codes of the frame, frame++ , max_len check
    OP_COPY       = 5,   // small f32 vector copy / add (prefill embedding assembly)
};

enum InKind : int32_t {
    IN_VEC       = 0,    // f32 vector(s) in global memory: in + t * in_stride
    IN_EMBED_ROW = 1,    // one row of an F16 embedding table; row id = tok_ptr (device memory) or,
                         // when tok_slot >= 0, the CTA-local feed_codes[tok_slot]
    IN_EMBED_SUM = 2,    // next Talker input: sum of the 16 code embeddings + text row
    IN_CP_FIRST  = 3,    // code-predictor positions 0 and 1 together (n_tok = 2):
                         // token 0 = f32 vector `in` (Talker hidden), token 1 = table[feed_codes[0]]
};

enum Epilogue : int32_t {
    EPI_STORE     = 0,   // out[r] = acc
    EPI_ADD       = 1,   // out[r] = acc + res[r]          (residual)
    EPI_SWIGLU    = 2,   // rows interleaved gate/up: out[p] = silu(acc[2p]) * acc[2p+1]
    EPI_BIAS      = 3,   // out[r] = acc + bias[r]
    EPI_BIAS_SILU = 4,   // out[r] = silu(acc + bias[r])
};

enum AttnMode : int32_t {
    ATTN_FLASH_F16SEQ = 0,  // ggml-cpu flash_attn_ext: F16 Q, sequential F16 V accumulator
    ATTN_SOFTMAX_P16  = 1,  // ggml-cpu mul_mat/soft_max/mul_mat: F16 Q, F32 softmax, F16 P
    ATTN_F32          = 2,  // code-predictor 2-token prefill: everything F32, unrounded K/V
    ATTN_FAST         = 3,  // F16 Q/K/V, FP32 softmax and FP32 P.V accumulation, parallel
};

// Per-stream device state (one per slot).
struct StreamState {
    int32_t n_past;          // Talker KV entries in use (= position of the next token)
    int32_t frame;           // frames emitted so far
    int32_t done;            // 1 once EOS was sampled or max_len reached
    int32_t eos;             // 1 if the stream ended by EOS
    int32_t max_len;
    int32_t n_text_tokens;   // drives the EOS boost (src/tts_transformer.cpp:2439)
    int32_t trailing_len;
    int32_t top_k;
    float   rep_pen;
    float   temperature;
    uint64_t seed;
    int32_t ignore_eos;      // benchmark switch: EOS can never be chosen (fixed-length utterances)
    int32_t forced_frames;   // teacher forcing (parity hook, qtts_stream_force): frames [0, forced_frames) feed back
    const int32_t * forced;  //   forced[frame][n_books] instead of the engine's own choice (which is still what is emitted)
    const float * uniforms;  // host-supplied uniforms [frame][n_books] (TRTCodePredictor::run_sampling_loop's rand_vals,
                             //   src/trt_code_predictor.cpp:602-726, extended to codebook 0); nullptr: Philox
    int32_t uniform_frames;  // rows of `uniforms`; later frames fall back to Philox
    int32_t sample_mode;     // 0: codebook-0 top-k threshold = exact k-th largest; 1: the reference's std::nth_element pick
    int32_t cur_codes[kMaxBooks];   // codes of the frame being produced (cb0..cb15): the engine's own choice, emitted
    int32_t feed_codes[kMaxBooks];  // codes fed back (code-predictor inputs, next Talker input, repetition set): forced or own
    uint32_t seen[kMaxVocab / 32];  // bitmap of distinct cb0 tokens emitted so far
};

struct EmbedSumArgs {
    const __half * tables[kMaxBooks];   // [0] = talker codec_embd, [c] = code_pred_embd[c-1]
    int32_t n_books;
    const float * trailing;             // [trailing_len][H]
    const float * pad;                  // [H]
};

struct Op {
    int32_t type;
    int32_t flags;           // F_* below
    // ---- OP_GEMV ----
    const wt_t * W;          // [rows][K] row-major, F16
    int32_t rows, K;
    int32_t in_kind, epi;
    int32_t n_tok;           // 1, or 2 for the merged code-predictor positions 0/1 (also OP_ATTN)
    int32_t tok_slot;        // IN_EMBED_ROW: index into the local feed_codes (-1: use tok_ptr)
    int32_t in_stride, out_stride;   // element strides between the n_tok vectors of in/res and out
    const float * in;        // IN_VEC
    const __half * table;    // IN_EMBED_ROW
    const int32_t * tok_ptr; // IN_EMBED_ROW: row id
    const EmbedSumArgs * es; // IN_EMBED_SUM
    const float * norm_w;    // RMSNorm weight (nullptr: no norm)
    float * x_out;           // CTA 0 stores the raw input vector here (residual stream init)
    float * hid_out;         // CTA 0 stores the normalised (unrounded) vector here
    const float * bias;
    const float * res;
    float * out;
    float eps;
    // flag-carrying copies ("LL": one 64-bit word = f32 value | 32-bit stage tag << 32) that let the
    // consumer stage poll the data itself instead of waiting at a grid barrier (see F_LL_OUT)
    unsigned long long * out_ll;         // written next to `out` with this stage's tag
    const unsigned long long * in_ll;    // IN_VEC (OP_ATTN: qkv) read from here, tag = this stage's tag - in_dist
    const unsigned long long * res_ll;   // EPI_ADD residual, tag = this stage's tag - res_dist
    unsigned long long * x_out_ll;       // flag-carrying copy of x_out (this stage's tag)
    int32_t in_dist, res_dist;
    // ---- OP_ATTN ----
    const float * qkv;       // [(n_heads + 2 n_kv_heads) * 128] raw projections
    const float * q_norm_w, * k_norm_w;
    __half * kcache, * vcache;   // [n_ctx][n_kv_heads*128] of this layer / slot
    float * kv0_f32;             // ATTN_F32 only: unrounded K,V of position 0 [2][n_kv_heads*128]
    const float * rope;          // [max_pos][64][2] cos,sin
    int32_t n_heads, n_kv_heads;
    int32_t pos_off;             // position = (use_n_past ? state.n_past : 0) + pos_off
    int32_t use_n_past;
    int32_t mode;
    int32_t ref_threads;         // thread count of the reference's split-KV path (>=512 keys)
    float   kq_scale;
    int32_t qkv_stride;          // element stride between the n_tok qkv vectors
    float * attn_part;           // ATTN_FAST split-KV partials [n_heads][kMaxSplits][kPartStride]
    unsigned int * attn_cnt;     // ATTN_FAST arrival counters [n_kv_heads]
    // ---- sampling ----
    const float * logits; int32_t n_vocab;
    int32_t book;                // OP_SAMPLE_CP: index 1..15 of the code written
    const float * uniforms;      // optional [.] host-supplied uniforms (nullptr: Philox)
    const int32_t * forced;      // optional teacher-forcing codes for the code predictor
    float * dump;                // optional copy of the logits (parity hooks)
    int32_t eos_id, suppress_start;
    float temperature_override;  // < -1e29: use stream state
    int32_t top_k_override;
    // ---- OP_FRAME_END ----
    int32_t * codes_out;         // [max_len][n_books]
    int32_t n_books;
    // ---- OP_COPY ----  out[i] = a[i] (+ b[i])
    const float * a, * b; int32_t n;
};

enum OpFlags : int32_t {
    F_INC_NPAST = 1,     // CTA 0 increments state.n_past when the stage completes
    F_LL_OUT    = 2,     // persistent mode: no grid barrier after this stage; every consumer of its
                         // output reads the flag-carrying copy (out_ll) and polls for this stage's tag
};

// Host-built list of the program's GEMV stages in execution order (one frame); the prefetcher
// walks it cyclically, so that issuing a chunk costs a handful of integer instructions.
struct WSched {
    const wt_t * W;
    int32_t K, pairs, ppc, pad;     // ppc = row pairs per ring chunk
    int32_t pad2[2];                // 32-byte entries: read as two 16-byte loads
};

static_assert(sizeof(WSched) == 32, "WSched is read as two int4");

struct GridBarrier {
    unsigned int count;   // monotonically increasing arrival counter
    unsigned int pad[31];
};

}  // namespace qtts
