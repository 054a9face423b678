// qtts_bops.cuh — stage descriptors of the BATCHED frame kernel (qtts_batched.cu): all concurrent
// utterance streams of one GPU advance in lock step, every projection is a real GEMM
// [streams x K] . [K x weight rows] on the tcgen05 tensor cores with the accumulator in TMEM.
//
// Shared-memory operand layouts (UMMA "K-major, no swizzle" canonical form: core matrix = 8 rows x
// 16 bytes, stored as 128 contiguous bytes; LBO = byte distance between core matrices adjacent
// in K, SBO = byte distance between 8-row groups):
//   A (activations, F16, written by the CTA's threads):  LBO = kALbo (144: the 16-byte pad makes
//       the threads' 8/16-byte stores bank-conflict free), SBO = kASbo, one buffer = 128 rows x
//       kKc K-elements;
//   B (weights, BF16 or F16, pre-tiled in HBM at load time so that one 1-D bulk copy lands a
//       ready-to-use operand): LBO = 128, SBO = 16 * (K-extent of the piece).
#pragma once
#include "qtts_ops.cuh"

namespace qtts {

constexpr int kKc         = 128;                 // K-elements per activation chunk (8 MMAs of K=16)
constexpr int kALbo       = 144;
constexpr int kASbo       = (kKc / 8) * kALbo;   // 2304
constexpr int kABufBytes  = 16 * kASbo;          // 128 rows -> 36864 B
constexpr int kBSlots     = 6;                   // weight ring slots
constexpr int kBSlotBytes = 24576;
constexpr int kBMaxRows   = 512;                 // activation rows per GEMM (streams x tokens)
constexpr int kBMaxGroup  = 256;                 // streams advanced in lock step by one launch
constexpr int kBMaxParts  = 256;                 // sum-of-squares partials per row (= out rows / 16)

enum BOpType : int32_t {
    BOP_GEMM    = 0,   // out[r][n] = epi( sum_k W[n][k] * f16(prologue(in[r][k])) )
    BOP_ATTN_CP = 1,   // code-predictor attention, one warp per (stream, KV group), <= 16 keys
    BOP_ATTN_TK = 2,   // Talker decode attention, one CTA pass per (stream, KV group)
    BOP_SAMPLE  = 3,   // per-stream sampling + embedding of the next input (+ frame end)
    BOP_NORM    = 4,   // RMSNorm of F32 rows -> F16 rows in the canonical chunk layout (the next GEMM's A operand,
                       // fetched by bulk copies); one row per CTA pass
};

enum BInKind : int32_t { BIN_F32 = 0, BIN_F16 = 1 };

enum BFlags : int32_t {
    BF_INC_NPAST = 1,      // GEMM: CTA 0 advances n_past of every live stream at the end of the stage
    BF_SSQ_OUT   = 2,      // GEMM epilogue writes per-row sum-of-squares partials of its output
    BF_ROWS_SIGNAL = 4,    // BOP_NORM: every finished row bumps the group's row counter; NO grid barrier after the stage
    BF_ROWS_WAIT   = 8,    // GEMM fed by such a stage: the A-operand producer waits for the row counter instead
    BF_ASYNC_OUT   = 16,   // the stage's output (F16 canonical chunks) is fetched by BULK COPIES (async proxy) in the next
                           // stage: the grid barrier behind it needs fence.proxy.async.global on the writer side; stages
                           // whose output is read with ordinary loads skip that fence
};

struct BOp {
    int32_t type, flags;
    int32_t n_streams;           // B
    int32_t rows_act;            // activation rows = n_tok * B (row = t * B + b)
    // ---- BOP_GEMM ----
    const unsigned char * Wt;    // tiled weights (see qtts_batched_host.cu: tile_weights)
    int32_t rows, K, n_tile, n_tiles, cta_off, kpp, n_pieces;
    int32_t in_kind, epi;
    const void * in;             // f32 or f16 rows
    int32_t in_row_stride, in_tok_stride;      // elements
    const float * norm_w;        // RMSNorm weight (BIN_F32 only; nullptr: no norm)
    const float * ssq_in;        // [rows_act][n_parts_in] partial sums of squares of the input rows
    int32_t n_parts_in;
    float eps;
    float * out32;
    __half * out16;              // EPI_SWIGLU output (F16 rows, next GEMM's A operand)
    const float * res;           // EPI_ADD (same strides as out32)
    const float * bias;
    int32_t out_row_stride, out_tok_stride;
    float * ssq_out;             // [rows_act][rows / 16]
    float * hid_out;             // normalised, unrounded input rows [rows_act][K] (Talker hidden)
    // split-K (F16-input stages with fewer tiles than half the SMs): the weight matrix is tiled as
    // ksplit "virtual tiles" per row block (virtual tile v = row block v / ksplit, K-half v % ksplit;
    // K above is the per-half extent).  Half 1 hands its partial accumulator to half 0 through
    // flag-carrying words (f32 | stage tag << 32); half 0 adds it in fixed order and runs the epilogue.
    int32_t ksplit;
    // split-M (stages with at least two M tiles and fewer N tiles than half the SMs): virtual tile v =
    // (N tile v / msplit, M-tile range v % msplit); both CTAs stream the same weight tile, each its own rows
    int32_t msplit;
    unsigned long long * part_ll;    // [row blocks][n_tile / 16][rows_act padded to 32][16]
    // deferred combine (ksplit with part32 set): half 0 runs the epilogue on its own partial sum, half 1 stores its
    // partial to part32 (layout of out32); the BOP_NORM stage that follows adds it (x = x + part, written back) and
    // computes the row's sum of squares itself - no CTA ever waits for its partner
    float * part32;
    // ---- attention ----
    const float * qkv; int32_t qkvw;
    const float * q_norm_w, * k_norm_w, * rope;
    int32_t n_heads, n_kv_heads, n_tok, pos, pair_f32;
    float kq_scale;
    __half * kc, * vc;           // layer base of stream 0
    int64_t kv_stream_stride;    // elements between streams
    __half * attn16; int32_t attn_w;           // output rows [rows_act][attn_w]
    // ---- BOP_SAMPLE ----
    int32_t kind;                // 0: codebook 0 (Talker logits), 1: code predictor codebook `book`
    int32_t book, n_books;
    const float * logits; int32_t logits_stride, n_vocab;
    int32_t eos_id, suppress_start;
    float * x_out; int32_t xw;   // embedding rows written for the next pass [rows][xw]
    float * ssq_e;               // [rows] complete sums of squares of those rows
    const float * hidden; int32_t hidden_stride;          // kind 0: token-0 row of the code predictor
    const __half * table;        // kind 0: codec_embd; kind 1, book < last: code_pred embd[book-1]
    const EmbedSumArgs * es;     // last book: tables of the next-Talker-input sum
    const float * trailing; int32_t trailing_stride; const float * pad; int32_t pad_stride;
    int32_t * codes_out; int32_t codes_stride;            // [stream][max_ctx][n_books]
    float * dump; int32_t dump_stride;                     // optional copy of the processed logits
};

// prefetcher schedule: the GEMM stages of one frame in execution order
struct BSched {
    const unsigned char * W;
    int32_t tile_bytes, piece_bytes, n_pieces, n_tiles, cta_off, msplit;     // n_tiles counts virtual tiles; weight tile = virtual tile / msplit
};
static_assert(sizeof(BSched) == 32, "BSched is read as two int4");

struct BRun {
    const BOp * ops; int32_t n_ops;
    const BSched * sched; int32_t n_sched;
    StreamState * st;            // [B]
    const int32_t * active;      // [B]
    GridBarrier * bar;
    int32_t n_streams, n_frames, check_done;
    uint32_t idesc_base;         // instruction descriptor without N (formats, M = 128)
    uint32_t tag_base;           // stage tag = tag_base + 1 + frame * n_ops + op index (never repeated)
    unsigned long long * prof;   // optional stage timers
    int32_t dbg;                 // bring-up switches (env QTTS_BDBG): 1 skip MMA issue, 2 skip TMEM loads
    uint32_t * trace;            // optional stage timeline of CTA trace_cta: 16 words per stage, store-only stamps written
    int32_t trace_cta, trace_cap;//   by the threads that own each step (see tools/trace_batched.py for the slots)
    int32_t cluster;             // 1: launched as thread-block clusters of 2 (CTA pairs share the A operand by multicast)
    int32_t a_bufs;              // full-size activation buffers: 2 (weight ring of kBSlots slots) or 4 (ring of kBSlots - 3): see bring_slots
};

}  // namespace qtts
