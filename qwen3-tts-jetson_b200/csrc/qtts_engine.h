// qtts_engine.h — host-side engine state behind the C ABI (include/qtts_b200.h).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <map>
#include <string>
#include <vector>
#include "../../include/qtts_b200.h"
#include "qtts_ops.cuh"
#include "qtts_bops.cuh"

namespace qtts {

struct DevTensor {
    void * raw = nullptr;          // as uploaded: F16 (2-D) or F32 (1-D)
    int dtype = 0;
    int64_t ne[4] = {1, 1, 1, 1};  // ggml order (ne[0] innermost)
    int n_dims = 0;
    size_t n_elem = 0;
};

struct LayerW {
    wt_t * wqkv = nullptr, * wo = nullptr, * wgu = nullptr, * wdown = nullptr;
    float * attn_norm = nullptr, * ffn_norm = nullptr, * q_norm = nullptr, * k_norm = nullptr;
};

struct Dims { int H, nh, nkv, ffn; };

// projection matrix re-tiled for the batched (tcgen05) kernel, see qtts_batched_host.cu
// K = K-extent of one (virtual) tile, n_tiles = number of virtual tiles = row blocks x ksplit
struct TiledW { unsigned char * p = nullptr; int rows = 0, K = 0, n_tile = 0, n_tiles = 0, kpp = 0, n_pieces = 0, ksplit = 1; };
struct TiledLayer { TiledW qkv, o, gu, down; };

// per-stream buffers shared by the batch-1 and the batched kernels: one contiguous array over all
// slots (slot i = base + i * stride), so the batched kernel indexes streams directly
struct BatchBufs {
    StreamState * st = nullptr;
    float * logits = nullptr, * hidden = nullptr, * trailing = nullptr, * pad = nullptr;
    __half * kc = nullptr, * vc = nullptr;
    int32_t * codes = nullptr;
    size_t kv_slab = 0, trailing_stride = 0, codes_stride = 0;
};

// one lock-step group (<= e->bgroup = 256 streams) advanced by one batched launch
struct BGroup {
    int s0 = 0, n = 0;
    float * x = nullptr, * qkv = nullptr, * ssq_e = nullptr, * ssq_p = nullptr, * emb = nullptr;
    __half * attn16 = nullptr, * act16 = nullptr;
    float * cp_x = nullptr, * cp_qkv = nullptr, * cp_ssq_e = nullptr, * cp_ssq_p = nullptr, * cp_logits = nullptr;
    __half * cp_attn16 = nullptr, * cp_act16 = nullptr, * cp_kc = nullptr, * cp_vc = nullptr;
    float * cb0_dump = nullptr, * cp_dump = nullptr;
    int32_t * active = nullptr;
    unsigned long long * part_ll = nullptr;   // split-K partial accumulators (flag-carrying)
    float * part32 = nullptr;                 // split-K, deferred combine: second half's partial sums (layout of x)
    __half * xn16 = nullptr;                  // BOP_NORM output (canonical F16 rows)
    size_t part_words = 0;
    BOp * ops = nullptr; int n_ops = 0;
    BSched * sched = nullptr; int n_sched = 0;
    std::vector<BOp> host_ops;
    // batched prefill (run_batched_prefill): the Talker half of the frame program over its own residual rows, so
    // streams of the group that are mid-utterance keep their pending input / logits / hidden
    float * pf_x = nullptr, * pf_ssq_e = nullptr, * pf_logits = nullptr, * pf_hidden = nullptr;
    const float ** pf_rows = nullptr;         // [n] prefill embedding rows of the streams being prefilled (else null)
    int32_t * pf_plen = nullptr, * pf_active = nullptr;
    BOp * pf_ops = nullptr; int pf_n_ops = 0;
    BSched * pf_sched = nullptr; int pf_n_sched = 0;
};

// per-slot device buffers
struct Slot {
    StreamState * st = nullptr;
    float * x = nullptr, * qkv = nullptr, * attn = nullptr, * act = nullptr;
    float * logits = nullptr, * hidden = nullptr, * in_embd = nullptr;
    float * cp_x = nullptr, * cp_qkv = nullptr, * cp_attn = nullptr, * cp_act = nullptr, * cp_logits = nullptr;
    // flag-carrying copies of the residual stream and the SwiGLU output (Op::out_ll)
    unsigned long long * x_ll = nullptr, * act_ll = nullptr, * cp_x_ll = nullptr, * cp_act_ll = nullptr;
    unsigned long long * qkv_ll = nullptr, * attn_ll = nullptr, * cp_qkv_ll = nullptr, * cp_attn_ll = nullptr;
    float * cp_hidden_in = nullptr;        // hidden handed to predict_codes by the API
    float * cp_dump = nullptr;             // [15][cp_vocab]
    float * cb0_dump = nullptr;            // [V]
    float * uniforms = nullptr;            // [max(16, max_len)]
    int32_t * forced = nullptr;            // [15]
    int32_t * force_codes = nullptr;       // [max_ctx][n_books] teacher-forcing codes (qtts_stream_force), allocated on first use
    float * force_u = nullptr;             // [max_ctx][n_books] host-supplied uniforms
    float * prefill_hidden = nullptr;      // [64][H] normalised hidden of every prefill row (forward_prefill's `output`), on first use
    __half * kc = nullptr, * vc = nullptr;           // [n_layers][max_ctx][kvw]
    __half * cp_kc = nullptr, * cp_vc = nullptr;     // [cp_layers][16][cp_kvw]
    float * attn_part = nullptr;                     // ATTN_FAST partials [n_heads][kMaxSplits][kPartStride]
    unsigned int * attn_cnt = nullptr;               // [n_kv_heads]
    float * prefill = nullptr;             // [max_prefill][H]
    float * trailing = nullptr;            // [max_text+1][H]
    float * pad = nullptr;                 // [H]
    float * tp_tmp = nullptr;              // [E]
    float * misc = nullptr;                // [8][H] scratch rows for the embedding assembly
    float * speaker = nullptr;             // [H]
    int32_t * tokens = nullptr;            // [max_text]
    int32_t * codes = nullptr;             // [max_ctx][n_books]
    EmbedSumArgs * es = nullptr;
    Op * frame_prog = nullptr; int frame_prog_n = 0;
    WSched * frame_sched = nullptr; int frame_sched_n = 0;
    WSched * scratch_sched = nullptr; int scratch_sched_cap = 0;
    std::vector<Op> frame_host;
    Op * scratch_prog = nullptr; int scratch_cap = 0;
    // host mirrors
    int32_t prefill_len = 0, trailing_len = 0, n_text_tokens = 0;
    bool active = false;
    bool pending = false;                  // prepared (embeddings + state) but the Talker prefill has not run yet
};

struct Engine {
    qtts_config cfg{};
    int device = 0, max_streams = 0, max_ctx = 0, max_text = 0;
    int n_sms = 0;
    size_t smem_bytes = 0;
    int slot_bytes = 0, n_slots = 0, xs_elems = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    std::string err;
    std::map<std::string, DevTensor> tensors;
    std::vector<void *> allocs;          // weights, tiled weights, folded text table: live as long as the engine
    std::vector<void *> stream_allocs;   // everything sized by max_streams / max_ctx / max_text: rebuilt by qtts_resize_kv
    bool stream_phase = false;           // which of the two lists dalloc appends to
    bool finalized = false;
    int debug_flags = 0;                 // qtts_set_debug
    int numerics = QTTS_NUMERICS_REFERENCE;
    int ref_threads = 4;          // ggml's default thread count (ggml/include/ggml.h:232)
    int persistent = 1;
    int ll = 1;                    // flag-carrying activations between GEMV stages instead of grid barriers (env QTTS_LL=0: off)
    uint32_t ll_seq = 0;           // stage tags handed out so far (monotonic across launches)
    uint32_t bll_seq = 0;          // same for the batched kernel's split-K hand-off
    int bgroup = 256;              // streams per lock-step group of the batched kernel (env QTTS_BGROUP, <= kBMaxGroup)
    int bcluster = -1;             // batched kernel launched as clusters of 2 CTAs (A operand multicast): -1 not probed yet, 0 off, 1 on (env QTTS_BCLUSTER=0 disables)
    int ignore_eos = 0;
    int64_t n_weights = 0, n_inexact = 0, weight_bytes = 0, kv_bytes = 0;
    int64_t launches = 0;
    double load_ms = 0.0; int64_t load_bytes = 0;       // qtts_load_gguf: wall time and bytes of the file -> HBM stage
    double last_loop_ms = 0.0;
    unsigned long long * d_inexact = nullptr;
    GridBarrier * bar = nullptr;
    unsigned long long * prof = nullptr; int prof_on = 0;   // [8][8] stage cycle counters (instrumentation)
    // qtts_step_frames staging: device gather buffer + pinned host mirror (codes [slot][n_frames][n_books] + meta)
    int32_t * step_dev = nullptr, * step_host = nullptr; size_t step_words = 0;
    long long * chunk_dev = nullptr; size_t chunk_words = 0;       // qtts_codes_chunk staging (host destination)
    uint32_t * trace = nullptr; int trace_on = 0, trace_cta = 0, trace_cap = 0;   // stage timeline of one CTA (qtts_trace)
    float * rope = nullptr;
    // fused weights
    std::vector<LayerW> talker, cp;
    wt_t * codec_head = nullptr, * fc1 = nullptr, * fc2 = nullptr, * mtp = nullptr;
    std::vector<wt_t *> lm_head;
    float * output_norm = nullptr, * cp_output_norm = nullptr, * fc1_b = nullptr, * fc2_b = nullptr, * mtp_b = nullptr;
    const __half * codec_embd = nullptr, * text_embd = nullptr;
    float * text_tab = nullptr;    // folded text table [text_vocab][H] (fold_text_table), nullptr: per-token projection
    std::vector<const __half *> cp_embd;
    std::vector<Slot> slots;
    // batched (tcgen05) path
    int bfmt = 0;                 // B-operand (weight) format: 0 = F16 (1 = BF16 only for the bring-up test)
    int batched = 1;              // qtts_step_frames uses the batched kernel
    bool batched_ready = false;
    int64_t weight_bytes_batched = 0;
    BatchBufs bb;
    std::vector<BGroup> groups;
    std::vector<TiledLayer> t_talker, t_cp;
    std::vector<TiledW> t_lm_head;
    TiledW t_codec_head, t_mtp;

    bool fail(const std::string & m) { err = m; return false; }
    template <typename T> T * dalloc(size_t n, bool zero = true) {
        void * p = nullptr;
        if (cudaMalloc(&p, n * sizeof(T)) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        if (zero) {
            // cudaMemset runs on the legacy stream and is asynchronous to the host; the engine's stream is
            // non-blocking, so without this wait a later copy / kernel on it could be overtaken by the memset
            cudaMemset(p, 0, n * sizeof(T));
            cudaStreamSynchronize(0);
        }
        (stream_phase ? stream_allocs : allocs).push_back(p);
        return (T *) p;
    }
    // host -> device on the engine's stream, complete on return (a synchronous cudaMemcpy from pageable memory may
    // return before the DMA has landed, and is not ordered with the non-blocking stream the kernels run on)
    bool h2d(void * dst, const void * src, size_t bytes) {
        if (cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, stream) != cudaSuccess ||
            cudaStreamSynchronize(stream) != cudaSuccess) {
            err = std::string("host to device copy failed: ") + cudaGetErrorString(cudaGetLastError());
            return false;
        }
        return true;
    }
};

// qtts_weights.cu
bool engine_upload(Engine * e, const char * name, int dtype, const int64_t * dims, int n_dims, const void * host);
bool engine_load_gguf(Engine * e, const char * path);
bool engine_finalize(Engine * e);
bool engine_build_streams(Engine * e);      // per-stream buffers, frame programs, lock-step groups (after the weights)
bool engine_resize(Engine * e, int n_ctx, int max_text);

// qtts_kernels.cu
size_t slot_bytes_for(int k_max);
size_t xs_elems_for(int k_max, int k2_max);
int n_slots_for(int k_max, int k2_max, size_t smem_limit);
size_t smem_required(int k_max, int k2_max, int n_slots);

// qtts_program.cu
void emit_talker_step(Engine * e, Slot & s, std::vector<Op> & P, int in_kind, const float * in_vec,
                      bool use_n_past, int pos_abs, int attn_mode, bool with_head, bool inc_npast);
void emit_code_predictor(Engine * e, Slot & s, std::vector<Op> & P, const float * hidden,
                         float * dump, const int32_t * forced, const float * uniforms,
                         bool override_sampling, float temperature, int top_k);
void emit_frame(Engine * e, Slot & s, std::vector<Op> & P);
void emit_cp_step(Engine * e, Slot & s, std::vector<Op> & P, const float * in_vec, const __half * table,
                  const int32_t * tok_ptr, int pos, int head, bool proj, float * hid_out);
bool emit_build_prefill(Engine * e, Slot & s, std::vector<Op> & P, const int32_t * tokens, int n_tokens,
                        bool has_speaker, int language_id);
bool run_program(Engine * e, Slot & s, const std::vector<Op> & P, int n_frames, bool check_done);
bool run_device_program(Engine * e, Slot & s, const Op * d_ops, const std::vector<Op> * host_ops, int n_ops,
                        const WSched * d_sched, int n_sched, int n_frames, bool check_done);
std::vector<WSched> build_sched(Engine * e, const std::vector<Op> & P);
bool fold_text_table(Engine * e);

// qtts_batched_host.cu
bool tile_matrix(Engine * e, const wt_t * src, int rows, int K, TiledW & T, bool allow_split = false);
bool batched_init(Engine * e);              // tiles the weights (once)
bool batched_build_groups(Engine * e);      // lock-step groups over the current slots
uint32_t batched_idesc(const Engine * e);
bool launch_batched(Engine * e, const BOp * d_ops, int n_ops, const BSched * d_sched, int n_sched, StreamState * st,
                    const int32_t * d_active, int n_streams, int n_frames, bool check_done);
bool run_batched_group(Engine * e, BGroup & G, int n_frames);
bool run_batched_prefill(Engine * e, BGroup & G, const std::vector<const float *> & rows, const std::vector<int32_t> & plen);

#define QTTS_CUDA(call) do { cudaError_t _e = (call); if (_e != cudaSuccess) { \
    e->err = std::string(#call) + ": " + cudaGetErrorString(_e); return false; } } while (0)

}  // namespace qtts
