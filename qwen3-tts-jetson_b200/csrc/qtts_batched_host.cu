// qtts_batched_host.cu — host side of the batched frame kernel: load-time re-tiling of the
// projection weights into the UMMA operand layout, the per-group frame program, the launcher.
#include "qtts_engine.h"
#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace qtts {

extern "C" __global__ void qtts_run_batched(const __grid_constant__ BRun run_param);
extern "C" __global__ void qtts_run_batched_ext(const __grid_constant__ BRun run_param);

// src: F16 [rows][K] row-major (the batch-1 kernel's fused layouts).  dst: for tile t (n_tile
// rows) and piece p (kpp K-elements, the last one shorter), the piece is one contiguous block in
// the K-major no-swizzle canonical layout: (row r, 8-element chunk c) at
//   (r / 8) * (kp * 16) + c * 128 + (r % 8) * 16,  kp = K-extent of the piece.
// With ksplit > 1 a row block is stored as ksplit virtual tiles, one per K-range of Kh = K / ksplit
// columns (virtual tile index = row block * ksplit + K-half); kpp / pieces refer to the virtual tile.
__global__ void k_tile_weights(const wt_t * __restrict__ src, unsigned char * __restrict__ dst,
                               int rows, int K, int n_tile, int kpp, int to_bf16, int ksplit) {
    const size_t n_units = (size_t) rows * (K / 8);
    const int Kh = K / ksplit;
    for (size_t u = (size_t) blockIdx.x * blockDim.x + threadIdx.x; u < n_units; u += (size_t) gridDim.x * blockDim.x) {
        const int r = (int) (u / (K / 8)), c = (int) (u % (K / 8));
        const int t = r / n_tile, rr = r % n_tile;
        const int kfull = c * 8, h = kfull / Kh, k = kfull - h * Kh;
        const int p = k / kpp, kin = k - p * kpp;
        const int kp = min(kpp, Kh - p * kpp);
        const size_t off = (size_t) (t * ksplit + h) * n_tile * Kh * 2 + (size_t) p * n_tile * kpp * 2 +
                           (size_t) (rr / 8) * (kp * 16) + (size_t) (kin / 8) * 128 + (size_t) (rr % 8) * 16;
        uint4 v = *reinterpret_cast<const uint4 *>(src + (size_t) r * K + kfull);
        if (to_bf16) {      // bring-up only (qtts_debug_bgemm with a BF16 B operand)
            const __half * h = reinterpret_cast<const __half *>(&v);
            __nv_bfloat16 b[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) b[i] = __float2bfloat16_rn(__half2float(h[i]));
            v = *reinterpret_cast<const uint4 *>(b);
        }
        *reinterpret_cast<uint4 *>(dst + off) = v;
    }
}

static int pick_n_tile(int rows, int grid) {
    const char * env = getenv("QTTS_NTILE");
    if (env) { const int v = atoi(env); if (v >= 16 && v <= 96 && v % 16 == 0 && rows % v == 0) return v; }
    int best = 16;
    for (int nt = 16; nt <= 96; nt += 16) {
        if (rows % nt) continue;
        best = nt;
        if (rows / nt <= grid) break;
    }
    return best;
}

bool tile_matrix(Engine * e, const wt_t * src, int rows, int K, TiledW & T, bool allow_split) {
    if (rows % 16 || K % kKc) return e->fail("batched path: projection shape must be a multiple of 16 x 128");
    T.rows = rows;
    T.n_tile = pick_n_tile(rows, e->n_sms);
    // split-K: a narrow projection (O, down) has fewer row blocks than half the SMs; two CTAs per row
    // block, each streaming half of K, double the bytes in flight of a latency-bound stage
    const char * env = getenv("QTTS_KSPLIT");
    // QTTS_KSPLIT=1: half 0 waits for half 1's partial (flag-carrying words) - measured: the hand-off costs what the
    // shorter loop saves.  QTTS_KSPLIT=2: deferred combine, the following norm stage adds the partial (BEmit::defer).
    // Only for groups of one M tile: wider groups use split-M, which needs unsplit weights.
    const int mode = env ? atoi(env) : 2;       // default: deferred combine (+9% at 64 streams, +12% at 128); 0 = off
    const bool want = allow_split && (mode == 1 || (mode == 2 && std::min(e->bgroup, e->max_streams) <= 128));
    T.ksplit = (want && 2 * (rows / T.n_tile) <= e->n_sms && K % (2 * kKc) == 0) ? 2 : 1;
    const int Kh = K / T.ksplit;
    T.K = Kh;
    T.n_tiles = (rows / T.n_tile) * T.ksplit;
    int kpp = (kBSlotBytes / (T.n_tile * 2)) / kKc * kKc;
    if (kpp > Kh) kpp = Kh;
    T.kpp = kpp;
    T.n_pieces = (Kh + kpp - 1) / kpp;
    T.p = e->dalloc<unsigned char>((size_t) rows * K * 2, false);
    if (!T.p) return e->fail("out of device memory for tiled weights");
    k_tile_weights<<<1184, 256, 0, e->stream>>>(src, T.p, rows, K, T.n_tile, kpp, e->bfmt == 0 ? 0 : 1, T.ksplit);
    QTTS_CUDA(cudaGetLastError());
    e->weight_bytes_batched += (int64_t) rows * K * 2;
    return true;
}

static BOp bblank(int B) {
    BOp o;
    memset(&o, 0, sizeof(o));
    o.n_streams = B; o.rows_act = B; o.n_tok = 1;
    return o;
}

struct BEmit {
    Engine * e;
    std::vector<BOp> P;
    std::vector<BSched> S;
    int B, cta_off = 0;
    unsigned long long * part_ll = nullptr;
    __half * xn16 = nullptr;     // normalised F16 rows (canonical chunks) handed from a BOP_NORM stage to its GEMM
    bool norm_stage = true;
    bool rows_signal = true;     // norm -> GEMM hand-off through a row counter instead of a grid barrier
    bool msplit_on = true;
    float * part32 = nullptr;            // deferred split-K: second half's partial sums, laid out like the output
    bool defer = false;
    const float * pending_x = nullptr;   // output of a deferred-split projection whose partial has not been added yet
    std::string bad;                     // first shape the kernel cannot run (checked by the caller: fail loudly, never fault)

    void gemm(const TiledW & T, int rows_act, int in_kind, const void * in, int in_row_stride, int in_tok_stride,
              const float * norm_w, const float * ssq_in, int n_parts_in, int epi, float * out32, __half * out16,
              const float * res, const float * bias, int out_row_stride, int out_tok_stride, float * ssq_out,
              float * hid_out = nullptr, int flags = 0) {
        if (in_kind == BIN_F32 && norm_w && norm_stage && xn16) {
            // RMSNorm once per row in its own small stage; the GEMM then takes F16 rows by bulk copy
            BOp n = bblank(B);
            n.type = BOP_NORM; n.rows_act = rows_act; n.K = T.K * T.ksplit;
            n.in = in; n.in_row_stride = in_row_stride; n.in_tok_stride = in_tok_stride;
            n.norm_w = norm_w; n.ssq_in = ssq_in; n.n_parts_in = n_parts_in; n.eps = e->cfg.rms_norm_eps;
            n.out16 = xn16; n.hid_out = hid_out;
            if (pending_x) {
                // this stage reads (a suffix of) the rows the split projection wrote: same strides, same offsets
                n.part32 = part32 + (reinterpret_cast<const float *>(in) - pending_x);
                pending_x = nullptr;
            }
            n.flags |= BF_ASYNC_OUT;            // xn16 is fetched by bulk copies (matters when the stage ends with a grid barrier)
            if (rows_signal) { n.flags |= BF_ROWS_SIGNAL; flags |= BF_ROWS_WAIT; }
            P.push_back(n);
            in_kind = BIN_F16; in = xn16; norm_w = nullptr; ssq_in = nullptr; n_parts_in = 0; hid_out = nullptr;
        }
        BOp o = bblank(B);
        o.type = BOP_GEMM; o.rows_act = rows_act; o.flags = flags | (ssq_out ? BF_SSQ_OUT : 0) | (out16 ? BF_ASYNC_OUT : 0);
        o.Wt = T.p; o.rows = T.rows; o.K = T.K; o.n_tile = T.n_tile; o.n_tiles = T.n_tiles; o.cta_off = cta_off;
        o.kpp = T.kpp; o.n_pieces = T.n_pieces;
        o.ksplit = T.ksplit; o.part_ll = (T.ksplit > 1) ? part_ll : nullptr;
        if (pending_x) { fprintf(stderr, "qtts: deferred split-K output consumed by a stage without a norm\n"); abort(); }
        if (T.ksplit > 1 && defer && epi == EPI_ADD && out32 && norm_stage && xn16) {
            o.part32 = part32; pending_x = out32;
            o.flags &= ~BF_SSQ_OUT;      // the norm stage squares the combined row itself
        }
        // split-M: a narrow projection over >= 2 M tiles gives each (N tile, M half) its own CTA
        const int m_total = (rows_act + 127) / 128;
        const int msplit = (msplit_on && in_kind == BIN_F16 && T.ksplit == 1 && m_total % 2 == 0 && 2 * T.n_tiles <= e->n_sms) ? 2 : 1;
        o.msplit = msplit; o.n_tiles = T.n_tiles * msplit;
        if (in_kind == BIN_F16) {
            // TMEM budget of bstage_gemm_tma: every issuer (one per activation buffer in use) keeps one accumulator of n_tile
            // columns per M tile of a work item; 512 columns are allocated.  Same rule as the device code.
            const int rg = (rows_act + 7) / 8, m_tiles = m_total / msplit;
            const int a_bufs = B > 64 ? 4 : 2;
            const bool four_full = rg > 8 && a_bufs == 4 && m_tiles <= 2 && 4 * m_tiles * T.n_tile <= 512;
            const int nab = (rg <= 8 || four_full) ? 4 : 2;
            if (nab * m_tiles * T.n_tile > 512 && bad.empty())
                bad = "batched kernel: a projection with " + std::to_string(T.n_tile) + "-row weight tiles over " + std::to_string(m_tiles) +
                      " M tiles (" + std::to_string(rows_act) + " activation rows) needs " + std::to_string(nab * m_tiles * T.n_tile) +
                      " TMEM columns (512 available); use smaller tiles (QTTS_NTILE) or a group size whose two-token pass has an even number of M tiles";
        }
        o.in_kind = in_kind; o.in = in; o.in_row_stride = in_row_stride; o.in_tok_stride = in_tok_stride;
        o.norm_w = norm_w; o.ssq_in = ssq_in; o.n_parts_in = n_parts_in; o.eps = e->cfg.rms_norm_eps;
        o.epi = epi; o.out32 = out32; o.out16 = out16; o.res = res; o.bias = bias;
        o.out_row_stride = out_row_stride; o.out_tok_stride = out_tok_stride; o.ssq_out = ssq_out; o.hid_out = hid_out;
        P.push_back(o);
        S.push_back(BSched{T.p, T.n_tile * T.K * 2, T.n_tile * T.kpp * 2, T.n_pieces, T.n_tiles * msplit, cta_off, msplit});
        cta_off = (cta_off + T.n_tiles * msplit) % e->n_sms;
    }
};

// one decoder layer: QKV | attention | O (+res) | gate/up (SwiGLU) | down (+res)
static void emit_blayer(BEmit & E, const TiledLayer & L, const LayerW & W, const Dims & D, int rows_act, int n_tok,
                        float * x, float * qkv, __half * attn16, __half * act16, const float * ssq_first, int parts_first,
                        float * ssq_p, BOp attn) {
    const int B = E.B;
    const int qkvw = (D.nh + 2 * D.nkv) * kHeadDim, aw = D.nh * kHeadDim, parts = D.H / 16;
    E.gemm(L.qkv, rows_act, BIN_F32, x, D.H, B * D.H, W.attn_norm, ssq_first, parts_first, EPI_STORE,
           qkv, nullptr, nullptr, nullptr, qkvw, B * qkvw, nullptr);
    attn.qkv = qkv; attn.qkvw = qkvw; attn.q_norm_w = W.q_norm; attn.k_norm_w = W.k_norm;
    attn.n_heads = D.nh; attn.n_kv_heads = D.nkv; attn.n_tok = n_tok; attn.rows_act = rows_act;
    attn.attn16 = attn16; attn.attn_w = aw;
    attn.flags |= BF_ASYNC_OUT;         // the O projection fetches attn16 by bulk copies
    E.P.push_back(attn);
    E.gemm(L.o, rows_act, BIN_F16, attn16, aw, B * aw, nullptr, nullptr, 0, EPI_ADD,
           x, nullptr, x, nullptr, D.H, B * D.H, ssq_p);
    E.gemm(L.gu, rows_act, BIN_F32, x, D.H, B * D.H, W.ffn_norm, ssq_p, parts, EPI_SWIGLU,
           nullptr, act16, nullptr, nullptr, D.ffn, B * D.ffn, nullptr);
    E.gemm(L.down, rows_act, BIN_F16, act16, D.ffn, B * D.ffn, nullptr, nullptr, 0, EPI_ADD,
           x, nullptr, x, nullptr, D.H, B * D.H, ssq_p);
}

// one Talker decode step of the group: 28 layers over the residual rows x (complete sums of squares in ssq_e),
// output_norm -> hidden_out, codec_head -> logits_out, n_past advanced for the launch's active streams
static void emit_btalker(Engine * e, BGroup & G, BEmit & E, float * x, float * ssq_e, float * logits_out, float * hidden_out) {
    const qtts_config & c = e->cfg;
    BatchBufs & bb = e->bb;
    const int B = G.n, s0 = G.s0, H = c.hidden_size;
    const Dims T{H, c.n_heads, c.n_kv_heads, c.intermediate_size};
    const float kq = 1.0f / sqrtf((float) kHeadDim);
    for (int l = 0; l < c.n_layers; ++l) {
        BOp a = bblank(B);
        a.type = BOP_ATTN_TK; a.rope = e->rope; a.eps = c.rms_norm_eps; a.kq_scale = kq;
        // warp-per-(stream, group) form (field reused as a switch: 2 = on); QTTS_BATTN=0 keeps the CTA-per-item form
        { const char * v = getenv("QTTS_BATTN"); a.pair_f32 = (v && atoi(v) == 0) ? 0 : 2; }
        a.kc = bb.kc + (size_t) s0 * bb.kv_slab + (size_t) l * e->max_ctx * T.nkv * kHeadDim;
        a.vc = bb.vc + (size_t) s0 * bb.kv_slab + (size_t) l * e->max_ctx * T.nkv * kHeadDim;
        a.kv_stream_stride = (int64_t) bb.kv_slab;
        emit_blayer(E, e->t_talker[l], e->talker[l], T, B, 1, x, G.qkv, G.attn16, G.act16,
                    l == 0 ? ssq_e : G.ssq_p, l == 0 ? 1 : H / 16, G.ssq_p, a);
    }
    E.gemm(e->t_codec_head, B, BIN_F32, x, H, 0, e->output_norm, G.ssq_p, H / 16, EPI_STORE,
           logits_out, nullptr, nullptr, nullptr, c.codec_vocab_size, 0, nullptr, hidden_out, BF_INC_NPAST);
}

// frame program of the stream group [s0, s0 + B)
static void emit_bframe(Engine * e, BGroup & G, BEmit & E) {
    const qtts_config & c = e->cfg;
    BatchBufs & bb = e->bb;
    const int B = G.n, s0 = G.s0;
    const int H = c.hidden_size, cpH = c.cp_hidden_size;
    const Dims T{H, c.n_heads, c.n_kv_heads, c.intermediate_size};
    const Dims C{cpH, c.cp_n_heads, c.cp_n_kv_heads, c.cp_intermediate_size};
    const float eps = c.rms_norm_eps;
    const float kq = 1.0f / sqrtf((float) kHeadDim);
    const bool mtp = e->mtp != nullptr;
    const int last = c.n_codebooks - 1;

    BOp s = bblank(B);
    s.type = BOP_SAMPLE; s.kind = 0; s.book = 0; s.n_books = c.n_codebooks;
    s.logits = bb.logits + (size_t) s0 * c.codec_vocab_size; s.logits_stride = c.codec_vocab_size; s.n_vocab = c.codec_vocab_size;
    s.eos_id = c.codec_eos_id; s.suppress_start = c.codec_vocab_size - 1024;
    s.x_out = mtp ? G.emb : G.cp_x; s.xw = H; s.ssq_e = G.cp_ssq_e;
    s.hidden = bb.hidden + (size_t) s0 * H; s.hidden_stride = H; s.table = e->codec_embd;
    s.dump = G.cb0_dump; s.dump_stride = c.codec_vocab_size;
    E.P.push_back(s);

    for (int p = 1; p <= last; ++p) {
        const int n_tok = (p == 1) ? 2 : 1, rows_act = n_tok * B;
        const float * ssq_first = G.cp_ssq_e; int parts_first = 1;
        if (mtp) {
            // talker space -> code-predictor space (src/tts_transformer.cpp:1709-1715)
            E.gemm(e->t_mtp, rows_act, BIN_F32, G.emb, H, B * H, nullptr, nullptr, 0, e->mtp_b ? EPI_BIAS : EPI_STORE,
                   G.cp_x, nullptr, nullptr, e->mtp_b, cpH, B * cpH, G.cp_ssq_p);
            ssq_first = G.cp_ssq_p; parts_first = cpH / 16;
        }
        for (int l = 0; l < c.cp_layers; ++l) {
            BOp a = bblank(B);
            a.type = BOP_ATTN_CP; a.rope = e->rope; a.eps = eps; a.kq_scale = kq;
            a.pos = (p == 1) ? 0 : p; a.pair_f32 = (p == 1) ? 1 : 0;
            a.kc = G.cp_kc + (size_t) l * 16 * C.nkv * kHeadDim; a.vc = G.cp_vc + (size_t) l * 16 * C.nkv * kHeadDim;
            a.kv_stream_stride = (int64_t) c.cp_layers * 16 * C.nkv * kHeadDim;
            emit_blayer(E, e->t_cp[l], e->cp[l], C, rows_act, n_tok, G.cp_x, G.cp_qkv, G.cp_attn16, G.cp_act16,
                        l == 0 ? ssq_first : G.cp_ssq_p, l == 0 ? parts_first : cpH / 16, G.cp_ssq_p, a);
        }
        // the head reads the LAST token of the pass
        E.gemm(e->t_lm_head[p - 1], B, BIN_F32, G.cp_x + (size_t) (n_tok - 1) * B * cpH, cpH, 0, e->cp_output_norm,
               G.cp_ssq_p + (size_t) (n_tok - 1) * B * (cpH / 16), cpH / 16, EPI_STORE,
               G.cp_logits, nullptr, nullptr, nullptr, c.cp_vocab_size, 0, nullptr);
        BOp sp = bblank(B);
        sp.type = BOP_SAMPLE; sp.kind = 1; sp.book = p; sp.n_books = c.n_codebooks;
        sp.logits = G.cp_logits; sp.logits_stride = c.cp_vocab_size; sp.n_vocab = c.cp_vocab_size;
        sp.xw = H;
        sp.dump = G.cp_dump ? G.cp_dump + (size_t) (p - 1) * B * c.cp_vocab_size : nullptr; sp.dump_stride = c.cp_vocab_size;
        if (p < last) {
            sp.table = e->cp_embd[p - 1]; sp.x_out = mtp ? G.emb : G.cp_x; sp.ssq_e = G.cp_ssq_e;
        } else {
            sp.es = e->slots[s0].es; sp.x_out = G.x; sp.ssq_e = G.ssq_e;
            sp.trailing = bb.trailing + (size_t) s0 * bb.trailing_stride; sp.trailing_stride = (int32_t) bb.trailing_stride;
            sp.pad = bb.pad + (size_t) s0 * H; sp.pad_stride = H;
            sp.codes_out = bb.codes + (size_t) s0 * bb.codes_stride; sp.codes_stride = (int32_t) bb.codes_stride;
        }
        E.P.push_back(sp);
    }
    emit_btalker(e, G, E, G.x, G.ssq_e, bb.logits + (size_t) s0 * c.codec_vocab_size, bb.hidden + (size_t) s0 * H);
}

uint32_t batched_idesc(const Engine * e) {
    // cute::UMMA::InstrDescriptor: c_format F32 [4,6) = 1, a_format [7,10) F16 = 0, b_format [10,13)
    // (F16 = 0 / BF16 = 1), a/b K-major (bits 15, 16 = 0), N >> 3 [17,23) (per stage), M >> 4 [24,29)
    return (1u << 4) | (0u << 7) | ((uint32_t) (e->bfmt ? 1 : 0) << 10) | ((128u >> 4) << 24);
}

bool batched_init(Engine * e) {
    const qtts_config & c = e->cfg;
    if (cudaFuncSetAttribute(qtts_run_batched, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) kBSmemBytes) != cudaSuccess ||
        cudaFuncSetAttribute(qtts_run_batched_ext, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) kBSmemBytes) != cudaSuccess)
        return e->fail(std::string("batched kernel: cudaFuncSetAttribute: ") + cudaGetErrorString(cudaGetLastError()));
    int occ = 0, occ_ext = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, qtts_run_batched, kThreads, kBSmemBytes);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_ext, qtts_run_batched_ext, kThreads, kBSmemBytes);
    if (occ < 1 || occ_ext < 1) return e->fail("batched kernel does not fit one CTA per SM");
    {   // clusters of 2 for the _ext kernel (opt-in, QTTS_BCLUSTER=1): can the driver co-schedule n_sms / 2 of them?
        const char * env = getenv("QTTS_BCLUSTER");
        e->bcluster = 0;
        if (env && atoi(env) != 0 && e->n_sms % 2 == 0) {        // measured neutral at 256 streams, -1 % at <= 128 (profiles/r02_summary.md section 12)
            cudaLaunchConfig_t lc;
            memset(&lc, 0, sizeof(lc));
            lc.gridDim = dim3(e->n_sms); lc.blockDim = dim3(kThreads); lc.dynamicSmemBytes = kBSmemBytes; lc.stream = e->stream;
            cudaLaunchAttribute at[2];
            at[0].id = cudaLaunchAttributeCooperative; at[0].val.cooperative = 1;
            at[1].id = cudaLaunchAttributeClusterDimension; at[1].val.clusterDim.x = 2; at[1].val.clusterDim.y = 1; at[1].val.clusterDim.z = 1;
            lc.attrs = at; lc.numAttrs = 2;
            int n_clusters = 0;
            if (cudaOccupancyMaxActiveClusters(&n_clusters, (const void *) qtts_run_batched_ext, &lc) == cudaSuccess && 2 * n_clusters >= e->n_sms)
                e->bcluster = 1;
            cudaGetLastError();
        }
    }
    const int rep = c.n_heads / c.n_kv_heads;
    if (rep > kMaxRep || e->max_ctx * rep > 2 * kMaxScores) return e->fail("batched path: context too long for the attention scratch");
    if (const char * v = getenv("QTTS_BGROUP")) e->bgroup = std::max(8, std::min(atoi(v), kBMaxGroup));
    // tiled weights
    e->t_talker.resize(c.n_layers); e->t_cp.resize(c.cp_layers); e->t_lm_head.resize(c.n_codebooks - 1);
    auto layer = [&](const LayerW & W, const Dims & D, TiledLayer & L) {
        return tile_matrix(e, W.wqkv, (D.nh + 2 * D.nkv) * kHeadDim, D.H, L.qkv) &&
               tile_matrix(e, W.wo, D.H, D.nh * kHeadDim, L.o, true) &&
               tile_matrix(e, W.wgu, 2 * D.ffn, D.H, L.gu) &&
               tile_matrix(e, W.wdown, D.H, D.ffn, L.down, true);
    };
    const Dims T{c.hidden_size, c.n_heads, c.n_kv_heads, c.intermediate_size};
    const Dims C{c.cp_hidden_size, c.cp_n_heads, c.cp_n_kv_heads, c.cp_intermediate_size};
    for (int i = 0; i < c.n_layers; ++i) if (!layer(e->talker[i], T, e->t_talker[i])) return false;
    for (int i = 0; i < c.cp_layers; ++i) if (!layer(e->cp[i], C, e->t_cp[i])) return false;
    for (int i = 0; i < c.n_codebooks - 1; ++i)
        if (!tile_matrix(e, e->lm_head[i], c.cp_vocab_size, c.cp_hidden_size, e->t_lm_head[i])) return false;
    if (!tile_matrix(e, e->codec_head, c.codec_vocab_size, c.hidden_size, e->t_codec_head)) return false;
    if (e->mtp && !tile_matrix(e, e->mtp, c.cp_hidden_size, c.hidden_size, e->t_mtp)) return false;
    QTTS_CUDA(cudaStreamSynchronize(e->stream));
    return true;
}

// lock-step groups of at most e->bgroup (<= 256) streams over the current slots (the merged code-predictor pass has
// 2 rows per stream: kBMaxRows = 512)
bool batched_build_groups(Engine * e) {
    const qtts_config & c = e->cfg;
    const int rep = c.n_heads / c.n_kv_heads;
    if (rep > kMaxRep || e->max_ctx * rep > 2 * kMaxScores) return e->fail("batched path: context too long for the attention scratch");
    e->groups.clear();
    const int H = c.hidden_size, cpH = c.cp_hidden_size;
    const int qkvw = (c.n_heads + 2 * c.n_kv_heads) * kHeadDim, cqkvw = (c.cp_n_heads + 2 * c.cp_n_kv_heads) * kHeadDim;
    for (int s0 = 0; s0 < e->max_streams; s0 += e->bgroup) {
        BGroup G;
        G.s0 = s0; G.n = std::min(e->bgroup, e->max_streams - s0);
        const size_t B = (size_t) G.n;
        const size_t B8 = (B + 7) & ~(size_t) 7, B28 = (2 * B + 7) & ~(size_t) 7;   // F16 activations: canonical chunks, rows padded to 8
        G.x = e->dalloc<float>(B * H); G.qkv = e->dalloc<float>(B * qkvw);
        G.attn16 = e->dalloc<__half>(B8 * c.n_heads * kHeadDim); G.act16 = e->dalloc<__half>(B8 * c.intermediate_size);
        G.ssq_e = e->dalloc<float>(B); G.ssq_p = e->dalloc<float>(B * (H / 16));
        G.emb = e->dalloc<float>(2 * B * H);
        G.cp_x = e->dalloc<float>(2 * B * std::max(H, cpH)); G.cp_qkv = e->dalloc<float>(2 * B * cqkvw);
        G.cp_attn16 = e->dalloc<__half>(B28 * c.cp_n_heads * kHeadDim); G.cp_act16 = e->dalloc<__half>(B28 * c.cp_intermediate_size);
        G.cp_ssq_e = e->dalloc<float>(2 * B); G.cp_ssq_p = e->dalloc<float>(2 * B * (cpH / 16));
        G.cp_logits = e->dalloc<float>(B * c.cp_vocab_size);
        G.cp_kc = e->dalloc<__half>(B * c.cp_layers * 16 * c.cp_n_kv_heads * kHeadDim);
        G.cp_vc = e->dalloc<__half>(B * c.cp_layers * 16 * c.cp_n_kv_heads * kHeadDim);
        G.cb0_dump = e->dalloc<float>(B * c.codec_vocab_size);
        G.cp_dump = e->dalloc<float>(B * 15 * c.cp_vocab_size);
        G.active = e->dalloc<int32_t>(B);
        if (!G.x || !G.qkv || !G.cp_x || !G.cp_qkv || !G.cp_logits || !G.cp_kc || !G.cp_vc || !G.cp_dump || !G.active)
            return e->fail("out of device memory for the batched activations");
        // split-K partials: one flag-carrying word per accumulator element of the widest split projection
        G.part_words = (size_t) std::max(H, cpH) * kBMaxRows;
        G.part_ll = e->dalloc<unsigned long long>(G.part_words);
        if (!G.part_ll) return e->fail("out of device memory for the split-K partials");
        G.xn16 = e->dalloc<__half>(B28 * (size_t) std::max(H, cpH));
        if (!G.xn16) return e->fail("out of device memory for the normalised activations");
        G.part32 = e->dalloc<float>(2 * B * (size_t) std::max(H, cpH));
        if (!G.part32) return e->fail("out of device memory for the split-K partials");
        BEmit E{e, {}, {}, G.n, 0, G.part_ll, G.xn16};
        E.part32 = G.part32;
        { const char * v = getenv("QTTS_KSPLIT"); E.defer = !v || atoi(v) == 2; }
        if (const char * v = getenv("QTTS_BNORM")) E.norm_stage = atoi(v) != 0;
        if (const char * v = getenv("QTTS_BROWS")) E.rows_signal = atoi(v) != 0;
        if (const char * v = getenv("QTTS_MSPLIT")) E.msplit_on = atoi(v) != 0;
        emit_bframe(e, G, E);
        if (!E.bad.empty()) return e->fail(E.bad);
        G.n_ops = (int) E.P.size(); G.n_sched = (int) E.S.size();
        G.ops = e->dalloc<BOp>(E.P.size(), false);
        G.sched = e->dalloc<BSched>(E.S.size(), false);
        if (!G.ops || !G.sched) return e->fail("out of device memory for the batched program");
        if (!e->h2d(G.ops, E.P.data(), E.P.size() * sizeof(BOp))) return false;
        if (!e->h2d(G.sched, E.S.data(), E.S.size() * sizeof(BSched))) return false;
        G.host_ops = E.P;
        // batched prefill program: one Talker step over private residual rows / outputs
        G.pf_x = e->dalloc<float>(B * H); G.pf_ssq_e = e->dalloc<float>(B);
        G.pf_logits = e->dalloc<float>(B * c.codec_vocab_size); G.pf_hidden = e->dalloc<float>(B * H);
        G.pf_rows = e->dalloc<const float *>(B); G.pf_plen = e->dalloc<int32_t>(B); G.pf_active = e->dalloc<int32_t>(B);
        if (!G.pf_x || !G.pf_ssq_e || !G.pf_logits || !G.pf_hidden || !G.pf_rows || !G.pf_plen || !G.pf_active)
            return e->fail("out of device memory for the batched prefill");
        BEmit F{e, {}, {}, G.n, 0, G.part_ll, G.xn16};
        F.part32 = G.part32; F.defer = E.defer; F.norm_stage = E.norm_stage; F.rows_signal = E.rows_signal; F.msplit_on = E.msplit_on;
        emit_btalker(e, G, F, G.pf_x, G.pf_ssq_e, G.pf_logits, G.pf_hidden);
        if (!F.bad.empty()) return e->fail(F.bad);
        G.pf_n_ops = (int) F.P.size(); G.pf_n_sched = (int) F.S.size();
        G.pf_ops = e->dalloc<BOp>(F.P.size(), false);
        G.pf_sched = e->dalloc<BSched>(F.S.size(), false);
        if (!G.pf_ops || !G.pf_sched) return e->fail("out of device memory for the batched prefill program");
        if (!e->h2d(G.pf_ops, F.P.data(), F.P.size() * sizeof(BOp))) return false;
        if (!e->h2d(G.pf_sched, F.S.data(), F.S.size() * sizeof(BSched))) return false;
        e->groups.push_back(G);
    }
    e->batched_ready = true;
    return true;
}

bool launch_batched(Engine * e, const BOp * d_ops, int n_ops, const BSched * d_sched, int n_sched, StreamState * st,
                    const int32_t * d_active, int n_streams, int n_frames, bool check_done) {
    QTTS_CUDA(cudaMemsetAsync(e->bar, 0, sizeof(GridBarrier), e->stream));     // arrival counter and row counter
    BRun run;
    memset(&run, 0, sizeof(run));
    run.ops = d_ops; run.n_ops = n_ops; run.sched = d_sched; run.n_sched = n_sched;
    run.st = st; run.active = d_active; run.bar = e->bar;
    run.n_streams = n_streams; run.n_frames = n_frames; run.check_done = check_done ? 1 : 0;
    run.idesc_base = batched_idesc(e);
    {   // stage tags never repeat across launches (stale partials can not satisfy a poll)
        const uint64_t n_tags = (uint64_t) n_ops * (uint64_t) n_frames + 1;
        if ((uint64_t) e->bll_seq + n_tags > 0xfff00000ull) {
            QTTS_CUDA(cudaStreamSynchronize(e->stream));
            for (BGroup & g : e->groups) QTTS_CUDA(cudaMemsetAsync(g.part_ll, 0, g.part_words * 8, e->stream));
            e->bll_seq = 0;
        }
        run.tag_base = e->bll_seq;
        e->bll_seq += (uint32_t) n_tags;
    }
    run.prof = e->prof_on ? e->prof + 2048 : nullptr;
    { const char * d = getenv("QTTS_BDBG"); run.dbg = d ? atoi(d) : 0; }
    run.trace = e->trace_on ? e->trace : nullptr; run.trace_cta = e->trace_cta; run.trace_cap = e->trace_cap / 2;
    void * args[] = {(void *) &run};
    // Thread-block clusters of 2 (148 = 2 x 74 packs): the CTAs of a pair share the activation operand of the GEMM stages
    // by multicast.  Still a cooperative launch (the grid barrier needs every CTA resident).  e->bcluster (set by batched_init):
    // 0 = off (default; odd SM count, or the driver cannot co-schedule n_sms / 2 clusters), 1 = on (QTTS_BCLUSTER=1 at load).
    cudaLaunchConfig_t lc;
    memset(&lc, 0, sizeof(lc));
    lc.gridDim = dim3(e->n_sms); lc.blockDim = dim3(kThreads); lc.dynamicSmemBytes = kBSmemBytes; lc.stream = e->stream;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeCooperative; at[0].val.cooperative = 1;
    at[1].id = cudaLaunchAttributeClusterDimension; at[1].val.clusterDim.x = 2; at[1].val.clusterDim.y = 1; at[1].val.clusterDim.z = 1;
    lc.attrs = at;
    run.cluster = e->bcluster > 0 ? 1 : 0;      // probed at load time (batched_init)
    // activation buffers vs weight-ring slots (bring_slots): groups of more than 64 rows take 4 buffers + 3 slots
    { const char * ab = getenv("QTTS_BABUFS"); run.a_bufs = ab ? (atoi(ab) == 4 ? 4 : 2) : (run.n_streams > 64 ? 4 : 2); }
    // qtts_run_batched: <= 64 streams, compile-time configuration; qtts_run_batched_ext: four activation buffers and / or clusters
    const bool ext = run.a_bufs == 4 || run.cluster;
    if (!ext) { run.a_bufs = 2; run.cluster = 0; }
    lc.numAttrs = run.cluster ? 2 : 1;
    QTTS_CUDA(cudaLaunchKernelExC(&lc, ext ? (const void *) qtts_run_batched_ext : (const void *) qtts_run_batched, args));
    e->launches += 1;
    return true;
}

bool run_batched_group(Engine * e, BGroup & G, int n_frames) {
    std::vector<int32_t> act(G.n);
    bool any = false;
    for (int i = 0; i < G.n; ++i) { act[i] = e->slots[G.s0 + i].active ? 1 : 0; any |= act[i] != 0; }
    if (!any) return true;
    QTTS_CUDA(cudaMemcpyAsync(G.active, act.data(), act.size() * sizeof(int32_t), cudaMemcpyHostToDevice, e->stream));
    QTTS_CUDA(cudaStreamSynchronize(e->stream));
    return launch_batched(e, G.ops, G.n_ops, G.sched, G.n_sched, e->bb.st + G.s0, G.active, G.n, n_frames, true);
}

// ------------------------------------------------------------------------------------------
// Batched prefill (forward_prefill, src/tts_transformer.cpp:1829-1920, for many utterances at once): position t of
// every stream being prefilled is one lock-step Talker step of the tcgen05 kernel (weights streamed once per position
// for the whole group instead of once per stream and position).  Causality comes from the KV cache exactly as in the
// token-by-token batch-1 prefill: position t attends to the F16 K/V rows 0..t written so far.
// Streams of the group that are mid-utterance are untouched: the program works on pf_x / pf_logits / pf_hidden, the
// launch's active mask covers only (stream being prefilled, t < its prefill length), and the K/V row a masked stream
// gets written at its own n_past is overwritten by that stream's next real step before anything reads it.
// ------------------------------------------------------------------------------------------
__global__ void k_prefill_gather(const float * const * __restrict__ rows, const int32_t * __restrict__ plen, int t,
                                 float * __restrict__ x, float * __restrict__ ssq_e, int32_t * __restrict__ active, int H) {
    __shared__ double red[32];
    const int b = blockIdx.x;
    const bool on = rows[b] != nullptr && t < plen[b];
    if (threadIdx.x == 0) active[b] = on ? 1 : 0;
    if (!on) return;
    const float * src = rows[b] + (size_t) t * H;
    double ss = 0.0;
    for (int i = threadIdx.x; i < H; i += blockDim.x) {
        const float v = src[i];
        x[(size_t) b * H + i] = v;
        ss += (double) (v * v);
    }
    for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    if (threadIdx.x == 0) {
        double tot = 0.0;
        for (int w = 0; w < (int) (blockDim.x >> 5); ++w) tot += red[w];
        ssq_e[b] = (float) tot;
    }
}

// logits / hidden of the streams whose LAST prefill position was t -> their rows of the per-stream slabs
__global__ void k_prefill_scatter(const float * const * __restrict__ rows, const int32_t * __restrict__ plen, int t,
                                  const float * __restrict__ lg, float * __restrict__ logits, int V,
                                  const float * __restrict__ hid, float * __restrict__ hidden, int H) {
    const int b = blockIdx.x;
    if (rows[b] == nullptr || plen[b] - 1 != t) return;
    for (int i = threadIdx.x; i < V; i += blockDim.x) logits[(size_t) b * V + i] = lg[(size_t) b * V + i];
    for (int i = threadIdx.x; i < H; i += blockDim.x) hidden[(size_t) b * H + i] = hid[(size_t) b * H + i];
}

bool run_batched_prefill(Engine * e, BGroup & G, const std::vector<const float *> & rows, const std::vector<int32_t> & plen) {
    const qtts_config & c = e->cfg;
    const int H = c.hidden_size, V = c.codec_vocab_size;
    int max_len = 0;
    for (int i = 0; i < G.n; ++i) if (rows[i]) max_len = std::max(max_len, (int) plen[i]);
    if (max_len == 0) return true;
    QTTS_CUDA(cudaMemcpyAsync(G.pf_rows, rows.data(), G.n * sizeof(const float *), cudaMemcpyHostToDevice, e->stream));
    QTTS_CUDA(cudaMemcpyAsync(G.pf_plen, plen.data(), G.n * sizeof(int32_t), cudaMemcpyHostToDevice, e->stream));
    QTTS_CUDA(cudaStreamSynchronize(e->stream));      // rows / plen are caller-owned host vectors
    for (int t = 0; t < max_len; ++t) {
        k_prefill_gather<<<G.n, 256, 0, e->stream>>>(G.pf_rows, G.pf_plen, t, G.pf_x, G.pf_ssq_e, G.pf_active, H);
        if (!launch_batched(e, G.pf_ops, G.pf_n_ops, G.pf_sched, G.pf_n_sched, e->bb.st + G.s0, G.pf_active, G.n, 1, false)) return false;
        k_prefill_scatter<<<G.n, 256, 0, e->stream>>>(G.pf_rows, G.pf_plen, t, G.pf_logits, e->bb.logits + (size_t) G.s0 * V, V,
                                                       G.pf_hidden, e->bb.hidden + (size_t) G.s0 * H, H);
        e->launches += 2;
    }
    QTTS_CUDA(cudaGetLastError());
    return true;
}

}  // namespace qtts
