// qtts_batched.cu — the BATCHED frame kernel: all concurrent utterance streams of a GPU advance
// in lock step inside ONE persistent cooperative launch (1 CTA per SM, 512 threads).
//
// Per frame the kernel walks the same stage list as the batch-1 kernel (qtts_kernels.cu), but
//   * every projection is a GEMM  D[rows_act x n_tile] = A[rows_act x K] . W_tile[n_tile x K]^T  issued
//     by ONE thread per CTA as tcgen05.mma (cta_group::1, kind::f16, M = 128, N = n_tile) with the
//     FP32 accumulator in TMEM; A = the streams' F16-rounded activations (the reference's
//     ggml_mul_mat rounding point, ggml/src/ggml-cpu/ggml-cpu.c:1229-1320), staged into shared memory
//     by all threads in 128-column chunks (double buffered, RMSNorm applied on the way); W tiles
//     are pre-arranged in HBM in the UMMA canonical layout and stream through a shared-memory
//     ring with cp.async.bulk (TMA, 1-D) + mbarrier complete_tx, prefetched across stage and
//     frame boundaries exactly like the batch-1 kernel's ring; ring slots are recycled by
//     tcgen05.commit -> mbarrier; the epilogue reads TMEM with tcgen05.ld (32x32b.x16) and fuses
//     residual add / SwiGLU / bias and the sum-of-squares partials the next RMSNorm needs;
//   * weight tiles of one stage go to CTAs (cta_off + t) mod grid with cta_off rotating from
//     stage to stage, so that the HBM stream is spread over all SMs across a frame although one
//     stage has fewer tiles than SMs;
//   * attention: code predictor = one warp per (stream, KV group) (<= 16 keys); Talker = one CTA
//     pass per (stream, KV group) with FP32 softmax / P.V (the batch-1 kernel's ATTN_FAST numerics);
//   * sampling / embedding / frame end run per stream, distributed over CTAs.
// Reference behaviour reproduced: src/tts_transformer.cpp:2416-2560 (frame loop), :2153-2340 (code
// predictor), :1376-1512 (Talker step), :2417-2499 (CB0 post-processing and sampling).
#ifndef QTTS_BAR_RED
#define QTTS_BAR_RED 0
#endif
#include "qtts_bops.cuh"

namespace qtts {

// ------------------------------------------------------------------------------------------
// tcgen05 / TMEM wrappers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t umma_smem_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    // cute::UMMA::SmemDescriptor: start >> 4 [0,14), LBO >> 4 [16,30), SBO >> 4 [32,46), version 1 [46,48),
    // layout type SWIZZLE_NONE [61,64) = 0
    return (uint64_t) ((addr & 0x3FFFFu) >> 4) | ((uint64_t) (lbo_bytes >> 4) << 16) |
           ((uint64_t) (sbo_bytes >> 4) << 32) | (1ull << 46);
}
__device__ __forceinline__ void umma_f16_ss(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 :: "r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t * bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                 :: "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after()  { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float * v) {
    uint32_t r[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ bool mbar_test(uint64_t * bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}

// ------------------------------------------------------------------------------------------
// CTA state.  With ~222 KB of shared memory per CTA there is almost no L1 left, so anything the
// compiler would put on the stack (structs passed by reference to the stage functions, spills)
// turns into L2 round trips; every piece of CTA state therefore lives in ONE shared-memory block
// at the start of dynamic shared memory, and the stage descriptor is copied there per stage.
//
// weight ring: piece number `seq` (counted from kernel start, in the CTA's own consumption order)
// lives in slot seq % kBSlots; full[slot] completes by bytes (bulk copy), empty[slot] by the
// tcgen05.commit that follows the last MMA reading the slot.
// ------------------------------------------------------------------------------------------
constexpr int kIssuersMax = 4;
struct BPrefetch {           // used by the prefetcher thread only
    int32_t idx, frame, t, p;
    int32_t n_tiles, n_pieces, tile_bytes, piece_bytes, msplit;
    const unsigned char * W;
    uint32_t seq;
    int32_t exhausted;
    uint64_t policy;
    // ring position kept incrementally (no division by the run-time slot count on the issue path)
    int32_t slot, nsl, paired;
    uint32_t par;            // parity to wait for on empty[slot]
    uint32_t ring;           // shared-memory address of the ring
};

struct BShared {
    uint64_t full[kBSlots], empty[kBSlots], abar[2], dbar;
    uint64_t tfull[4], tempty[4];     // A pieces brought by bulk copies (F16-input stages)
    uint32_t n_fill[4], fill_base[4];  // fills issued per A buffer (producer's counter) / its value at the start of the stage
    BRun run;                // copy of the kernel parameter
    BOp op;                  // the stage being executed
    Op item_op;              // Talker attention: per-item view for the batch-1 helpers
    BPrefetch pf;
    double red[kWarps];
    float scale[kBMaxRows];
    uint32_t tmem_base;
    uint32_t cons_seq;       // ring pieces consumed so far
    uint32_t a_use[2];       // commits issued on abar[i]
    uint32_t d_use;          // commits issued on dbar
    uint32_t tag;            // this stage's tag (split-K hand-off)
    uint32_t rows_target;    // rows signalled by all BF_ROWS_SIGNAL stages of this launch so far (uniform over the grid)
    int32_t all_done;
    uint32_t * tr;           // this stage's 16-word trace slot (nullptr: not traced)
};
#define BTR(k) do { if (S->tr) S->tr[k] = (uint32_t) clock64(); } while (0)
constexpr size_t kBSharedBytes = (sizeof(BShared) + 127) & ~(size_t) 127;

// [BShared][A buffers 2 x kABufBytes][weight ring]: an M = 128 MMA reads 128 rows although a piece
// may hold fewer; the over-read then falls into the ring (valid shared memory), never past the end
__device__ __forceinline__ unsigned char * babuf_base(BShared * S) { return reinterpret_cast<unsigned char *>(S) + kBSharedBytes; }
// Shared memory behind BShared: [activation buffers: a_bufs x kABufBytes][weight ring: bring_slots x kBSlotBytes].  Small groups
// (<= 64 rows) are bound by how far the weights run ahead: 2 buffers (used as 4 halves) + 6 ring slots.  Large groups move
// 8x more activation bytes than weight bytes per CTA and stage and were bound by the turnaround of two buffers
// (copy -> MMA -> commit -> next copy, ~0.9 us, profiles/r02_summary.md section 11): 4 buffers + 3 ring slots.
#ifndef QTTS_BRING_RUNTIME
#define QTTS_BRING_RUNTIME 1
#endif
#ifndef QTTS_BCLUSTER_CODE
#define QTTS_BCLUSTER_CODE 1
#endif
__device__ __forceinline__ int bring_slots(const BShared * S) { return (QTTS_BRING_RUNTIME && S->run.a_bufs == 4) ? kBSlots - 3 : kBSlots; }
__device__ __forceinline__ unsigned char * bring_base(BShared * S) { return babuf_base(S) + (size_t) ((QTTS_BRING_RUNTIME && S->run.a_bufs == 4) ? 4 : 2) * kABufBytes; }
static_assert(3 * kBSlotBytes == 2 * kABufBytes, "two more activation buffers replace exactly three ring slots");
// byte offset of element (row r, column col) of an F16 activation matrix stored in global memory
// as UMMA-canonical chunks of 128 columns (chunk_bytes = ceil(rows / 8) * 2048): LBO 128, SBO 2048
__device__ __forceinline__ size_t canon_off(int r, int col, int chunk_bytes) {
    return (size_t) (col >> 7) * chunk_bytes + (size_t) (r >> 3) * 2048 + (size_t) ((col & 127) >> 3) * 128 + (r & 7) * 16 + (col & 7) * 2;
}
__device__ __forceinline__ int canon_chunk_bytes(int rows) { return ((rows + 7) >> 3) << 11; }

__device__ __forceinline__ int first_tile(int cta_off) {
    const int G = (int) gridDim.x;
    int f = ((int) blockIdx.x - cta_off) % G;
    return f < 0 ? f + G : f;
}

__device__ __forceinline__ int first_tile_of(int block, int cta_off) {
    const int G = (int) gridDim.x;
    int f = (block - cta_off) % G;
    return f < 0 ? f + G : f;
}
// split-M work items: virtual tile t -> (weight tile, M half).  Without clusters adjacent items are the two halves of one
// weight tile; in cluster mode the two CTAs of a pair (adjacent items) must read the SAME activation rows to share them
// by multicast, so adjacent items are two weight tiles of the same half: t = 4a + 2h + b -> (tile 2a + b, half h).
__device__ __forceinline__ bool msplit_paired(int cluster, int msplit, int n_tiles) { return cluster && msplit == 2 && (n_tiles & 3) == 0; }
__device__ __forceinline__ int tile_weight(int t, int msplit, bool paired) { return paired ? (((t >> 2) << 1) | (t & 1)) : t / msplit; }
__device__ __forceinline__ int tile_half(int t, int msplit, bool paired) { return paired ? ((t >> 1) & 1) : t % msplit; }

// cluster mode: bulk copy multicast to the CTAs of `mask` (same shared-memory offsets, each CTA's own mbarrier counts the bytes)
__device__ __forceinline__ void bulk_g2s_mc(void * dst, const void * src, uint32_t bytes, uint64_t * bar, uint16_t mask, uint64_t policy) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster.L2::cache_hint [%0], [%1], %2, [%3], %4, %5;"
                 :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)), "h"(mask), "l"(policy) : "memory");
}
// tcgen05.commit that arrives on the mbarrier at this offset in every CTA of `mask`
__device__ __forceinline__ void umma_commit_mc(uint64_t * bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 :: "r"(smem_u32(bar)), "h"(mask) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// advance to the next piece of this CTA; false when the launch needs no more weights
__device__ __forceinline__ bool bpf_advance(BShared * S) {
    BPrefetch & pf = S->pf;
    if (pf.exhausted) return false;
    if (pf.p >= pf.n_pieces) { pf.t += (int) gridDim.x; pf.p = 0; }
    while (pf.t >= pf.n_tiles) {
        if (++pf.idx >= S->run.n_sched) { pf.idx = 0; if (++pf.frame >= S->run.n_frames) { pf.exhausted = 1; return false; } }
        const int4 e0 = __ldg(reinterpret_cast<const int4 *>(S->run.sched + pf.idx));
        const int4 e1 = __ldg(reinterpret_cast<const int4 *>(S->run.sched + pf.idx) + 1);
        pf.W = reinterpret_cast<const unsigned char *>(((unsigned long long) (unsigned) e0.y << 32) | (unsigned) e0.x);
        pf.tile_bytes = e0.z; pf.piece_bytes = e0.w; pf.n_pieces = e1.x; pf.n_tiles = e1.y; pf.msplit = e1.w > 1 ? e1.w : 1;
        pf.t = first_tile(e1.z); pf.p = 0;
        pf.paired = msplit_paired(S->run.cluster, pf.msplit, pf.n_tiles) ? 1 : 0;
    }
    return true;
}

// issue the bulk copy of the next piece; blocking = wait for the slot to be recycled
__device__ __forceinline__ bool bpf_issue(BShared * S, bool blocking) {
    if (!bpf_advance(S)) return false;
    BPrefetch & pf = S->pf;
    const int slot = pf.slot;
    const uint32_t par = pf.par;
    if (blocking) mbar_wait(S->empty + slot, par);
    else if (!mbar_test(S->empty + slot, par)) return false;
    const int off = pf.p * pf.piece_bytes;
    const uint32_t bytes = (uint32_t) min(pf.piece_bytes, pf.tile_bytes - off);
    mbar_expect_tx(S->full + slot, bytes);
    const int wt = pf.paired ? tile_weight(pf.t, 2, true) : pf.t / pf.msplit;
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 :: "r"(pf.ring + (uint32_t) slot * (uint32_t) kBSlotBytes), "l"(pf.W + (size_t) wt * pf.tile_bytes + off), "r"(bytes),
                    "r"(smem_u32(S->full + slot)), "l"(pf.policy) : "memory");
    if (slot + 1 == pf.nsl) { pf.slot = 0; pf.par = par ^ 1u; } else pf.slot = slot + 1;
    pf.p++; pf.seq++;
    return true;
}
__device__ __forceinline__ void bpf_poll(BShared * S) {
    while (bpf_issue(S, false)) {}
}
__device__ __forceinline__ void bpf_ensure(BShared * S, uint32_t seq_needed) {
    while ((int32_t) (S->pf.seq - seq_needed) <= 0) if (!bpf_issue(S, true)) break;
}

__device__ __forceinline__ void bgrid_barrier(BShared * S, unsigned int & epoch, bool async_out) {
    // F16 activations are re-read by bulk copies (async proxy): writer-side proxy fence, only behind such stages
    if (async_out) asm volatile("fence.proxy.async.global;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {
        BTR(15);
        const unsigned int target = epoch + gridDim.x;
        epoch = target;
        unsigned int * cnt = &S->run.bar->count;
#if QTTS_BAR_RED
        // release-reduction: no return value to wait for; cumulative over the CTA's writes ordered by the bar.sync above
        asm volatile("red.release.gpu.global.add.u32 [%0], %1;" :: "l"(cnt), "r"(1u) : "memory");
#else
        __threadfence();
        atomicAdd(cnt, 1u);
#endif
        BTR(9);
        unsigned int v;
        do {
            asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(cnt) : "memory");
        } while ((int) (v - target) < 0);
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(cnt) : "memory");
        BTR(10);
    } else if (threadIdx.x == kPfThread) {
        bpf_poll(S);
    }
    __syncthreads();
}

// ------------------------------------------------------------------------------------------
// BOP_GEMM
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ int row_off(int r, int B, int row_stride, int tok_stride) {
    const int t = (r >= B) ? 1 : 0;
    return (r - t * B) * row_stride + t * tok_stride;
}

// A "load unit" is what one pass of the 512 threads moves with 4 x 16-byte loads each:
//   BIN_F32: 64 rows x 128 K-elements (unit u of a chunk = rows 64u .. 64u+63),
//   BIN_F16: 128 rows x 128 K-elements.
// global -> registers for unit (kc, mt, u)
template <int IN_KIND>
__device__ __forceinline__ void bload(const BOp & op, int kc, int mt, int u, uint4 * v, float4 & nw) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (IN_KIND == BIN_F32) {
        const float * in = reinterpret_cast<const float *>(op.in);
        const int k = kc * kKc + lane * 4;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int r = mt * 128 + u * 64 + warp + 16 * i;
            if (r < op.rows_act)
                v[i] = __ldcg(reinterpret_cast<const uint4 *>(in + row_off(r, op.n_streams, op.in_row_stride, op.in_tok_stride) + k));
        }
        if (op.norm_w) nw = __ldg(reinterpret_cast<const float4 *>(op.norm_w + k));
    } else {
        const __half * in = reinterpret_cast<const __half *>(op.in);
        const int k = kc * kKc + (lane & 15) * 8;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int r = mt * 128 + warp * 2 + (lane >> 4) + 32 * i;
            if (r < op.rows_act)
                v[i] = __ldcg(reinterpret_cast<const uint4 *>(in + row_off(r, op.n_streams, op.in_row_stride, op.in_tok_stride) + k));
        }
    }
}

// registers -> A buffer (F16, UMMA canonical K-major layout), RMSNorm applied for BIN_F32
template <int IN_KIND>
__device__ __forceinline__ void bstore(const BOp & op, BShared * S, unsigned char * A, int kc, int mt, int u, const uint4 * v,
                                       const float4 & nw, bool write_hid) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (IN_KIND == BIN_F32) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int rl = u * 64 + warp + 16 * i, r = mt * 128 + rl;
            if (r < op.rows_act) {
                float x0 = __uint_as_float(v[i].x), x1 = __uint_as_float(v[i].y);
                float x2 = __uint_as_float(v[i].z), x3 = __uint_as_float(v[i].w);
                if (op.norm_w) {
                    const float s = S->scale[r];
                    x0 = (x0 * s) * nw.x; x1 = (x1 * s) * nw.y; x2 = (x2 * s) * nw.z; x3 = (x3 * s) * nw.w;
                    if (write_hid)
                        *reinterpret_cast<float4 *>(op.hid_out + (size_t) r * op.K + kc * kKc + lane * 4) = make_float4(x0, x1, x2, x3);
                }
                const __half2 h0 = __floats2half2_rn(x0, x1), h1 = __floats2half2_rn(x2, x3);
                uint2 pk;
                pk.x = *reinterpret_cast<const uint32_t *>(&h0);
                pk.y = *reinterpret_cast<const uint32_t *>(&h1);
                *reinterpret_cast<uint2 *>(A + (rl >> 3) * kASbo + (lane >> 1) * kALbo + (rl & 7) * 16 + (lane & 1) * 8) = pk;
            }
        }
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int rl = warp * 2 + (lane >> 4) + 32 * i, r = mt * 128 + rl;
            if (r < op.rows_act)
                *reinterpret_cast<uint4 *>(A + (rl >> 3) * kASbo + (lane & 15) * kALbo + (rl & 7) * 16) = v[i];
        }
    }
}

// epilogue of one 16-column group of one m-tile: thread = activation row
__device__ __forceinline__ void bepilogue16(const BOp & op, int r, int n0, const float * acc, const float4 * side) {
    const int B = op.n_streams;
    if (op.epi == EPI_SWIGLU) {
        // weight rows interleaved gate_i, up_i (qtts_weights.cu): out[n0/2 + i] = silu(acc[2i]) * acc[2i+1]
        __half2 h[4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
            h[i] = __floats2half2_rn(silu_f(acc[4 * i]) * acc[4 * i + 1], silu_f(acc[4 * i + 2]) * acc[4 * i + 3]);
        uint4 pk;
        pk.x = *reinterpret_cast<const uint32_t *>(&h[0]); pk.y = *reinterpret_cast<const uint32_t *>(&h[1]);
        pk.z = *reinterpret_cast<const uint32_t *>(&h[2]); pk.w = *reinterpret_cast<const uint32_t *>(&h[3]);
        *reinterpret_cast<uint4 *>(reinterpret_cast<unsigned char *>(op.out16) + canon_off(r, n0 >> 1, canon_chunk_bytes(op.rows_act))) = pk;
        return;
    }
    float * out = op.out32 + row_off(r, B, op.out_row_stride, op.out_tok_stride) + n0;
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float4 o = make_float4(acc[4 * i], acc[4 * i + 1], acc[4 * i + 2], acc[4 * i + 3]);
        if (op.epi == EPI_ADD || op.epi == EPI_BIAS) { o.x += side[i].x; o.y += side[i].y; o.z += side[i].z; o.w += side[i].w; }
        *reinterpret_cast<float4 *>(out + 4 * i) = o;
        ss += o.x * o.x; ss += o.y * o.y; ss += o.z * o.z; ss += o.w * o.w;
    }
    if (op.flags & BF_SSQ_OUT) op.ssq_out[(size_t) r * (op.rows >> 4) + (n0 >> 4)] = ss;
}

// MMA issue is spread over kIssuers threads (lane 0 of warps 0..3, one per SM sub-partition): issuing
// one tcgen05.mma costs ~60 cycles of a thread's time while a 128 x 16..96 x 16 MMA occupies the
// tensor pipe for far less, so a single issuer was the bottleneck of a stage (measured: 765 cycles
// per 8-MMA chunk).  Issuer w takes K-steps [w * 8 / n_iss, (w + 1) * 8 / n_iss) of every chunk and
// accumulates into its OWN TMEM accumulator (columns (w * m_tiles + mt) * n_tile ..); the epilogue
// adds the n_iss accumulators in fixed order.  Every issuer commits to the same mbarriers (count
// kIssuers): tcgen05.commit tracks the MMAs of the executing thread only.
constexpr int kIssuers = 4;
__device__ __forceinline__ int n_issuers(int n_tile, int m_tiles) {
    int n = kIssuers;
    while (n > 1 && n * m_tiles * n_tile > 512) n >>= 1;
    return n;
}

__device__ __forceinline__ void bgemm_issue(BShared * S, int w, int n_iss, int kc, int mt, uint32_t a0, uint32_t a_lbo, uint32_t a_sbo,
                                         uint64_t * a_free, int piece, int koff, int kp, uint32_t cons_seq, uint32_t idesc,
                                         int m_tiles, bool last) {
    const BOp & op = S->op;
    tc_fence_after();
    const uint32_t seq = cons_seq + (uint32_t) piece;
    const uint32_t nsl = (uint32_t) bring_slots(S);
    const int slot = (int) (seq % nsl);
    if (mt == 0 && koff == 0) mbar_wait(S->full + slot, (seq / nsl) & 1u);
    if (w < n_iss && !(S->run.dbg & 1)) {
        const uint32_t w0 = smem_u32(bring_base(S) + (size_t) slot * kBSlotBytes) + (uint32_t) (koff >> 3) * 128u;
        const uint32_t d = S->tmem_base + (uint32_t) ((w * m_tiles + mt) * op.n_tile);
        const int per = (kKc / 16) / n_iss;
        const uint64_t ad0 = umma_smem_desc(a0, a_lbo, a_sbo), bd0 = umma_smem_desc(w0, 128, (uint32_t) kp * 16u);
        for (int i = 0; i < per; ++i) {
            const int ks = w * per + i;
            umma_f16_ss(d, ad0 + (uint64_t) ((ks * 2 * a_lbo) >> 4), bd0 + (uint64_t) (ks * 16), idesc, (kc > 0 || i > 0) ? 1u : 0u);
        }
    }
    umma_commit(a_free);
    if (mt == m_tiles - 1 && koff + kKc == kp) umma_commit(S->empty + slot);
    if (last) umma_commit(&S->dbar);
}

// epilogue of tile t: warp w reads TMEM lanes 32*(w%4).. (its sub-partition), 16 columns per pass
template <bool EXT>
__device__ __noinline__ void bgemm_epilogue(BShared * S, int t_virtual, int m_tiles, uint32_t d_use, int n_acc = 0, int mt0 = 0) {
    const BOp & op = S->op;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t tmem_base = S->tmem_base;
    const int dbg = S->run.dbg;
    const int q = warp & 3;
    const int ncg = op.n_tile >> 4;
    // n_acc > 0: accumulators of issuers 0..n_acc-1 are summed; -nab: m-tile mt has one accumulator, issuer (mt % nab)'s
    const int n_iss = n_acc > 0 ? n_acc : (n_acc < 0 ? 1 : n_issuers(op.n_tile, m_tiles));
    // split-K: virtual tile -> (row block t, K-half); half 1 only publishes its partial accumulator
    const int ks = op.ksplit > 1 ? op.ksplit : 1;
    const int ms = op.msplit > 1 ? op.msplit : 1;       // split-M: m_tiles / mt below are local, rows start at M tile mt0
    const int t = (EXT && ms > 1) ? tile_weight(t_virtual, ms, msplit_paired(S->run.cluster, ms, op.n_tiles)) : t_virtual / (ks * ms);
    const int half = (ks > 1) ? t_virtual - t * ks : 0;      // (split-M and split-K are never combined)
    const int rows_pad = (op.rows_act + 31) & ~31;
    const uint32_t tag = S->tag;
    bool waited = false;
    for (int cg = warp >> 2; cg < ncg; cg += 4) {
        const int n0 = t * op.n_tile + cg * 16;
        for (int mt = 0; mt < m_tiles; ++mt) {
            // M = 128: accumulator row i sits in TMEM lane i; M = 64 (rows_act <= 64): row i sits in lane
            // (i % 16) + 32 * (i / 16), i.e. lanes 0..15 of each warp's sub-partition
            const bool m64 = op.rows_act <= 64;
            const int r = m64 ? q * 16 + lane : (mt0 + mt) * 128 + q * 32 + lane;
            const bool valid = r < op.rows_act && (!m64 || lane < 16);
            float4 side[4];
            if (valid && op.epi == EPI_ADD && half == 0) {
                const float * rs = op.res + row_off(r, op.n_streams, op.out_row_stride, op.out_tok_stride) + n0;
#pragma unroll
                for (int i = 0; i < 4; ++i) side[i] = __ldcg(reinterpret_cast<const float4 *>(rs + 4 * i));
            } else if (op.epi == EPI_BIAS) {
#pragma unroll
                for (int i = 0; i < 4; ++i) side[i] = __ldg(reinterpret_cast<const float4 *>(op.bias + n0 + 4 * i));
            }
            if (!waited) { mbar_wait(&S->dbar, d_use & 1u); tc_fence_after(); waited = true; if (threadIdx.x == 32) BTR(6); }
            __syncwarp();
            float acc[16];
            if (!(dbg & 2)) {
                // issuers whose accumulators hold m-tile mt: n_acc > 0: 0 .. n_acc-1; n_acc = -nab: step s = kc * m_tiles + mt lands in
                // buffer (= issuer) s % nab, i.e. issuer mt % nab alone when m_tiles is a multiple of nab, issuers mt, mt + m_tiles,
                // ... < nab when nab is a multiple of m_tiles.  Fixed order: results do not depend on timing.
                int wa = 0, wstep = 1, wn = n_iss;
                if (n_acc < 0) {
                    const int nab = -n_acc;
                    if (!EXT || m_tiles >= nab) { wa = mt & (nab - 1); wn = 1; }
                    else { wa = mt; wstep = m_tiles; wn = nab / m_tiles; }
                }
                tmem_ld16(tmem_base + ((uint32_t) (q * 32) << 16) + (uint32_t) ((wa * m_tiles + mt) * op.n_tile + cg * 16), acc);
                for (int j = 1; j < wn; ++j) {
                    const int w = wa + j * wstep;
                    float more[16];
                    tmem_ld16(tmem_base + ((uint32_t) (q * 32) << 16) + (uint32_t) ((w * m_tiles + mt) * op.n_tile + cg * 16), more);
#pragma unroll
                    for (int i = 0; i < 16; ++i) acc[i] += more[i];
                }
            } else {
#pragma unroll
                for (int i = 0; i < 16; ++i) acc[i] = 0.f;
            }
            if (ks > 1 && valid && op.part32) {
                if (half == 1) {        // deferred combine: the following norm stage adds this partial
                    float * pd = op.part32 + row_off(r, op.n_streams, op.out_row_stride, op.out_tok_stride) + n0;
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        *reinterpret_cast<float4 *>(pd + 4 * i) = make_float4(acc[4 * i], acc[4 * i + 1], acc[4 * i + 2], acc[4 * i + 3]);
                    continue;
                }
            } else if (ks > 1 && valid) {
                unsigned long long * pp = op.part_ll + (((size_t) t * ncg + cg) * rows_pad + r) * 16;
                if (half == 1) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const unsigned long long w = ((unsigned long long) tag << 32) | (unsigned long long) __float_as_uint(acc[i]);
                        asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" :: "l"(pp + i), "l"(w) : "memory");
                    }
                    continue;
                }
                // half 0: all 16 words requested together, repeated until every tag matches
                uint32_t spins = 0;
                bool ok;
                do {
                    unsigned long long w[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(w[i]) : "l"(pp + i) : "memory");
                    ok = true;
#pragma unroll
                    for (int i = 0; i < 16; ++i) ok = ok && ((uint32_t) (w[i] >> 32) == tag);
                    if (ok) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) acc[i] += __uint_as_float((uint32_t) w[i]);
                    } else if (++spins > (1u << 22)) __trap();
                } while (!ok);
            }
            if (valid) bepilogue16(op, r, n0, acc, side);
        }
    }
}

template <int IN_KIND, bool EXT>
__device__ __noinline__ void bstage_gemm(BShared * S) {
    const BOp & op = S->op;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int G = (int) gridDim.x;
    const int t_first = first_tile(op.cta_off);
    if (t_first < op.n_tiles) {
        const int m_tiles = (op.rows_act + 127) >> 7;
        const int nkc = op.K / kKc;
        const int steps = nkc * m_tiles;
        const int upc = (IN_KIND == BIN_F32 && op.rows_act > 64) ? 2 : 1;     // load units per chunk
        const int n_iss = n_issuers(op.n_tile, m_tiles);
        const int total = steps * upc;
        const uint32_t idesc = (op.rows_act <= 64 ? (S->run.idesc_base & ~(0x1Fu << 24)) | (4u << 24) : S->run.idesc_base) | ((uint32_t) (op.n_tile >> 3) << 17);
        const bool write_hid = (op.hid_out != nullptr) && (t_first == 0);
        const uint32_t tmem_base = S->tmem_base;
        uint32_t a_use0 = S->a_use[0], a_use1 = S->a_use[1], d_use = S->d_use, cons_seq = S->cons_seq;
        unsigned char * abuf = babuf_base(S);
        unsigned char * ring = bring_base(S);
        const int dbg = S->run.dbg;

        // per-row RMSNorm scale from the producers' partial sums (double sum like ggml's rms_norm,
        // ops.cpp:3738-3747)
        if (IN_KIND == BIN_F32 && op.norm_w) {
            // 4 threads per row, all partial loads of a thread in flight together (one L2 round trip per
            // 128 partials instead of one per 8: this prologue was a quarter of the stage)
            const int np = op.n_parts_in, sub = tid & 3;
            for (int r0 = 0; r0 < op.rows_act; r0 += kThreads / 4) {
                const int r = r0 + (tid >> 2);
                double ss = 0.0;
                if (r < op.rows_act) {
                    const float * p = op.ssq_in + (size_t) r * np;
                    for (int base = 0; base < np; base += 128) {
                        float q[32];
#pragma unroll
                        for (int u = 0; u < 32; ++u) { const int i = base + sub + 4 * u; q[u] = (i < np) ? __ldcg(p + i) : 0.f; }
#pragma unroll
                        for (int u = 0; u < 32; ++u) ss += (double) q[u];
                    }
                }
                ss += __shfl_xor_sync(0xffffffffu, ss, 1);
                ss += __shfl_xor_sync(0xffffffffu, ss, 2);
                if (sub == 0 && r < op.rows_act) {
                    const float mean = (float) (ss / (double) op.K);
                    S->scale[r] = 1.0f / sqrtf(mean + op.eps);
                }
            }
            __syncthreads();
        }

        for (int t = t_first; t < op.n_tiles; t += G) {
            // two load units in flight (two named register sets, loop unrolled by two): the loads of
            // unit j+2 are issued as soon as unit j went to shared memory
            uint4 v0[4], v1[4];
            float4 nw0 = make_float4(1.f, 1.f, 1.f, 1.f), nw1 = nw0;
#pragma unroll
            for (int i = 0; i < 4; ++i) { v0[i] = make_uint4(0u, 0u, 0u, 0u); v1[i] = v0[i]; }
            auto load_unit = [&](int j, uint4 * v, float4 & nw) {
                const int s = j / upc, u = j - s * upc;
                const int kc = s / m_tiles, mt = s - kc * m_tiles;
                bload<IN_KIND>(op, kc, mt, u, v, nw);
            };
            // a load unit goes to shared memory (after the MMAs that read this buffer two chunks ago are done)
            auto stage_unit = [&](int j, const uint4 * v, const float4 & nw) {
                const int s = j / upc, u = j - s * upc;
                const int kc = s / m_tiles, mt = s - kc * m_tiles;
                const int buf = s & 1;
                if (u == 0) {
                    const uint32_t au = buf ? a_use1 : a_use0;
                    if (au > 0) mbar_wait(S->abar + buf, (au - 1) & 1u);
                }
                bstore<IN_KIND>(op, S, abuf + (size_t) buf * kABufBytes, kc, mt, u, v, nw, write_hid);
                if (u == upc - 1) {
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    __syncthreads();
                }
            };
            // ... and its MMAs are issued AFTER the loads of the unit after next were requested, so the
            // issuers' ~300 cycles per chunk overlap with the load latency instead of delaying it
            auto issue_unit = [&](int j) {
                const int s = j / upc, u = j - s * upc;
                if (u != upc - 1) return;
                const int kc = s / m_tiles, mt = s - kc * m_tiles;
                const int buf = s & 1;
                const int kin = kc * kKc;
                const int piece = kin / op.kpp, koff = kin - piece * op.kpp;
                const int kp = min(op.kpp, op.K - piece * op.kpp);
                if (lane == 0 && warp < kIssuers) {
                    bgemm_issue(S, warp, n_iss, kc, mt, smem_u32(abuf + (size_t) buf * kABufBytes), kALbo, kASbo, S->abar + buf,
                                piece, koff, kp, cons_seq, idesc, m_tiles, s == steps - 1);
                } else if (tid == kPfThread) {
                    // the piece the NEXT chunk needs must be in flight before the issuers wait for it
                    const int kin1 = min(kin + kKc, op.K - kKc);
                    bpf_ensure(S, cons_seq + (uint32_t) (kin1 / op.kpp));
                    bpf_poll(S);
                }
                if (buf) a_use1++; else a_use0++;
            };
            load_unit(0, v0, nw0);
            if (total > 1) load_unit(1, v1, nw1);
            if (tid == kPfThread) bpf_ensure(S, cons_seq);
            for (int j = 0; j < total; j += 2) {
                stage_unit(j, v0, nw0);
                if (j + 2 < total) load_unit(j + 2, v0, nw0);
                issue_unit(j);
                if (j + 1 < total) {
                    stage_unit(j + 1, v1, nw1);
                    if (j + 3 < total) load_unit(j + 3, v1, nw1);
                    issue_unit(j + 1);
                }
            }
            bgemm_epilogue<EXT>(S, t, m_tiles, d_use);
            d_use++;
            tc_fence_before();
            __syncthreads();
            cons_seq += (uint32_t) op.n_pieces;
        }
        if (tid == 0) { S->a_use[0] = a_use0; S->a_use[1] = a_use1; S->d_use = d_use; S->cons_seq = cons_seq; }
    }
    if ((op.flags & BF_INC_NPAST) && blockIdx.x == 0 && tid < op.n_streams) {
        StreamState * st = S->run.st + tid;
        if (S->run.active[tid] && !st->done) st->n_past += 1;
    }
}

// ------------------------------------------------------------------------------------------
// BOP_GEMM with F16 input (O and down projections): the producing stage wrote the activations to
// global memory already in UMMA-canonical 128-column chunks, so the whole stage is asynchronous:
// thread kAProd bulk-copies (row block, chunk) pieces L2 -> shared memory (2 or 4 buffers), thread 0
// waits on the pieces' mbarriers and issues the MMAs, tcgen05.commit recycles the buffers; the other
// threads only run the epilogue.  No per-chunk CTA barrier, no register staging.
// ------------------------------------------------------------------------------------------
constexpr int kAProd = 4 * 32;      // lane 0 of warp 4 (warps 0..3 hold the MMA issuers)

// EXT = false: the configuration of groups of <= 64 rows without clusters (two activation buffers, kBSlots ring slots) with
// every choice a compile-time constant - the issue path of this stage is latency-critical to the point where a run-time
// slot count and two shared-memory flag reads per chunk cost 7 % of a 64-stream frame-step (profiles/r02_summary.md
// section 14).  EXT = true: four buffers / 3 slots and / or cluster multicast, chosen per launch.
template <bool EXT>
__device__ __noinline__ void bstage_gemm_tma(BShared * S) {
    const BOp & op = S->op;
    const int tid = threadIdx.x;
    const int G = (int) gridDim.x;
    const int t_first = first_tile(op.cta_off);
    if (t_first < op.n_tiles) {
        const int msplit = op.msplit > 1 ? op.msplit : 1;
        const int m_tiles = ((op.rows_act + 127) >> 7) / msplit;     // M tiles of ONE work item
        const int nkc = op.K / kKc;
        const int steps = nkc * m_tiles;
        const int Rg = (op.rows_act + 7) >> 3, cb = Rg << 11;
        // buffers in use: 4 halves of the two buffers for <= 64 rows; 4 full buffers when the launch has them (a_bufs = 4) and
        // the accumulators of 4 issuers fit in TMEM; else the two full buffers
        const bool four_full = EXT && QTTS_BRING_RUNTIME && Rg > 8 && S->run.a_bufs == 4 && m_tiles <= 2 && 4 * m_tiles * op.n_tile <= 512;
        const int nab = (Rg <= 8 || four_full) ? 4 : 2, bstep = 4 / nab;          // buffers in use, barrier index step
        const int lane = tid & 31, warp = tid >> 5;
        const int abstride = (Rg <= 8) ? kABufBytes / 2 : kABufBytes;
        // accumulators: issuer w (< nab) writes (w, mt) for the steps that land in its buffer; with two
        // buffers and two m-tiles step kc * 2 + mt lands in buffer mt, i.e. m-tile mt has ONE accumulator, issuer mt's
        // (generally: with m_tiles a multiple of the buffer count, m-tile mt always lands in buffer mt % nab)
        // (with 4 buffers and 2 m-tiles: m-tile mt has two accumulators, issuers mt and mt + 2 - the epilogue's -nab rule)
        const int acc_mode = ((m_tiles >= nab && m_tiles % nab == 0) || (EXT && m_tiles > 1 && nab % m_tiles == 0)) ? -nab : min(nab, steps);
        const uint32_t tmem_base = S->tmem_base;
        const uint32_t idesc = (op.rows_act <= 64 ? (S->run.idesc_base & ~(0x1Fu << 24)) | (4u << 24) : S->run.idesc_base) | ((uint32_t) (op.n_tile >> 3) << 17);
        uint32_t d_use = S->d_use, cons_seq = S->cons_seq;
        unsigned char * abuf = babuf_base(S);
        const bool paired = EXT && msplit_paired(S->run.cluster, msplit, op.n_tiles);
        // cluster mode: the other CTA of the pair, and whether it walks the same (K range, M tiles) as this one in a
        // given round - then the two share every activation piece (each copies half of it, multicast to both)
        const int t_peer_first = (EXT && S->run.cluster) ? first_tile_of((int) (blockIdx.x ^ 1u), op.cta_off) : op.n_tiles;
        const uint32_t crank = blockIdx.x & 1u;
        for (int t = t_first; t < op.n_tiles; t += G) {
            // split-K: virtual tile t reads the activation chunks of its K-half
            const int kc_base = (op.ksplit > 1) ? (t % op.ksplit) * nkc : 0;
            const int mt0 = tile_half(t, msplit, paired) * m_tiles;   // split-M: first M tile of this work item
            const int t_peer = t_peer_first + (t - t_first);
            const bool shared_a = EXT && QTTS_BCLUSTER_CODE && S->run.cluster && !(S->run.dbg & 4) && t_peer < op.n_tiles &&
                                  ((op.ksplit > 1) ? (t_peer % op.ksplit) * nkc : 0) == kc_base &&
                                  tile_half(t_peer, msplit, paired) * m_tiles == mt0;
            if (tid == kAProd) {
                uint64_t pol;
                asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
                if (op.flags & BF_ROWS_WAIT) {
                    // the rows come from a BOP_NORM stage that ended without a grid barrier: wait for its row counter
                    const uint32_t target = S->rows_target;
                    const unsigned int * ctr = &S->run.bar->pad[15];
                    unsigned int v;
                    uint32_t spins = 0;
                    do {
                        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
                        if (++spins > (1u << 24)) __trap();
                    } while ((int) (v - target) < 0);
                }
                asm volatile("fence.proxy.async.global;" ::: "memory");
                if (t == t_first) BTR(2);
                const unsigned char * in = reinterpret_cast<const unsigned char *>(op.in);
                const int pdbg = S->run.dbg;                 // (read once: this loop decides how early the first pieces are in flight)
                int kc = 0, mt = 0;
                for (int s = 0; s < steps; ++s, ++mt) {
                    if (mt == m_tiles) { mt = 0; ++kc; }
                    const int buf = s & (nab - 1), bi = buf * bstep;
                    const uint32_t nf = S->n_fill[bi];
                    mbar_wait(S->tempty + bi, (nf & 1u) ^ 1u);
                    const uint32_t bytes = (uint32_t) min(16, Rg - 16 * (mt0 + mt)) << 11;
                    if (pdbg & 4) {      // bring-up: pipeline overhead without the copy
                        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(S->tfull + bi)) : "memory");
                    } else {
                        const int kcs = kc_base + ((pdbg & 8) ? (kc + (int) blockIdx.x) % nkc : kc);   // (dbg 8: bring-up, staggered chunk order)
                        mbar_expect_tx(S->tfull + bi, bytes);
                        if (shared_a) {
                            // this CTA brings its half of the piece for both CTAs; the other half arrives from the peer
                            const uint32_t hb = bytes >> 1, ho = crank * hb;
                            bulk_g2s_mc(abuf + (size_t) buf * abstride + ho, in + (size_t) kcs * cb + (size_t) (mt0 + mt) * 32768 + ho, hb,
                                        S->tfull + bi, (uint16_t) 3, pol);
                        } else {
                            bulk_g2s(abuf + (size_t) buf * abstride, in + (size_t) kcs * cb + (size_t) (mt0 + mt) * 32768, bytes, S->tfull + bi, pol);
                        }
                    }
                    S->n_fill[bi] = nf + 1;
                }
                if (t == t_first) BTR(3);
            } else if (lane == 0 && warp < kIssuers) {
                // Chunk-alternating issue: issuer w takes the K-chunks kc = w, w + n_alt, ... and issues all
                // 8 K-steps of each into its OWN accumulators (summed by the epilogue).  A tcgen05.mma costs
                // its issuing thread ~60 cycles and a commit ~200; with the K-steps of every chunk spread
                // over four issuers each of them paid wait + fence + commit per chunk for two MMAs (measured
                // 1290 cycles per chunk); alternating whole chunks keeps the tensor pipe fed instead.
                const bool pr = S->run.prof != nullptr && warp == 0 && t == 0;     // stage timers (tile 0's CTA)
                long long c_wait = 0, c_issue = 0, c_first = 0;
                const long long c0 = pr ? clock64() : 0;
                const int w = warp;
                const uint32_t ring0 = EXT ? smem_u32(bring_base(S)) : smem_u32(babuf_base(S)) + 2u * (uint32_t) kABufBytes;
                const uint32_t nsl = EXT ? (uint32_t) bring_slots(S) : (uint32_t) kBSlots;
                const uint32_t abuf0 = smem_u32(abuf);
                const bool skip_mma = (S->run.dbg & 1) != 0;
                const int nab_shift = (nab == 4) ? 2 : 1;
                uint32_t started = 0;       // accumulators (w, mt) this issuer has written
                const uint32_t my_fill_base = S->fill_base[(w & (nab - 1)) * bstep];      // issuer w only ever waits on buffer w
                int kc = 0;
                for (int piece = 0; piece < op.n_pieces; ++piece) {
                    const int kp = min(op.kpp, op.K - piece * op.kpp);
                    const int chunks = kp / kKc;
                    const uint32_t seq = cons_seq + (uint32_t) piece;
                    const int slot = (int) (seq % nsl);
                    mbar_wait(S->full + slot, (seq / nsl) & 1u);     // every issuer: keeps the slot phases aligned
                    if (warp == 0 && piece == 0 && t == t_first) BTR(11);
                    const uint32_t w0 = ring0 + (uint32_t) slot * kBSlotBytes;
                    bool mine = false;
                    for (int c = 0; c < chunks; ++c, ++kc) {
                        for (int mt = 0; mt < m_tiles; ++mt) {
                            const int s = kc * m_tiles + mt;
                            const int buf = s & (nab - 1), bi = buf * bstep;
                            if (buf != w) continue;             // issuer w owns A buffer w: its fills are consumed in order by one thread
                            mine = true;
                            const uint32_t fill = my_fill_base + (uint32_t) (s >> nab_shift);
                            const long long ca = pr ? clock64() : 0;
                            mbar_wait(S->tfull + bi, fill & 1u);
                            if (warp == 0 && s == 0 && t == t_first) BTR(4);
                            const long long cb2 = pr ? clock64() : 0;
                            tc_fence_after();
                            if (!skip_mma) {
                                const uint32_t d = tmem_base + (uint32_t) ((w * m_tiles + mt) * op.n_tile);
                                const uint64_t ad0 = umma_smem_desc(abuf0 + (uint32_t) buf * (uint32_t) abstride, 128, 2048);
                                const uint64_t bd0 = umma_smem_desc(w0 + (uint32_t) c * 2048u, 128, (uint32_t) kp * 16u);
                                const bool fresh = !((started >> mt) & 1u);
                                started |= 1u << mt;
#pragma unroll
                                for (int i = 0; i < kKc / 16; ++i)
                                    umma_f16_ss(d, ad0 + (uint64_t) (i * 16), bd0 + (uint64_t) (i * 16), idesc, (!fresh || i > 0) ? 1u : 0u);
                            }
                            // buffer recycling: in cluster mode tempty counts two arrivals per fill - both CTAs of the pair must
                            // have consumed the buffer before either half of the next piece is multicast into it
                            if (shared_a) umma_commit_mc(S->tempty + bi, (uint16_t) 3);
                            else {
                                umma_commit(S->tempty + bi);
                                if (EXT && QTTS_BCLUSTER_CODE && S->run.cluster) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(S->tempty + bi)) : "memory");
                            }
                            if (pr) { const long long cc = clock64(); c_wait += cb2 - ca; c_issue += cc - cb2; if (s == 0) c_first = cb2 - c0; }
                        }
                    }
                    // release the weight slot: after my MMAs that read it, or at once when I had none in this piece
                    if (mine) umma_commit(S->empty + slot);
                    else asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(S->empty + slot)) : "memory");
                }
                umma_commit(&S->dbar);
                if (warp == 0 && t == t_first) BTR(5);
                if (pr) {
                    unsigned long long * P = S->run.prof;
                    P[40] += 1; P[41] += (unsigned long long) c_first; P[42] += (unsigned long long) c_wait;
                    P[43] += (unsigned long long) c_issue; P[44] += (unsigned long long) (clock64() - c0);
                }
            } else if (tid == kPfThread) {
                bpf_ensure(S, cons_seq + (uint32_t) (op.n_pieces - 1));
                bpf_poll(S);
            }
            __syncwarp();
            const bool pr1 = S->run.prof != nullptr && tid == 32 && t == 0;      // a thread without issue duties
            const long long e0 = pr1 ? clock64() : 0;
            bgemm_epilogue<EXT>(S, t, m_tiles, d_use, acc_mode, mt0);
            if (tid == 32 && t == t_first) BTR(7);
            const long long e1 = pr1 ? clock64() : 0;
            d_use++;
            tc_fence_before();
            __syncthreads();
            if (pr1) { S->run.prof[45] += (unsigned long long) (e1 - e0); S->run.prof[46] += (unsigned long long) (clock64() - e1); }
            if (tid < 4) S->fill_base[tid] = S->n_fill[tid];      // all fills of this tile were issued and consumed
            if (t + G < op.n_tiles) __syncthreads();
            cons_seq += (uint32_t) op.n_pieces;
        }
        if (tid == 0) { S->d_use = d_use; S->cons_seq = cons_seq; }
    }
    if ((op.flags & BF_INC_NPAST) && blockIdx.x == 0 && tid < op.n_streams) {
        StreamState * st = S->run.st + tid;
        if (S->run.active[tid] && !st->done) st->n_past += 1;
    }
}

// ------------------------------------------------------------------------------------------
// BOP_NORM: x[r][:] (F32) -> f16( (x * scale_r) * w ) in the canonical chunk layout.  Doing this
// once per row instead of inside every consumer CTA removes the dominant cost of the normalised
// GEMM stages (every CTA re-read all rows as F32 from L2 and converted them: 38 MB of L2 traffic
// per stage at 64 streams) and lets those stages run on the bulk-copy path.
// Same arithmetic and rounding point as the in-stage prologue (ggml rms_norm, ops.cpp:3709-3758).
// ------------------------------------------------------------------------------------------
__device__ __noinline__ void bstage_norm(BShared * S) {
    const BOp & op = S->op;
    const int tid = threadIdx.x, K = op.K, np = op.n_parts_in;
    const int cb = canon_chunk_bytes(op.rows_act);
    const bool signal = (op.flags & BF_ROWS_SIGNAL) != 0;
    if (signal && tid == 0) S->rows_target += (uint32_t) op.rows_act;       // every CTA keeps the same running target
    for (int r = blockIdx.x; r < op.rows_act; r += (int) gridDim.x) {
        const float * x = reinterpret_cast<const float *>(op.in) + row_off(r, op.n_streams, op.in_row_stride, op.in_tok_stride);
        // everything this thread needs is requested before the reduction
        const int k = tid * 8;
        float4 xa = make_float4(0.f, 0.f, 0.f, 0.f), xb = xa, wa = xa, wb = xa;
        if (k < K) {
            xa = __ldcg(reinterpret_cast<const float4 *>(x + k)); xb = __ldcg(reinterpret_cast<const float4 *>(x + k + 4));
            wa = __ldg(reinterpret_cast<const float4 *>(op.norm_w + k)); wb = __ldg(reinterpret_cast<const float4 *>(op.norm_w + k + 4));
        }
        double ss = 0.0;
        if (op.part32) {
            // the producing projection was split over K: x = x + (second half's partial), in this fixed order;
            // the row is written back (it is the residual of the next projection) and squared here
            const float * pr = op.part32 + (x - reinterpret_cast<const float *>(op.in));
            float * xw = const_cast<float *>(x);
            for (int kk = k; kk < K; kk += kThreads * 8) {
                float4 a = xa, b = xb;
                if (kk != k) { a = __ldcg(reinterpret_cast<const float4 *>(x + kk)); b = __ldcg(reinterpret_cast<const float4 *>(x + kk + 4)); }
                const float4 pa = __ldcg(reinterpret_cast<const float4 *>(pr + kk)), pb = __ldcg(reinterpret_cast<const float4 *>(pr + kk + 4));
                a.x += pa.x; a.y += pa.y; a.z += pa.z; a.w += pa.w; b.x += pb.x; b.y += pb.y; b.z += pb.z; b.w += pb.w;
                *reinterpret_cast<float4 *>(xw + kk) = a; *reinterpret_cast<float4 *>(xw + kk + 4) = b;
                if (kk == k) { xa = a; xb = b; }
                ss += (double) (a.x * a.x); ss += (double) (a.y * a.y); ss += (double) (a.z * a.z); ss += (double) (a.w * a.w);
                ss += (double) (b.x * b.x); ss += (double) (b.y * b.y); ss += (double) (b.z * b.z); ss += (double) (b.w * b.w);
            }
        } else {
            for (int i = tid; i < np; i += kThreads) ss += (double) __ldcg(op.ssq_in + (size_t) r * np + i);
        }
        ss = block_sum_d(ss, S->red);
        const float mean = (float) (ss / (double) K);
        const float sc = 1.0f / sqrtf(mean + op.eps);
        for (int kk = k; kk < K; kk += kThreads * 8) {
            if (kk != k) {
                xa = __ldcg(reinterpret_cast<const float4 *>(x + kk)); xb = __ldcg(reinterpret_cast<const float4 *>(x + kk + 4));
                wa = __ldg(reinterpret_cast<const float4 *>(op.norm_w + kk)); wb = __ldg(reinterpret_cast<const float4 *>(op.norm_w + kk + 4));
            }
            const float v0 = (xa.x * sc) * wa.x, v1 = (xa.y * sc) * wa.y, v2 = (xa.z * sc) * wa.z, v3 = (xa.w * sc) * wa.w;
            const float v4 = (xb.x * sc) * wb.x, v5 = (xb.y * sc) * wb.y, v6 = (xb.z * sc) * wb.z, v7 = (xb.w * sc) * wb.w;
            if (op.hid_out) {
                *reinterpret_cast<float4 *>(op.hid_out + (size_t) r * K + kk) = make_float4(v0, v1, v2, v3);
                *reinterpret_cast<float4 *>(op.hid_out + (size_t) r * K + kk + 4) = make_float4(v4, v5, v6, v7);
            }
            const __half2 h0 = __floats2half2_rn(v0, v1), h1 = __floats2half2_rn(v2, v3);
            const __half2 h2 = __floats2half2_rn(v4, v5), h3 = __floats2half2_rn(v6, v7);
            uint4 pk;
            pk.x = *reinterpret_cast<const uint32_t *>(&h0); pk.y = *reinterpret_cast<const uint32_t *>(&h1);
            pk.z = *reinterpret_cast<const uint32_t *>(&h2); pk.w = *reinterpret_cast<const uint32_t *>(&h3);
            *reinterpret_cast<uint4 *>(reinterpret_cast<unsigned char *>(op.out16) + canon_off(r, kk, cb)) = pk;
        }
        if (signal) asm volatile("fence.proxy.async.global;" ::: "memory");    // the row is read by bulk copies (async proxy)
        __syncthreads();        // S->red is reused by the next row of this CTA
        if (signal && tid == 0) {
            __threadfence();
            atomicAdd(&S->run.bar->pad[15], 1u);
        }
    }
}

// BOP_NORM at warp granularity: one WARP per row instead of one CTA per row, rows dealt over CTAs first (row r -> CTA
// r % grid, warp r / grid), so every row of a stage has its own warp somewhere and no CTA barrier or block reduction is on
// the path: 256 rows were two passes of ~2 us per CTA, now one warp-latency.  Same arithmetic and rounding point; the
// double-precision sum runs lane-strided + xor tree instead of warp-partials in warp order (both fixed orders).
__device__ __noinline__ void bstage_norm_warp(BShared * S) {
    const BOp & op = S->op;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, K = op.K, np = op.n_parts_in;
    const int cb = canon_chunk_bytes(op.rows_act);
    const bool signal = (op.flags & BF_ROWS_SIGNAL) != 0;
    const int G = (int) gridDim.x;
    if (signal && tid == 0) S->rows_target += (uint32_t) op.rows_act;       // every CTA keeps the same running target
    for (int r = (int) blockIdx.x + G * warp; r < op.rows_act; r += G * kWarps) {
        const float * x = reinterpret_cast<const float *>(op.in) + row_off(r, op.n_streams, op.in_row_stride, op.in_tok_stride);
        double ss = 0.0;
        if (op.part32) {
            // the producing projection was split over K: x = x + (second half's partial), in this fixed order; the row is
            // written back (it is the residual of the next projection) and squared here
            const float * pr = op.part32 + (x - reinterpret_cast<const float *>(op.in));
            float * xw = const_cast<float *>(x);
            for (int kk = lane * 8; kk < K; kk += 256) {
                float4 a = __ldcg(reinterpret_cast<const float4 *>(x + kk)), b = __ldcg(reinterpret_cast<const float4 *>(x + kk + 4));
                const float4 pa = __ldcg(reinterpret_cast<const float4 *>(pr + kk)), pb = __ldcg(reinterpret_cast<const float4 *>(pr + kk + 4));
                a.x += pa.x; a.y += pa.y; a.z += pa.z; a.w += pa.w; b.x += pb.x; b.y += pb.y; b.z += pb.z; b.w += pb.w;
                *reinterpret_cast<float4 *>(xw + kk) = a; *reinterpret_cast<float4 *>(xw + kk + 4) = b;
                ss += (double) (a.x * a.x); ss += (double) (a.y * a.y); ss += (double) (a.z * a.z); ss += (double) (a.w * a.w);
                ss += (double) (b.x * b.x); ss += (double) (b.y * b.y); ss += (double) (b.z * b.z); ss += (double) (b.w * b.w);
            }
        } else {
            for (int i = lane; i < np; i += 32) ss += (double) __ldcg(op.ssq_in + (size_t) r * np + i);
        }
        ss = warp_sum_d(ss);
        const float mean = (float) (ss / (double) K);
        const float sc = 1.0f / sqrtf(mean + op.eps);
        for (int kk = lane * 8; kk < K; kk += 256) {
            // (after the split-K combine above these are this lane's own writes)
            const float4 xa = __ldcg(reinterpret_cast<const float4 *>(x + kk)), xb = __ldcg(reinterpret_cast<const float4 *>(x + kk + 4));
            const float4 wa = __ldg(reinterpret_cast<const float4 *>(op.norm_w + kk)), wb = __ldg(reinterpret_cast<const float4 *>(op.norm_w + kk + 4));
            const float v0 = (xa.x * sc) * wa.x, v1 = (xa.y * sc) * wa.y, v2 = (xa.z * sc) * wa.z, v3 = (xa.w * sc) * wa.w;
            const float v4 = (xb.x * sc) * wb.x, v5 = (xb.y * sc) * wb.y, v6 = (xb.z * sc) * wb.z, v7 = (xb.w * sc) * wb.w;
            if (op.hid_out) {
                *reinterpret_cast<float4 *>(op.hid_out + (size_t) r * K + kk) = make_float4(v0, v1, v2, v3);
                *reinterpret_cast<float4 *>(op.hid_out + (size_t) r * K + kk + 4) = make_float4(v4, v5, v6, v7);
            }
            const __half2 h0 = __floats2half2_rn(v0, v1), h1 = __floats2half2_rn(v2, v3);
            const __half2 h2 = __floats2half2_rn(v4, v5), h3 = __floats2half2_rn(v6, v7);
            uint4 pk;
            pk.x = *reinterpret_cast<const uint32_t *>(&h0); pk.y = *reinterpret_cast<const uint32_t *>(&h1);
            pk.z = *reinterpret_cast<const uint32_t *>(&h2); pk.w = *reinterpret_cast<const uint32_t *>(&h3);
            *reinterpret_cast<uint4 *>(reinterpret_cast<unsigned char *>(op.out16) + canon_off(r, kk, cb)) = pk;
        }
        if (signal) {
            asm volatile("fence.proxy.async.global;" ::: "memory");    // the row is read by bulk copies (async proxy)
            __syncwarp();
            if (lane == 0) {
                __threadfence();
                atomicAdd(&S->run.bar->pad[15], 1u);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// BOP_ATTN_CP — one warp per (stream, KV group); lane holds dims lane + 32 j (j < 4), so the NEOX
// pairs (i, i + 64) sit in one lane.  pair_f32: positions 0 and 1 in one pass with the all-F32
// arithmetic of the reference's 2-token prefill graph (src/tts_transformer.cpp:1613-1626);
// otherwise the step graph's F16 cache / F32 softmax / F16-rounded P (:1666-1827).
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void warp_norm_rope(float * x, const float * w, const float * rope, int pos, float eps, int lane, bool do_norm) {
    if (do_norm) {
        double ss = 0.0;
#pragma unroll
        for (int j = 0; j < 4; ++j) ss += (double) (x[j] * x[j]);
        ss = warp_sum_d(ss);
        const float mean = (float) (ss / (double) kHeadDim);
        const float scale = 1.0f / sqrtf(mean + eps);
#pragma unroll
        for (int j = 0; j < 4; ++j) x[j] = (x[j] * scale) * __ldg(w + lane + 32 * j);
    }
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        const int i = lane + 32 * j;
        const float c = __ldg(rope + ((size_t) pos * 64 + i) * 2), s = __ldg(rope + ((size_t) pos * 64 + i) * 2 + 1);
        const float x0 = x[j], x1 = x[j + 2];
        x[j] = x0 * c - x1 * s;
        x[j + 2] = x0 * s + x1 * c;
    }
}

// same arithmetic with the norm weights and the RoPE factors already in registers
__device__ __forceinline__ void warp_norm_rope_pre(float * x, const float * w, const float * rc, const float * rs, float eps, bool do_norm) {
    if (do_norm) {
        double ss = 0.0;
#pragma unroll
        for (int j = 0; j < 4; ++j) ss += (double) (x[j] * x[j]);
        ss = warp_sum_d(ss);
        const float mean = (float) (ss / (double) kHeadDim);
        const float scale = 1.0f / sqrtf(mean + eps);
#pragma unroll
        for (int j = 0; j < 4; ++j) x[j] = (x[j] * scale) * w[j];
    }
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        const float x0 = x[j], x1 = x[j + 2];
        x[j] = x0 * rc[j] - x1 * rs[j];
        x[j + 2] = x0 * rs[j] + x1 * rc[j];
    }
}

template <bool PAIR>
__device__ __noinline__ void bstage_attn_cp(BShared * S) {
    const BOp & op = S->op;
    const int lane = threadIdx.x & 31;
    const int gw = blockIdx.x * kWarps + (threadIdx.x >> 5), nw = gridDim.x * kWarps;
    const int B = op.n_streams, nkv = op.n_kv_heads, rep = op.n_heads / nkv;
    const int kvw = nkv * kHeadDim;
    // per-warp scratch in the (idle) A-buffer area: 16 scores + one query head (F16-rounded)
    float * scw = reinterpret_cast<float *>(babuf_base(S)) + (threadIdx.x >> 5) * (16 + kHeadDim);
    float * qs = scw + 16;
    // single-token pass: position, norm weights and RoPE factors are the same for every item of the stage
    float wq[4], wk[4], rc[2], rs[2];
    if (!PAIR) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            wq[j] = op.q_norm_w ? __ldg(op.q_norm_w + lane + 32 * j) : 1.f;
            wk[j] = op.k_norm_w ? __ldg(op.k_norm_w + lane + 32 * j) : 1.f;
        }
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const float2 cs = __ldg(reinterpret_cast<const float2 *>(op.rope) + (size_t) op.pos * 64 + lane + 32 * j);
            rc[j] = cs.x; rs[j] = cs.y;
        }
    }
    for (int item = gw; item < B * nkv; item += nw) {
        const int b = item / nkv, g = item - b * nkv;
        __half * kc = op.kc + (size_t) b * op.kv_stream_stride;
        __half * vc = op.vc + (size_t) b * op.kv_stream_stride;
        float k0f[4], v0f[4];
        for (int t = 0; t < op.n_tok; ++t) {
            const int pos = op.pos + t;
            const float * base = op.qkv + (size_t) (t * B + b) * op.qkvw;
            float k[4], v[4], qn[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                k[j] = __ldcg(base + (size_t) (op.n_heads + g) * kHeadDim + lane + 32 * j);
                v[j] = __ldcg(base + (size_t) (op.n_heads + nkv + g) * kHeadDim + lane + 32 * j);
                qn[j] = __ldcg(base + (size_t) (g * rep) * kHeadDim + lane + 32 * j);       // first query head of the group
            }
            if (PAIR) warp_norm_rope(k, op.k_norm_w, op.rope, pos, op.eps, lane, op.k_norm_w != nullptr);
            else      warp_norm_rope_pre(k, wk, rc, rs, op.eps, op.k_norm_w != nullptr);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                kc[(size_t) pos * kvw + g * kHeadDim + lane + 32 * j] = __float2half_rn(k[j]);
                vc[(size_t) pos * kvw + g * kHeadDim + lane + 32 * j] = __float2half_rn(v[j]);
            }
            if (PAIR && t == 0) {
#pragma unroll
                for (int j = 0; j < 4; ++j) { k0f[j] = k[j]; v0f[j] = v[j]; }
            }
            __syncwarp();
            for (int r = 0; r < rep; ++r) {
                const int h = g * rep + r;
                float q[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) q[j] = qn[j];
                if (r + 1 < rep) {      // the next head's query is requested while this one is processed
#pragma unroll
                    for (int j = 0; j < 4; ++j) qn[j] = __ldcg(base + (size_t) (h + 1) * kHeadDim + lane + 32 * j);
                }
                if (PAIR) warp_norm_rope(q, op.q_norm_w, op.rope, pos, op.eps, lane, op.q_norm_w != nullptr);
                else      warp_norm_rope_pre(q, wq, rc, rs, op.eps, op.q_norm_w != nullptr);
                float o[4];
                if (PAIR) {
                    if (t == 0) {
#pragma unroll
                        for (int j = 0; j < 4; ++j) o[j] = v[j];      // softmax over one key = 1
                    } else {
                        float s0 = 0.f, s1 = 0.f;
#pragma unroll
                        for (int j = 0; j < 4; ++j) { s0 = fmaf(k0f[j], q[j], s0); s1 = fmaf(k[j], q[j], s1); }
                        s0 = warp_sum(s0) * op.kq_scale; s1 = warp_sum(s1) * op.kq_scale;
                        const float mx = fmaxf(s0, s1);
                        const float e0 = expf(s0 - mx), e1 = expf(s1 - mx);
                        const float inv = (float) (1.0 / ((double) e0 + (double) e1));
                        const float p0 = e0 * inv, p1 = e1 * inv;
#pragma unroll
                        for (int j = 0; j < 4; ++j) o[j] = fmaf(v[j], p1, v0f[j] * p0);
                    }
                } else {
                    // F16 Q, keys/values from the F16 cache (the current token included: just written).
                    // All key rows are requested at once: half-warp `hw` takes the keys of its parity, lane
                    // l16 of it 8 consecutive dims (one 16-byte load per key) - a key loop with a dependent
                    // load + warp reduction per key was 16 L2 round trips per head.
                    const int n_kv = pos + 1;       // <= 16
                    const int hw = lane >> 4, l16 = lane & 15;
#pragma unroll
                    for (int j = 0; j < 4; ++j) qs[lane + 32 * j] = f16r(q[j]);
                    __syncwarp();
                    float qa[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) qa[i] = qs[8 * l16 + i];
                    const __half * kbase = kc + g * kHeadDim + 8 * l16;
                    const __half * vbase = vc + g * kHeadDim + 8 * l16;
                    uint4 kv[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int jk = 2 * i + hw;
                        kv[i] = (jk < n_kv) ? *reinterpret_cast<const uint4 *>(kbase + (size_t) jk * kvw) : make_uint4(0u, 0u, 0u, 0u);
                    }
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const __half2 * hk = reinterpret_cast<const __half2 *>(&kv[i]);
                        float sdot = 0.f;
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const float2 f = __half22float2(hk[u]);
                            sdot = fmaf(f.x, qa[2 * u], sdot);
                            sdot = fmaf(f.y, qa[2 * u + 1], sdot);
                        }
#pragma unroll
                        for (int o2 = 8; o2 > 0; o2 >>= 1) sdot += __shfl_xor_sync(0xffffffffu, sdot, o2);
                        if (l16 == 0 && 2 * i + hw < n_kv) scw[2 * i + hw] = sdot * op.kq_scale;
                    }
                    // the value rows are requested before the softmax needs the scores
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int jk = 2 * i + hw;
                        kv[i] = (jk < n_kv) ? *reinterpret_cast<const uint4 *>(vbase + (size_t) jk * kvw) : make_uint4(0u, 0u, 0u, 0u);
                    }
                    __syncwarp();
                    float mx = -CUDART_INF_F;
                    for (int jk = 0; jk < n_kv; ++jk) mx = fmaxf(mx, scw[jk]);
                    double sum = 0.0;
                    for (int jk = 0; jk < n_kv; ++jk) sum += (double) expf(scw[jk] - mx);
                    const float inv = (float) (1.0 / sum);
                    // partial sums per key class jk mod 4, combined as (p0 + p1) + (p2 + p3) like the batch-1 kernel:
                    // this half-warp holds classes hw (even i) and hw + 2 (odd i)
                    float accA[8], accB[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) { accA[u] = 0.f; accB[u] = 0.f; }
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int jk = 2 * i + hw;
                        if (jk < n_kv) {
                            const float pk = f16r(expf(scw[jk] - mx) * inv);       // the P.V mul_mat rounds P to F16
                            const __half2 * hv = reinterpret_cast<const __half2 *>(&kv[i]);
#pragma unroll
                            for (int u = 0; u < 4; ++u) {
                                const float2 f = __half22float2(hv[u]);
                                if (i & 1) { accB[2 * u] = fmaf(f.x, pk, accB[2 * u]); accB[2 * u + 1] = fmaf(f.y, pk, accB[2 * u + 1]); }
                                else       { accA[2 * u] = fmaf(f.x, pk, accA[2 * u]); accA[2 * u + 1] = fmaf(f.y, pk, accA[2 * u + 1]); }
                            }
                        }
                    }
                    __half2 oh[4];
#pragma unroll
                    for (int u = 0; u < 8; u += 2) {
                        // lanes of half-warp 0 end up with (p0 + p1) + (p2 + p3) of dims 8 * l16 + u, u + 1
                        const float a0 = accA[u] + __shfl_xor_sync(0xffffffffu, accA[u], 16);
                        const float b0 = accB[u] + __shfl_xor_sync(0xffffffffu, accB[u], 16);
                        const float a1 = accA[u + 1] + __shfl_xor_sync(0xffffffffu, accA[u + 1], 16);
                        const float b1 = accB[u + 1] + __shfl_xor_sync(0xffffffffu, accB[u + 1], 16);
                        oh[u >> 1] = __floats2half2_rn(a0 + b0, a1 + b1);
                    }
                    if (hw == 0) {
                        uint4 pk4;
                        pk4.x = *reinterpret_cast<const uint32_t *>(&oh[0]); pk4.y = *reinterpret_cast<const uint32_t *>(&oh[1]);
                        pk4.z = *reinterpret_cast<const uint32_t *>(&oh[2]); pk4.w = *reinterpret_cast<const uint32_t *>(&oh[3]);
                        *reinterpret_cast<uint4 *>(reinterpret_cast<unsigned char *>(op.attn16) +
                                                   canon_off(t * B + b, h * kHeadDim + 8 * l16, canon_chunk_bytes(op.rows_act))) = pk4;
                    }
                    __syncwarp();
                    continue;
                }
                unsigned char * out = reinterpret_cast<unsigned char *>(op.attn16);
                const int cb = canon_chunk_bytes(op.rows_act);
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    *reinterpret_cast<__half *>(out + canon_off(t * B + b, h * kHeadDim + lane + 32 * j, cb)) = __float2half_rn(o[j]);
            }
        }
    }
}

// Single-token code-predictor attention, one warp per (stream, QUERY HEAD): twice the items of the per-group form
// above (1024 at 64 streams: every warp of the grid still gets at most one), half the dependent chain per warp,
// and the softmax over the <= 16 keys is evaluated by the lanes in parallel (lane j owns key j; the double sum runs
// through the same xor tree as the batch-1 kernel's block_sum_d over one warp).  Both head-warps of a KV group
// compute the group's new key themselves; the first one appends it to the cache, and neither reads row `pos` back
// from global memory (the current key / value travel through the warp's own shared-memory scratch).
__device__ __noinline__ void bstage_attn_cp_head(BShared * S) {
    const BOp & op = S->op;
    const int lane = threadIdx.x & 31, hw = lane >> 4, l16 = lane & 15;
    const int wib = threadIdx.x >> 5;
    const int gw = blockIdx.x * kWarps + wib, nw = gridDim.x * kWarps;
    const int B = op.n_streams, nh = op.n_heads, nkv = op.n_kv_heads, rep = nh / nkv;
    const int kvw = nkv * kHeadDim;
    const int pos = op.pos, n_kv = pos + 1;          // <= 16
    // per-warp scratch in the (idle) A-buffer area: q (F16-rounded, as floats), current key / value (F16), scores
    unsigned char * wsm = babuf_base(S) + (size_t) wib * 1280;
    float * qs = reinterpret_cast<float *>(wsm);                       // [128]
    __half * ks = reinterpret_cast<__half *>(wsm + 512);               // [128]
    __half * vs = reinterpret_cast<__half *>(wsm + 768);               // [128]
    float * scw = reinterpret_cast<float *>(wsm + 1024);               // [16] scores, then probabilities
    float wq[4], wk[4], rc[2], rs[2];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        wq[j] = op.q_norm_w ? __ldg(op.q_norm_w + lane + 32 * j) : 1.f;
        wk[j] = op.k_norm_w ? __ldg(op.k_norm_w + lane + 32 * j) : 1.f;
    }
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        const float2 cs = __ldg(reinterpret_cast<const float2 *>(op.rope) + (size_t) pos * 64 + lane + 32 * j);
        rc[j] = cs.x; rs[j] = cs.y;
    }
    const int cb = canon_chunk_bytes(op.rows_act);
    for (int item = gw; item < B * nh; item += nw) {
        const int b = item / nh, h = item - b * nh, g = h / rep;
        __half * kc = op.kc + (size_t) b * op.kv_stream_stride + g * kHeadDim;
        __half * vc = op.vc + (size_t) b * op.kv_stream_stride + g * kHeadDim;
        const float * base = op.qkv + (size_t) b * op.qkvw;
        // cached rows of the earlier positions: requested first, they do not depend on this token
        uint4 kv[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int jk = 2 * i + hw;
            kv[i] = (jk < pos) ? __ldcg(reinterpret_cast<const uint4 *>(kc + (size_t) jk * kvw + 8 * l16)) : make_uint4(0u, 0u, 0u, 0u);
        }
        float k[4], v[4], q[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            k[j] = __ldcg(base + (size_t) (nh + g) * kHeadDim + lane + 32 * j);
            v[j] = __ldcg(base + (size_t) (nh + nkv + g) * kHeadDim + lane + 32 * j);
            q[j] = __ldcg(base + (size_t) h * kHeadDim + lane + 32 * j);
        }
        warp_norm_rope_pre(k, wk, rc, rs, op.eps, op.k_norm_w != nullptr);
        warp_norm_rope_pre(q, wq, rc, rs, op.eps, op.q_norm_w != nullptr);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const __half kh = __float2half_rn(k[j]), vh = __float2half_rn(v[j]);
            ks[lane + 32 * j] = kh; vs[lane + 32 * j] = vh;
            qs[lane + 32 * j] = f16r(q[j]);
            if (h == g * rep) {
                kc[(size_t) pos * kvw + lane + 32 * j] = kh;
                vc[(size_t) pos * kvw + lane + 32 * j] = vh;
            }
        }
        __syncwarp();
        float qa[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) qa[i] = qs[8 * l16 + i];
        {
            const uint4 kcur = *reinterpret_cast<const uint4 *>(ks + 8 * l16);
#pragma unroll
            for (int i = 0; i < 8; ++i) if (2 * i + hw == pos) kv[i] = kcur;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const __half2 * hk = reinterpret_cast<const __half2 *>(&kv[i]);
            float sdot = 0.f;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const float2 f = __half22float2(hk[u]);
                sdot = fmaf(f.x, qa[2 * u], sdot);
                sdot = fmaf(f.y, qa[2 * u + 1], sdot);
            }
#pragma unroll
            for (int o2 = 8; o2 > 0; o2 >>= 1) sdot += __shfl_xor_sync(0xffffffffu, sdot, o2);
            if (l16 == 0 && 2 * i + hw < n_kv) scw[2 * i + hw] = sdot * op.kq_scale;
        }
        // the value rows are requested before the softmax needs the scores
        {
            const uint4 vcur = *reinterpret_cast<const uint4 *>(vs + 8 * l16);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int jk = 2 * i + hw;
                kv[i] = (jk < pos) ? __ldcg(reinterpret_cast<const uint4 *>(vc + (size_t) jk * kvw + 8 * l16))
                                   : (jk == pos ? vcur : make_uint4(0u, 0u, 0u, 0u));
            }
        }
        __syncwarp();
        // softmax: lane j owns key j (ggml soft_max: exp(x - max), double sum, scale by (float) (1 / sum); the P.V
        // mul_mat then rounds P to F16)
        const float sj = (lane < n_kv) ? scw[lane] : -CUDART_INF_F;
        const float mx = warp_max(sj);
        const float ej = (lane < n_kv) ? expf(sj - mx) : 0.f;
        const double sum = warp_sum_d((double) ej);
        const float inv = (float) (1.0 / sum);
        __syncwarp();
        if (lane < n_kv) scw[lane] = f16r(ej * inv);
        __syncwarp();
        float accA[8], accB[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) { accA[u] = 0.f; accB[u] = 0.f; }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int jk = 2 * i + hw;
            if (jk < n_kv) {
                const float pk = scw[jk];
                const __half2 * hv = reinterpret_cast<const __half2 *>(&kv[i]);
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const float2 f = __half22float2(hv[u]);
                    if (i & 1) { accB[2 * u] = fmaf(f.x, pk, accB[2 * u]); accB[2 * u + 1] = fmaf(f.y, pk, accB[2 * u + 1]); }
                    else       { accA[2 * u] = fmaf(f.x, pk, accA[2 * u]); accA[2 * u + 1] = fmaf(f.y, pk, accA[2 * u + 1]); }
                }
            }
        }
        __half2 oh[4];
#pragma unroll
        for (int u = 0; u < 8; u += 2) {
            // half-warp 0 holds key classes 0 (A) and 2 (B), half-warp 1 classes 1 and 3: (p0 + p1) + (p2 + p3)
            const float a0 = accA[u] + __shfl_xor_sync(0xffffffffu, accA[u], 16);
            const float b0 = accB[u] + __shfl_xor_sync(0xffffffffu, accB[u], 16);
            const float a1 = accA[u + 1] + __shfl_xor_sync(0xffffffffu, accA[u + 1], 16);
            const float b1 = accB[u + 1] + __shfl_xor_sync(0xffffffffu, accB[u + 1], 16);
            oh[u >> 1] = __floats2half2_rn(a0 + b0, a1 + b1);
        }
        if (hw == 0) {
            uint4 pk4;
            pk4.x = *reinterpret_cast<const uint32_t *>(&oh[0]); pk4.y = *reinterpret_cast<const uint32_t *>(&oh[1]);
            pk4.z = *reinterpret_cast<const uint32_t *>(&oh[2]); pk4.w = *reinterpret_cast<const uint32_t *>(&oh[3]);
            *reinterpret_cast<uint4 *>(reinterpret_cast<unsigned char *>(op.attn16) + canon_off(b, h * kHeadDim + 8 * l16, cb)) = pk4;
        }
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------------
// BOP_ATTN_TK — Talker decode attention, one CTA pass per (stream, KV group): QK-norm + NEOX RoPE
// + F16 KV append, FP32 softmax and P.V over the F16 cache (ATTN_FAST numerics of qtts_kernels.cu).
// ------------------------------------------------------------------------------------------
template <int REP>
__device__ void battn_tk_item(const Op & iop, const float * qkv, unsigned char * out16, int row, int cb, int pos, AttnSmem * sm, float * sc) {
    const int tid = threadIdx.x;
    const int n_kv = pos + 1, kvw = iop.n_kv_heads * kHeadDim;
    const int g = iop.book;                      // KV group of this item (field reused)
    const int stride = (n_kv + 3) & ~3;
    attn_prepare(iop, qkv, g * REP, REP, g, pos, sm);
    attn_append(iop, g, pos, kvw, sm);
    attn_scores<REP>(iop, sm, g, kvw, pos, 0, n_kv, sc, stride);
    __syncthreads();
    float l[REP];
#pragma unroll
    for (int r = 0; r < REP; ++r) {
        float mx = -CUDART_INF_F;
        for (int j = tid; j < n_kv; j += kThreads) mx = fmaxf(mx, sc[r * stride + j]);
        mx = block_max_f(mx, sm->redf);
        double sum = 0.0;
        for (int j = tid; j < n_kv; j += kThreads) {
            const float e = __expf(sc[r * stride + j] - mx);
            sc[r * stride + j] = e;
            sum += (double) e;
        }
        sum = block_sum_d(sum, sm->red);
        l[r] = (float) sum;
    }
    __syncthreads();
    {
        const int d = tid & (kHeadDim - 1), part = tid >> 7;
        float acc[REP];
#pragma unroll
        for (int r = 0; r < REP; ++r) acc[r] = 0.f;
        for (int j = part; j < n_kv; j += 4) {
            const float v = __half2float((j < pos) ? __ldcg(iop.vcache + (size_t) j * kvw + g * kHeadDim + d) : sm->v16[d]);
#pragma unroll
            for (int r = 0; r < REP; ++r) acc[r] = fmaf(v, sc[r * stride + j], acc[r]);
        }
#pragma unroll
        for (int r = 0; r < REP; ++r) sm->part[part][r][d] = acc[r];
    }
    __syncthreads();
    for (int idx = tid; idx < REP * kHeadDim; idx += kThreads) {
        const int r = idx / kHeadDim, d = idx % kHeadDim;
        const float o = (sm->part[0][r][d] + sm->part[1][r][d]) + (sm->part[2][r][d] + sm->part[3][r][d]);
        const float lr = (r == 0) ? l[0] : l[REP - 1 < r ? REP - 1 : r];
        *reinterpret_cast<__half *>(out16 + canon_off(row, (g * REP + r) * kHeadDim + d, cb)) = __float2half_rn(o / lr);
    }
    __syncthreads();
}

// Talker attention, one WARP per (stream, KV group) - the form the code-predictor attention has, with
// an online softmax over blocks of 16 keys so that any context length fits: half-warp `hw` takes the
// keys of its parity, lane l16 of it 8 consecutive dims (one 16-byte load per key and operand).
// ATTN_FAST numerics: F16 Q / K / V, FP32 scores, softmax and P.V.  All (stream, group) items of the
// stage run concurrently (2048 warps at 256 streams) instead of 14 CTA-wide items in sequence per SM.
// KSM = true (the _ext kernel with four activation buffers: 9 KB of shared memory per warp): the K rows of a block are staged
// by cp.async like the V rows and the loop is software-pipelined - K of block j+1 is requested as soon as block j's scores
// are done, V of block j+1 as soon as its P.V is done - so every warp always has 4 - 8 KB of KV in flight and no K load
// sits on the dependency chain.  KSM = false: K through registers (4.5 KB per warp in the two-buffer layout).
template <bool KSM>
__device__ __noinline__ void bstage_attn_tk_warp(BShared * S) {
    constexpr int REP = 2;
    const BOp & op = S->op;
    const BRun & run = S->run;
    const int lane = threadIdx.x & 31, hw = lane >> 4, l16 = lane & 15;
    const int gw = blockIdx.x * kWarps + (threadIdx.x >> 5), nw = gridDim.x * kWarps;
    const int B = op.n_streams, nkv = op.n_kv_heads, kvw = nkv * kHeadDim;
    // per-warp shared memory (4608 B of the activation-buffer region, free during an attention stage), used in turn
    // as: Q redistribution (1 KB) -> staging of the block's V rows (4 KB, written by cp.async) -> split-KV hand-over
    constexpr int kWarpRegion = (KSM ? 4 : 2) * kABufBytes / kWarps;
    static_assert(kWarpRegion >= (KSM ? 8192 : 4096) && kWarpRegion >= 20 * 32 * 4, "attention warp region");
    unsigned char * wreg = babuf_base(S) + (threadIdx.x >> 5) * kWarpRegion;
    float * qs = reinterpret_cast<float *>(wreg);
    const int cb = canon_chunk_bytes(op.rows_act);
    float wq[4], wk[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        wq[j] = op.q_norm_w ? __ldg(op.q_norm_w + lane + 32 * j) : 1.f;
        wk[j] = op.k_norm_w ? __ldg(op.k_norm_w + lane + 32 * j) : 1.f;
    }
    // split-KV over `spl` adjacent warps of the CTA when there are fewer items than warps (<= 128 streams: 2, <= 64: 4):
    // the key range is dealt in blocks of 16, the LAST warp of the group owns the tail (and the new K/V row), the
    // others hand (m, l, acc) over through shared memory and it merges them in fixed order.  A long context is
    // latency-bound per warp (two dependent L2/DRAM round trips per 16 keys), so bytes in flight scale with the warps.
    const int wib = threadIdx.x >> 5;
    int spl = 1;
    if (4 * B * nkv <= nw) spl = 4; else if (2 * B * nkv <= nw) spl = 2;
    if (S->run.dbg & 16) spl = 1;      // A/B switch (QTTS_BDBG=16): one warp per item
    const int part = gw % spl, ngrp = nw / spl;
    const bool tail = part == spl - 1;
    const uint32_t vst = smem_u32(wreg) + (uint32_t) (threadIdx.x & 31) * 16u;       // this lane's slot of V row i: + i * 512
    for (int item = gw / spl; item < B * nkv; item += ngrp) {
        const int b = item / nkv, g = item - b * nkv;
        const int pos = __ldcg(&run.st[b].n_past), n_kv = pos + 1;
        __half * kc = op.kc + (size_t) b * op.kv_stream_stride + g * kHeadDim;
        __half * vc = op.vc + (size_t) b * op.kv_stream_stride + g * kHeadDim;
        const float * base = op.qkv + (size_t) b * op.qkvw;
        float k[4], v[4], q0[4], q1[4], rc[2], rs[2];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            k[j] = __ldcg(base + (size_t) (op.n_heads + g) * kHeadDim + lane + 32 * j);
            v[j] = __ldcg(base + (size_t) (op.n_heads + nkv + g) * kHeadDim + lane + 32 * j);
            q0[j] = __ldcg(base + (size_t) (g * REP) * kHeadDim + lane + 32 * j);
            q1[j] = __ldcg(base + (size_t) (g * REP + 1) * kHeadDim + lane + 32 * j);
        }
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const float2 cs = __ldg(reinterpret_cast<const float2 *>(op.rope) + (size_t) pos * 64 + lane + 32 * j);
            rc[j] = cs.x; rs[j] = cs.y;
        }
        if (tail) {
            warp_norm_rope_pre(k, wk, rc, rs, op.eps, op.k_norm_w != nullptr);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                kc[(size_t) pos * kvw + lane + 32 * j] = __float2half_rn(k[j]);
                vc[(size_t) pos * kvw + lane + 32 * j] = __float2half_rn(v[j]);
            }
        }
        warp_norm_rope_pre(q0, wq, rc, rs, op.eps, op.q_norm_w != nullptr);
        warp_norm_rope_pre(q1, wq, rc, rs, op.eps, op.q_norm_w != nullptr);
#pragma unroll
        for (int j = 0; j < 4; ++j) { qs[lane + 32 * j] = f16r(q0[j]); qs[kHeadDim + lane + 32 * j] = f16r(q1[j]); }
        __syncwarp();
        float qa[REP][8];
#pragma unroll
        for (int r = 0; r < REP; ++r)
#pragma unroll
            for (int i = 0; i < 8; ++i) qa[r][i] = qs[r * kHeadDim + 8 * l16 + i];
        float m[REP], l[REP], acc[REP][8];
#pragma unroll
        for (int r = 0; r < REP; ++r) {
            m[r] = -CUDART_INF_F; l[r] = 0.f;
#pragma unroll
            for (int u = 0; u < 8; ++u) acc[r][u] = 0.f;
        }
        const __half * kb = kc + 8 * l16, * vb = vc + 8 * l16;
        const int per = ((n_kv + 15) / 16) / spl;                   // 16-key blocks of every warp but the last
        const int jb0 = part * per * 16, jb1 = tail ? n_kv : (part + 1) * per * 16;
        // Per 16-key block: K rows to registers, V rows by cp.async into the warp's shared-memory staging buffer at the same
        // time (they land while the scores are computed; profiles/r01p: 14.5 % of the samples waited on the V rows with
        // nothing else pending), and both operands of the blocks further ahead prefetched into L2.
        // Q stays packed F16 and the dot product runs on the mixed-precision FMA: same products, same order, same
        // single rounding as fmaf on the converted values (both conversions are exact), without the conversions.
        uint32_t qh[REP][4];
        __syncwarp();       // every lane has read its Q values: the region becomes the V staging buffer
#pragma unroll
        for (int r = 0; r < REP; ++r)
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const __half2 h = __floats2half2_rn(qa[r][2 * u], qa[r][2 * u + 1]);      // exact: qa holds F16 values
                qh[r][u] = *reinterpret_cast<const uint32_t *>(&h);
            }
        uint4 kA[8];
        const uint4 z4 = make_uint4(0u, 0u, 0u, 0u);
        const uint32_t kst = vst + 4096u;        // KSM: this lane's slot of K row i: + i * 512
        auto stage_rows = [&](const __half * base, uint32_t dst, int jb_) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int jk = jb_ + 2 * i + hw;
                const bool ok = jk < n_kv;
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;"
                             :: "r"(dst + (uint32_t) i * 512u), "l"(base + (size_t) (ok ? jk : 0) * kvw), "r"(ok ? 16 : 0) : "memory");
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        };
        if (KSM && jb0 < jb1) { stage_rows(kb, kst, jb0); stage_rows(vb, vst, jb0); }
        // Optional L2 prefetch `pfd` blocks ahead (A/B switch, off): lane l asks for line (l & 1) of key (l >> 1)'s K and V
        // segments (256 B each).  1.7B, 128 streams, 2054 keys: 14.60 ms per frame-step without, 14.46 at depth 2, 15.65 at
        // depth 3, 16.84 at depth 5 (profiles/r02q_1p7b_attention_ab.log) - the 4096 attention warps already keep HBM busy.
        const int pfd = (S->run.dbg >> 8) & 7;      // QTTS_BDBG = depth << 8; measured: 0 / 2 equal, 3 and 5 slower -> off by default
        const size_t pf_off = (size_t) (lane >> 1) * kvw + (size_t) (lane & 1) * 64;      // in halves
        for (int d = 0; d < pfd; ++d) {
            const int jk = jb0 + d * 16 + (lane >> 1);
            if (jk < jb1) {
                asm volatile("prefetch.global.L2 [%0];" :: "l"(kc + (size_t) (jb0 + d * 16) * kvw + pf_off));
                asm volatile("prefetch.global.L2 [%0];" :: "l"(vc + (size_t) (jb0 + d * 16) * kvw + pf_off));
            }
        }
        for (int jb = jb0; jb < jb1; jb += 16) {
            if (pfd > 0 && jb + pfd * 16 + (lane >> 1) < jb1) {
                asm volatile("prefetch.global.L2 [%0];" :: "l"(kc + (size_t) (jb + pfd * 16) * kvw + pf_off));
                asm volatile("prefetch.global.L2 [%0];" :: "l"(vc + (size_t) (jb + pfd * 16) * kvw + pf_off));
            }
            const bool more = jb + 16 < jb1;
            if (KSM) {
                // pending groups here: K(jb), V(jb), in this order: K(jb) has landed once at most one is pending
                asm volatile("cp.async.wait_group 1;" ::: "memory");
            } else {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int jk = jb + 2 * i + hw;
                    kA[i] = (jk < n_kv) ? *reinterpret_cast<const uint4 *>(kb + (size_t) jk * kvw) : z4;
                }
                stage_rows(vb, vst, jb);
            }
            float sc[REP][8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (KSM) asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(kA[i].x), "=r"(kA[i].y), "=r"(kA[i].z), "=r"(kA[i].w) : "r"(kst + (uint32_t) i * 512u) : "memory");
                float d0 = 0.f, d1 = 0.f;
                d0 = fhfma2(kA[i].x, qh[0][0], d0); d1 = fhfma2(kA[i].x, qh[1][0], d1);
                d0 = fhfma2(kA[i].y, qh[0][1], d0); d1 = fhfma2(kA[i].y, qh[1][1], d1);
                d0 = fhfma2(kA[i].z, qh[0][2], d0); d1 = fhfma2(kA[i].z, qh[1][2], d1);
                d0 = fhfma2(kA[i].w, qh[0][3], d0); d1 = fhfma2(kA[i].w, qh[1][3], d1);
#pragma unroll
                for (int o2 = 8; o2 > 0; o2 >>= 1) {
                    d0 += __shfl_xor_sync(0xffffffffu, d0, o2);
                    d1 += __shfl_xor_sync(0xffffffffu, d1, o2);
                }
                const bool ok = jb + 2 * i + hw < n_kv;
                sc[0][i] = ok ? d0 * op.kq_scale : -CUDART_INF_F;
                sc[1][i] = ok ? d1 * op.kq_scale : -CUDART_INF_F;
            }
#pragma unroll
            for (int r = 0; r < REP; ++r) {
                float mb = sc[r][0];
#pragma unroll
                for (int i = 1; i < 8; ++i) mb = fmaxf(mb, sc[r][i]);
                mb = fmaxf(mb, __shfl_xor_sync(0xffffffffu, mb, 16));
                const float mn = fmaxf(m[r], mb);                   // finite: key jb exists
                const float scale = __expf(m[r] - mn);              // 0 on the first block
                float ls = 0.f;
#pragma unroll
                for (int i = 0; i < 8; ++i) { sc[r][i] = __expf(sc[r][i] - mn); ls += sc[r][i]; }
                ls += __shfl_xor_sync(0xffffffffu, ls, 16);
                l[r] = fmaf(l[r], scale, ls);
                m[r] = mn;
#pragma unroll
                for (int u = 0; u < 8; ++u) acc[r][u] *= scale;
            }
            if (KSM && more) {
                stage_rows(kb, kst, jb + 16);                          // the K staging buffer is free: its rows are in the scores
                asm volatile("cp.async.wait_group 1;" ::: "memory");   // pending: V(jb), K(jb + 16) -> V(jb) has landed
            } else {
                asm volatile("cp.async.wait_group 0;" ::: "memory");
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                uint4 vv;
                asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(vv.x), "=r"(vv.y), "=r"(vv.z), "=r"(vv.w) : "r"(vst + (uint32_t) i * 512u) : "memory");
                const __half2 * hv = reinterpret_cast<const __half2 *>(&vv);
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const float2 f = __half22float2(hv[u]);
                    acc[0][2 * u] = fmaf(f.x, sc[0][i], acc[0][2 * u]); acc[0][2 * u + 1] = fmaf(f.y, sc[0][i], acc[0][2 * u + 1]);
                    acc[1][2 * u] = fmaf(f.x, sc[1][i], acc[1][2 * u]); acc[1][2 * u + 1] = fmaf(f.y, sc[1][i], acc[1][2 * u + 1]);
                }
            }
            if (KSM && more) stage_rows(vb, vst, jb + 16);         // the V staging buffer is free
        }
        if (spl > 1) {
            if (!tail) {
                float * pw = reinterpret_cast<float *>(wreg) + lane;
#pragma unroll
                for (int r = 0; r < REP; ++r) {
                    pw[(r * 10 + 0) * 32] = m[r]; pw[(r * 10 + 1) * 32] = l[r];
#pragma unroll
                    for (int u = 0; u < 8; ++u) pw[(r * 10 + 2 + u) * 32] = acc[r][u];
                }
            }
            asm volatile("bar.sync %0, %1;" :: "r"(1 + wib / spl), "r"(32 * spl) : "memory");
            if (tail) {
                for (int p = 0; p < spl - 1; ++p) {
                    const float * pw = reinterpret_cast<const float *>(wreg - (size_t) (spl - 1 - p) * kWarpRegion) + lane;
#pragma unroll
                    for (int r = 0; r < REP; ++r) {
                        const float mp = pw[(r * 10 + 0) * 32], lp = pw[(r * 10 + 1) * 32];
                        const float mn = fmaxf(m[r], mp);            // finite: the tail warp always has a block
                        const float st = __expf(m[r] - mn), sp = __expf(mp - mn);     // an empty part has m = -inf -> 0
                        l[r] = fmaf(l[r], st, lp * sp);
                        m[r] = mn;
#pragma unroll
                        for (int u = 0; u < 8; ++u) acc[r][u] = fmaf(acc[r][u], st, pw[(r * 10 + 2 + u) * 32] * sp);
                    }
                }
            }
            asm volatile("bar.sync %0, %1;" :: "r"(1 + wib / spl), "r"(32 * spl) : "memory");
        }
#pragma unroll
        for (int r = 0; r < REP; ++r) {
            if (!tail) break;
            __half2 oh[4];
            const float inv = 1.0f / l[r];
#pragma unroll
            for (int u = 0; u < 8; u += 2) {
                const float a0 = acc[r][u] + __shfl_xor_sync(0xffffffffu, acc[r][u], 16);
                const float a1 = acc[r][u + 1] + __shfl_xor_sync(0xffffffffu, acc[r][u + 1], 16);
                oh[u >> 1] = __floats2half2_rn(a0 * inv, a1 * inv);
            }
            if (hw == 0) {
                uint4 pk4;
                pk4.x = *reinterpret_cast<const uint32_t *>(&oh[0]); pk4.y = *reinterpret_cast<const uint32_t *>(&oh[1]);
                pk4.z = *reinterpret_cast<const uint32_t *>(&oh[2]); pk4.w = *reinterpret_cast<const uint32_t *>(&oh[3]);
                *reinterpret_cast<uint4 *>(reinterpret_cast<unsigned char *>(op.attn16) +
                                           canon_off(b, (g * REP + r) * kHeadDim + 8 * l16, cb)) = pk4;
            }
        }
        __syncwarp();
    }
}

__device__ __noinline__ void bstage_attn_tk(BShared * S) {
    const BOp & op = S->op;
    const BRun & run = S->run;
    unsigned char * scratch = babuf_base(S);
    Op * iop = &S->item_op;
    AttnSmem * sm = reinterpret_cast<AttnSmem *>(scratch);
    float * sc = reinterpret_cast<float *>(scratch + ((sizeof(AttnSmem) + 15) & ~(size_t) 15));
    const int B = op.n_streams, nkv = op.n_kv_heads, rep = op.n_heads / nkv;
    for (int item = blockIdx.x; item < B * nkv; item += gridDim.x) {
        const int b = item / nkv, g = item - b * nkv;
        const int pos = __ldcg(&run.st[b].n_past);
        __syncthreads();
        if (threadIdx.x == 0) {
            iop->n_heads = op.n_heads; iop->n_kv_heads = nkv;
            iop->q_norm_w = op.q_norm_w; iop->k_norm_w = op.k_norm_w;
            iop->eps = op.eps; iop->rope = op.rope; iop->mode = ATTN_FAST; iop->kq_scale = op.kq_scale;
            iop->kcache = op.kc + (size_t) b * op.kv_stream_stride;
            iop->vcache = op.vc + (size_t) b * op.kv_stream_stride;
            iop->book = g;
        }
        __syncthreads();
        const float * qkv = op.qkv + (size_t) b * op.qkvw;
        unsigned char * out16 = reinterpret_cast<unsigned char *>(op.attn16);
        const int cb = canon_chunk_bytes(op.rows_act);
        if (rep == 1)      battn_tk_item<1>(*iop, qkv, out16, b, cb, pos, sm, sc);
        else if (rep == 2) battn_tk_item<2>(*iop, qkv, out16, b, cb, pos, sm, sc);
        else               battn_tk_item<4>(*iop, qkv, out16, b, cb, pos, sm, sc);
    }
}

// ------------------------------------------------------------------------------------------
// BOP_SAMPLE — per stream (CTA b mod grid): logits post-processing + sampling
// (src/tts_transformer.cpp:2417-2499 for codebook 0, :2198-2236 for the code predictor), then the
// embedding rows of the NEXT pass, and for the last codebook the frame end (:2500-2527) and the
// next Talker input sum (:2529-2553).
// ------------------------------------------------------------------------------------------
__device__ __noinline__ void bstage_sample(BShared * S) {
    const BOp & op = S->op;
    const BRun & run = S->run;
    unsigned char * scratch = babuf_base(S);
    double * red = S->red;
    SampleSmem * sm = reinterpret_cast<SampleSmem *>(scratch);
    float * pbuf = reinterpret_cast<float *>(scratch + ((sizeof(SampleSmem) + 15) & ~(size_t) 15));
    const int B = op.n_streams, n = op.n_vocab, tid = threadIdx.x;
    for (int b = blockIdx.x; b < B; b += gridDim.x) {
        StreamState * st = run.st + b;
        __syncthreads();
        const bool live = run.active[b] && !__ldcg(&st->done);
        const int frame = __ldcg(&st->frame);
        const float temperature = st->temperature;
        const int top_k = st->top_k;
        const float * lg = op.logits + (size_t) b * op.logits_stride;
        int tok;
        if (op.kind == 0) {
            const float rep = st->rep_pen;
            for (int i = tid; i < n; i += kThreads) {
                float v = __ldcg(lg + i);
                if (i >= op.suppress_start && i != op.eos_id) v = -CUDART_INF_F;
                if (rep != 1.0f && ((__ldcg(&st->seen[i >> 5]) >> (i & 31)) & 1u)) v = (v > 0.0f) ? v / rep : v * rep;
                sm->l[i] = v;
            }
            __syncthreads();
            const int expected = max(20, st->n_text_tokens * 4);
            if (frame >= expected) {
                float mx = -CUDART_INF_F;
                for (int i = tid; i < n; i += kThreads) mx = fmaxf(mx, sm->l[i]);
                mx = block_max_f(mx, sm->redf);
                if (tid == 0) {
                    const float ramp = fminf(1.0f, (float) (frame - expected) / (float) expected);
                    const float target = mx + 5.0f;
                    // one FMA, like the reference as compiled (GCC contracts `l += ramp * (target - l)`)
                    sm->l[op.eos_id] = __fmaf_rn(ramp, target - sm->l[op.eos_id], sm->l[op.eos_id]);
                }
                __syncthreads();
            }
            if (op.dump) for (int i = tid; i < n; i += kThreads) op.dump[(size_t) b * op.dump_stride + i] = sm->l[i];
            const int ignore_eos = st->ignore_eos;
            if (ignore_eos) {
                if (tid == 0) sm->l[op.eos_id] = -CUDART_INF_F;
                __syncthreads();
            }
            if (temperature <= 0.0f) tok = block_argmax(sm->l, n, sm);
            else {
                const float * su = st->uniforms;
                const float u = (su && frame < st->uniform_frames) ? __ldcg(su + (size_t) frame * op.n_books)
                                                                   : philox_uniform(st->seed, (uint32_t) frame, 0u, 0x0cb0u);
                tok = block_sample(n, temperature, top_k, u, ignore_eos ? -1 : op.eos_id, sm, pbuf, st->sample_mode);
            }
            __syncthreads();
            // teacher forcing (parity hook): the own choice is emitted, the forced code is fed back
            const int32_t * fc = st->forced;
            const bool forcing = fc != nullptr && frame < st->forced_frames;
            const int fed = forcing ? __ldcg(fc + (size_t) frame * op.n_books) : tok;
            if (tid == 0 && live) {
                if (tok == op.eos_id && !forcing) { st->done = 1; st->eos = 1; }
                else {
                    st->cur_codes[0] = tok; st->feed_codes[0] = fed;
                    st->seen[fed >> 5] |= 1u << (fed & 31);
                }
            }
            // code-predictor input rows: token 0 = Talker hidden, token 1 = codec_embd[cb0]
            const int feed = (live && (tok != op.eos_id || forcing)) ? fed : __ldcg(&st->feed_codes[0]);
            const __half * row = op.table + (size_t) feed * op.xw;
            double s0 = 0.0, s1 = 0.0;
            for (int i = tid; i < op.xw; i += kThreads) {
                const float h = __ldcg(op.hidden + (size_t) b * op.hidden_stride + i);
                const float e = __half2float(row[i]);
                op.x_out[(size_t) b * op.xw + i] = h;
                op.x_out[(size_t) (B + b) * op.xw + i] = e;
                s0 += (double) (h * h); s1 += (double) (e * e);
            }
            s0 = block_sum_d(s0, red);
            s1 = block_sum_d(s1, red);
            if (tid == 0) { op.ssq_e[b] = (float) s0; op.ssq_e[B + b] = (float) s1; }
        } else {
            for (int i = tid; i < n; i += kThreads) {
                const float v = __ldcg(lg + i);
                sm->l[i] = v;
                if (op.dump) op.dump[(size_t) b * op.dump_stride + i] = v;
            }
            __syncthreads();
            if (temperature <= 0.0f) tok = block_argmax(sm->l, n, sm);
            else {
                const float * su = st->uniforms;
                const float u = (su && frame < st->uniform_frames) ? __ldcg(su + (size_t) frame * op.n_books + op.book)
                                                                   : philox_uniform(st->seed, (uint32_t) frame, (uint32_t) op.book, 0x0c9du);
                tok = block_sample(n, temperature, top_k, u, -1, sm, pbuf, 0);
            }
            __syncthreads();
            const int32_t * fc = st->forced;
            const int fed = (fc && frame < st->forced_frames) ? __ldcg(fc + (size_t) frame * op.n_books + op.book) : tok;
            if (tid == 0) { st->cur_codes[op.book] = tok; st->feed_codes[op.book] = fed; }
            if (op.book < op.n_books - 1) {
                const __half * row = op.table + (size_t) fed * op.xw;
                double s0 = 0.0;
                for (int i = tid; i < op.xw; i += kThreads) {
                    const float e = __half2float(row[i]);
                    op.x_out[(size_t) b * op.xw + i] = e;
                    s0 += (double) (e * e);
                }
                s0 = block_sum_d(s0, red);
                if (tid == 0) op.ssq_e[b] = (float) s0;
            } else {
                // frame end + next Talker input
                __syncthreads();
                // idx[0..n_books): the frame's own codes (emitted); idx[kMaxBooks..): the fed-back codes (next Talker input)
                if (tid < op.n_books) {
                    sm->idx[tid] = (unsigned short) ((tid == op.book) ? tok : __ldcg(&st->cur_codes[tid]));
                    sm->idx[kMaxBooks + tid] = (unsigned short) ((tid == op.book) ? fed : __ldcg(&st->feed_codes[tid]));
                }
                __syncthreads();
                int f1 = frame;
                if (live) {
                    if (tid < op.n_books)
                        op.codes_out[(size_t) b * op.codes_stride + (size_t) frame * op.n_books + tid] = (int32_t) sm->idx[tid];
                    f1 = frame + 1;
                    if (tid == 0) {
                        st->frame = f1;
                        if (f1 >= st->max_len) st->done = 1;
                    }
                }
                const int ft = max(f1 - 1, 0);
                const float * trow = (ft < st->trailing_len) ? op.trailing + (size_t) b * op.trailing_stride + (size_t) ft * op.xw
                                                             : op.pad + (size_t) b * op.pad_stride;
                const EmbedSumArgs * es = op.es;
                double s0 = 0.0;
                for (int i = tid; i < op.xw; i += kThreads) {
                    float acc = __half2float(es->tables[0][(size_t) sm->idx[kMaxBooks] * op.xw + i]);
                    for (int c = 1; c < es->n_books; ++c)
                        acc += __half2float(es->tables[c][(size_t) sm->idx[kMaxBooks + c] * op.xw + i]);
                    acc += __ldcg(trow + i);
                    op.x_out[(size_t) b * op.xw + i] = acc;
                    s0 += (double) (acc * acc);
                }
                s0 = block_sum_d(s0, red);
                if (tid == 0) op.ssq_e[b] = (float) s0;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// the batched program runner
// ------------------------------------------------------------------------------------------
constexpr size_t kBSmemBytes = kBSharedBytes + (size_t) kBSlots * kBSlotBytes + 2 * (size_t) kABufBytes;
static_assert(kBSmemBytes <= 232448, "batched kernel exceeds the 227 KB of shared memory per CTA");
static_assert(sizeof(AttnSmem) + 64 + 2 * kMaxScores * sizeof(float) <= 2 * (size_t) kABufBytes, "attention scratch overlays the A buffers");
static_assert(sizeof(SampleSmem) + 64 + kMaxVocab * sizeof(float) <= 2 * (size_t) kABufBytes, "sampling scratch overlays the A buffers");

// Two kernels from one body.  qtts_run_batched (EXT = false) is the configuration of groups of <= 64 streams: two
// activation buffers, kBSlots ring slots, no clusters, every such choice a compile-time constant.  qtts_run_batched_ext
// (EXT = true) takes them from the launch (BRun::a_bufs, BRun::cluster).  One kernel with run-time switches cost the
// small groups 3 - 7 % (profiles/r02_summary.md section 14): the GEMM stage's issue path is that latency-critical.
template <bool EXT>
__device__ __forceinline__ void run_batched_body(const BRun & run_param) {
    extern __shared__ __align__(128) unsigned char bsm[];
    BShared * S = reinterpret_cast<BShared *>(bsm);

    for (int i = threadIdx.x; i < (int) (sizeof(BRun) / 4); i += kThreads)
        reinterpret_cast<uint32_t *>(&S->run)[i] = reinterpret_cast<const uint32_t *>(&run_param)[i];
    if (threadIdx.x == 0) {
        for (int i = 0; i < kBSlots; ++i) { mbar_init(S->full + i, 1); mbar_init(S->empty + i, kIssuers); }
        mbar_init(S->abar, kIssuers); mbar_init(S->abar + 1, kIssuers); mbar_init(&S->dbar, kIssuers);
        for (int i = 0; i < 4; ++i) {
            mbar_init(S->tfull + i, 1); mbar_init(S->tempty + i, (EXT && run_param.cluster) ? 2 : 1); S->n_fill[i] = 0; S->fill_base[i] = 0;
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        S->cons_seq = 0; S->a_use[0] = 0; S->a_use[1] = 0; S->d_use = 0; S->all_done = 0; S->rows_target = 0;
        BPrefetch & pf = S->pf;
        pf.idx = -1; pf.frame = 0; pf.t = 0; pf.p = 0; pf.n_tiles = 0; pf.n_pieces = 0; pf.tile_bytes = 0; pf.piece_bytes = 0; pf.msplit = 1;
        pf.W = nullptr; pf.seq = 0; pf.exhausted = (run_param.n_sched <= 0) ? 1 : 0;
        pf.policy = policy_evict_first();
        pf.slot = 0; pf.par = 1u; pf.paired = 0;
        pf.nsl = (EXT && QTTS_BRING_RUNTIME && run_param.a_bufs == 4) ? kBSlots - 3 : kBSlots;
        pf.ring = smem_u32(babuf_base(S)) + (uint32_t) ((EXT && QTTS_BRING_RUNTIME && run_param.a_bufs == 4) ? 4 : 2) * (uint32_t) kABufBytes;
    }
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(&S->tmem_base)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc_fence_before();
    __syncthreads();
    if (EXT && run_param.cluster) cluster_sync_all();      // the peer's mbarriers exist before anything is multicast to them
    tc_fence_after();
    if (threadIdx.x == kPfThread) bpf_poll(S);
    unsigned int epoch = 0;
    uint32_t next_word = 0;
    bool have_next = false;
    const int n_frames = run_param.n_frames, n_ops = run_param.n_ops, n_streams = run_param.n_streams;
    const bool prof = run_param.prof != nullptr && blockIdx.x == 0 && threadIdx.x == 0;

    for (int f = 0; f < n_frames; ++f) {
        if (run_param.check_done) {
            // uniform across the grid: the states were published by the previous frame's barriers
            if (threadIdx.x == 0) S->all_done = 1;
            __syncthreads();
            for (int b = threadIdx.x; b < n_streams; b += kThreads)
                if (run_param.active[b] && !__ldcg(&run_param.st[b].done)) S->all_done = 0;
            __syncthreads();
            if (S->all_done) break;
        }
        for (int i = 0; i < n_ops; ++i) {
            // stage descriptor -> shared memory.  It was requested before the previous stage's grid
            // barrier (see below), never across a call: a pending global load makes the first
            // non-inlined call of the stage wait a full L2 round trip.
            static_assert(sizeof(BOp) / 4 <= kThreads, "one word per thread");
            if (threadIdx.x < (int) (sizeof(BOp) / 4)) {
                if (!have_next) next_word = __ldg(reinterpret_cast<const uint32_t *>(run_param.ops + i) + threadIdx.x);
                reinterpret_cast<uint32_t *>(&S->op)[threadIdx.x] = next_word;
                have_next = true;
            }
            if (threadIdx.x == kThreads - 1) {
                S->tag = run_param.tag_base + 1u + (uint32_t) (f * n_ops + i);
                uint32_t * tr = nullptr;
                if (run_param.trace && (int) blockIdx.x == run_param.trace_cta && f * n_ops + i < run_param.trace_cap)
                    tr = run_param.trace + (size_t) (f * n_ops + i) * 16;
                S->tr = tr;
            }
            __syncthreads();
            const int type = S->op.type, in_kind = S->op.in_kind;
            if (threadIdx.x == 0 && S->tr) {
                S->tr[0] = (uint32_t) type | ((uint32_t) (in_kind & 1) << 4) | ((uint32_t) (S->op.flags & 0xff) << 8) | ((uint32_t) i << 16);
                S->tr[12] = (uint32_t) S->op.K; S->tr[13] = (uint32_t) S->op.rows; S->tr[14] = (uint32_t) f;
                S->tr[1] = (uint32_t) clock64();
            }
            const long long t0 = prof ? clock64() : 0;
            switch (type) {
                case BOP_GEMM:
                    if (in_kind == BIN_F32) bstage_gemm<BIN_F32, EXT>(S);
                    else                    bstage_gemm_tma<EXT>(S);
                    break;
                case BOP_ATTN_CP:
                    if (S->op.pair_f32) bstage_attn_cp<true>(S);
                    else if (S->run.dbg & 64) bstage_attn_cp<false>(S);        // A/B switch: the per-group form
                    else bstage_attn_cp_head(S);
                    break;
                case BOP_ATTN_TK:
                    if (S->op.pair_f32 == 2 && S->op.n_heads == 2 * S->op.n_kv_heads) {
                        if (EXT && S->run.a_bufs == 4 && !(S->run.dbg & 8192)) bstage_attn_tk_warp<true>(S);      // QTTS_BDBG=8192: K through registers (A/B)
                        else bstage_attn_tk_warp<false>(S);
                    }
                    else bstage_attn_tk(S);
                    break;
                case BOP_SAMPLE:  bstage_sample(S); break;
                case BOP_NORM:
                    // more rows than CTAs: one warp per row (one pass instead of two, 9.24 -> 9.13 ms per frame-step at 256
                    // streams); otherwise a CTA per row is the lower latency (a single warp was 6 % slower at 64 / 128 streams,
                    // profiles/r02z_normwarp_ab.log).  QTTS_BDBG=2048 forces the CTA form, 4096 the warp form.
                    if (EXT && (((int) S->op.rows_act > (int) gridDim.x && !(S->run.dbg & 2048)) || (S->run.dbg & 4096))) bstage_norm_warp(S);
                    else bstage_norm(S);
                    break;
            }
            const long long t1 = prof ? clock64() : 0;
            if (threadIdx.x == 0) BTR(8);
            if (threadIdx.x < (int) (sizeof(BOp) / 4)) {     // next stage's descriptor lands during the barrier
                const int ni = (i + 1 < n_ops) ? i + 1 : 0;
                next_word = __ldg(reinterpret_cast<const uint32_t *>(run_param.ops + ni) + threadIdx.x);
            }
            if (type == BOP_NORM && (S->op.flags & BF_ROWS_SIGNAL)) {
                if (threadIdx.x == kPfThread) bpf_poll(S);      // no barrier: the consumer waits for the row counter
                __syncthreads();
            } else {
                bgrid_barrier(S, epoch, (S->op.flags & BF_ASYNC_OUT) != 0 || (S->run.dbg & 32) != 0);
            }
            if (prof) {
                const long long t2 = clock64();
                const int slot = (type == BOP_NORM) ? 12 : type * 2 + ((type == BOP_GEMM && in_kind == BIN_F16) ? 8 : 0);
                run_param.prof[slot * 4 + 0] += 1;
                run_param.prof[slot * 4 + 1] += (unsigned long long) (t1 - t0);
                run_param.prof[slot * 4 + 2] += (unsigned long long) (t2 - t1);
            }
        }
    }
    // drain: every bulk copy issued into this CTA's shared memory must land before it exits
    __syncthreads();
    if (threadIdx.x == kPfThread) {
        for (uint32_t s = S->cons_seq; s != S->pf.seq; ++s)
            mbar_wait(S->full + (s % (uint32_t) S->pf.nsl), (s / (uint32_t) S->pf.nsl) & 1u);
    }
    tc_fence_before();
    __syncthreads();
    if (EXT && run_param.cluster) cluster_sync_all();      // no CTA exits while its peer can still signal its mbarriers
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(S->tmem_base), "r"(512));
    }
}

extern "C" __global__ void __launch_bounds__(kThreads, 1)
qtts_run_batched(const __grid_constant__ BRun run_param) { run_batched_body<false>(run_param); }

extern "C" __global__ void __launch_bounds__(kThreads, 1)
qtts_run_batched_ext(const __grid_constant__ BRun run_param) { run_batched_body<true>(run_param); }

}  // namespace qtts
