// qtts_kernels.cu — sm_100a device stages of the Qwen3-TTS codec-frame loop, batch-1 /
// small-batch regime (HBM-bound GEMV path).  See DESIGN.md for the stage list and rooflines.
//
// Structure: ONE persistent cooperative kernel (1 CTA per SM, 512 threads) walks a linear program
// of stage descriptors (Op) for n_frames codec frames, with a grid-wide barrier after every
// all-to-all stage.  Weight bytes never wait for a barrier: thread 0 of every CTA streams the
// CTA's row slices of the NEXT projections into a shared-memory ring with bulk asynchronous
// copies (cp.async.bulk + mbarrier complete_tx, the TMA engine), up to kNumSlots*slot_bytes
// ahead of execution and across stage / frame boundaries, so that by the time a barrier opens the
// stage's weights are already on chip and the critical path is barrier + input vector + FMAs.
//
// Reference behaviour each stage reproduces (file:line into the reference tree):
//   stage_gemv  : ggml_mul_mat with F16 weights and F32 activations rounded to F16
//                 (ggml/src/ggml-cpu/ggml-cpu.c:1229-1320) + ggml_rms_norm/ggml_mul prologue
//                 (ops.cpp:3709-3758; src/tts_transformer.cpp:1413-1414) + residual / SwiGLU
//                 epilogues (src/tts_transformer.cpp:1477-1493).
//   stage_attn  : q/k RMSNorm, NEOX RoPE, F16 KV append, attention
//                 (src/tts_transformer.cpp:1420-1475; ops.cpp:8108-8344 for the flash variant,
//                 :5224-5330 soft_max for the code-predictor / prefill variant).
//   stage_sample_*: src/tts_transformer.cpp:2198-2236 and :2417-2499.
#include "qtts_ops.cuh"
#include "qtts_nth.cuh"
#ifndef QTTS_OPT_RES
#define QTTS_OPT_RES 1
#endif
#ifndef QTTS_OPT_ATTN
#define QTTS_OPT_ATTN 1
#endif
#ifndef QTTS_OPT_FHFMA
#define QTTS_OPT_FHFMA 1          // GEMV dot on the mixed-precision FMA (0: convert + FFMA, for A/B runs)
#endif
#ifndef QTTS_OPT_SPLITK
#define QTTS_OPT_SPLITK 0         // 1: narrow projections (<= 8 row pairs per CTA): 2 or 4 warps share a row pair, each a K slice.
                                  // Measured SLOWER (2.834 -> 2.995 ms/frame, profiles/r02t_ab_splitk.log): the extra CTA barrier and the
                                  // shared-memory hand-over cost more than the shorter walk over weights that already sit in shared memory
#endif
#ifndef QTTS_OPT_ATTN_WARP
#define QTTS_OPT_ATTN_WARP 0      // 1: warp-per-head attention for <= 16 keys.  Measured SLOWER than the CTA-wide path
                                  // (3.005 vs 2.820 ms/frame, profiles/r02_ab.md), so it stays off
#endif
#include <math_constants.h>

namespace qtts {

// ------------------------------------------------------------------------------------------
// small helpers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float f16r(float v) { return __half2float(__float2half_rn(v)); }
// d = a * b + c with a, b F16 (low / high half of a 32-bit register: ptxas folds the selection into the .H1 operand
// modifier) and c, d F32, one rounding: the sm_100 mixed-precision FMA (SASS FHFMA).  Same real arithmetic as
// fmaf((float) a, (float) b, c) - both conversions are exact - at one instruction per multiply-add.
__device__ __forceinline__ float fhfma(uint16_t a, uint16_t b, float c) {
    float d;
    asm("fma.rn.f32.f16 %0, %1, %2, %3;" : "=f"(d) : "h"(a), "h"(b), "f"(c));
    return d;
}
__device__ __forceinline__ float fhfma2(uint32_t w, uint32_t x, float c) {      // c + w.lo * x.lo, then + w.hi * x.hi
#if !QTTS_OPT_FHFMA
    const float2 wf = __half22float2(*reinterpret_cast<const __half2 *>(&w)), xf = __half22float2(*reinterpret_cast<const __half2 *>(&x));
    return fmaf(wf.y, xf.y, fmaf(wf.x, xf.x, c));
#endif
    uint16_t w0, w1, x0, x1;
    asm("mov.b32 {%0, %1}, %2;" : "=h"(w0), "=h"(w1) : "r"(w));
    asm("mov.b32 {%0, %1}, %2;" : "=h"(x0), "=h"(x1) : "r"(x));
    return fhfma(w1, x1, fhfma(w0, x0, c));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// block-wide reductions through a small shared scratch (kWarps entries); all threads call.
__device__ double block_sum_d(double v, double * scratch) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    v = warp_sum_d(v);
    __syncthreads();
    if (lane == 0) scratch[warp] = v;
    __syncthreads();
    double t = 0.0;
    for (int i = 0; i < kWarps; ++i) t += scratch[i];   // fixed order: deterministic
    return t;
}
__device__ float block_max_f(float v, float * scratch) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) scratch[warp] = v;
    __syncthreads();
    float t = -CUDART_INF_F;
    for (int i = 0; i < kWarps; ++i) t = fmaxf(t, scratch[i]);
    return t;
}

// exp with the accuracy of a correctly rounded expf (glibc's expf is < 0.502 ulp): evaluate in
// double and round once.
__device__ __forceinline__ float expf_cr(float x) { return (float) exp((double) x); }

__device__ __forceinline__ float silu_f(float x) { return x / (1.0f + expf(-x)); }

// ------------------------------------------------------------------------------------------
// CTA-local copy of the mutable stream state.  Sampling and bookkeeping stages are executed by
// EVERY CTA redundantly (deterministic code, identical results), so no grid barrier is needed
// between a sampling stage and the stage that consumes the sampled id; CTA 0 mirrors the state
// to global memory for the host.
// ------------------------------------------------------------------------------------------
struct LocalState {
    int32_t n_past, frame, done, eos;
    int32_t cur_codes[kMaxBooks];
    int32_t feed_codes[kMaxBooks];
    uint32_t seen[kMaxVocab / 32];
};

struct RunCfg {
    int32_t slot_bytes;     // bytes per ring slot (multiple of 1024, >= 2 weight rows)
    int32_t n_slots;        // ring slots in use (<= kNumSlots)
    int32_t xs_elems;       // floats in the input-vector area
    int32_t persistent;     // 1: cooperative launch with grid barriers; 0: one op per launch
    int32_t check_done;
    int32_t prof_cta;
    int32_t n_sched;        // GEMV stages per frame
    const WSched * sched;   // [n_sched]
    uint32_t tag_base;      // stage tags of this launch are tag_base + 1 + (frame * n_ops + op index)
    unsigned long long * prof;   // optional [8 op types][8] cycle counters filled by CTA 0 (nullptr: off)
    uint32_t * trace;            // optional stage timeline of CTA prof_cta, thread 0: 8 words per stage, store-only (no
    int32_t trace_cap;           //   read-modify-write on the critical path): [type | op << 8, start, own poll done,
                                 //   input staged, rows done, stage end (after the barrier if any), frame, spare]
};

// ------------------------------------------------------------------------------------------
// mbarrier + bulk copy (TMA engine, 1-D) wrappers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t * bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t * bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t * bar, uint32_t parity) {
    const uint32_t a = smem_u32(bar);
    uint32_t ok = 0;
    while (!ok) {
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(a), "r"(parity) : "memory");
    }
}
// global -> shared bulk copy, completion counted in bytes on `bar`; weights are streamed once
// per use, so they are marked evict-first in L2.
__device__ __forceinline__ void bulk_g2s(void * dst, const void * src, uint32_t bytes, uint64_t * bar, uint64_t policy) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy) : "memory");
}
__device__ __forceinline__ uint64_t policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}

// ------------------------------------------------------------------------------------------
// the weight ring: every GEMV op contributes, per CTA, ceil(pairs_cta / pairs_per_chunk) chunks
// (a chunk = whole row pairs, <= slot_bytes); chunk number `seq` (counted from kernel start)
// lives in slot seq % kNumSlots and completes phase (seq / kNumSlots) & 1 of that slot's mbarrier.
// ------------------------------------------------------------------------------------------
struct Ring {
    unsigned char * base;
    uint64_t * full;        // [kNumSlots]
    int32_t slot_bytes;
    int32_t n_slots;
};

constexpr int kPfThread = kThreads - 32;   // lane 0 of the last warp issues the bulk copies

// Prefetcher state.  It lives in SHARED memory: only one thread uses it, and as a local variable of
// the kernel it cost every thread ~20 registers (spilled around the stages; with ~225 KB of shared
// memory there is no L1 to catch spills).  Upcoming schedule entries are pulled into `ent` by
// cp.async two entries ahead, so advancing never waits for an L2 round trip.
struct Prefetch {
    int32_t idx, frame;     // schedule entry, frame being prefetched
    int32_t q, p1;          // next row pair / end of this CTA's slice of the entry
    int32_t K, ppc, slot;
    uint32_t seq;
    const wt_t * W;
    uint64_t policy;
    int4 ent[4][2];         // schedule entries idx .. idx+2 (slot = entry index & 3)
};

__device__ __forceinline__ void sched_fetch(Prefetch & pf, const WSched * __restrict__ sched, int n_sched, int k) {
    // k = running entry counter; entry index = k % n_sched
    const WSched * src = sched + (k % n_sched);
    const uint32_t dst = smem_u32(&pf.ent[k & 3][0]);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" :: "r"(dst), "l"(src) : "memory");
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" :: "r"(dst + 16), "l"(reinterpret_cast<const char *>(src) + 16) : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
}

__device__ __forceinline__ void cta_pairs(const Op & op, int & p0, int & p1) {
    const unsigned pairs = (unsigned) op.rows >> 1;
    p0 = (int) (pairs * blockIdx.x / gridDim.x);
    p1 = (int) (pairs * (blockIdx.x + 1) / gridDim.x);
}

// prefetcher thread only: issue bulk copies for upcoming chunks while ring slots are free.
// pf.idx counts entries monotonically (entry = idx % n_sched, frame = idx / n_sched).
__device__ __noinline__ void prefetch_issue(const WSched * __restrict__ sched, int n_sched, int n_frames, Prefetch & pf,
                                            uint32_t cons_seq, const Ring ring) {
    while (pf.seq - cons_seq < (uint32_t) ring.n_slots && pf.frame < n_frames) {
        if (pf.q >= pf.p1) {
            const int k = ++pf.idx;
            if (k >= (pf.frame + 1) * n_sched) { if (++pf.frame >= n_frames) break; }
            asm volatile("cp.async.wait_group 1;" ::: "memory");      // entry k landed (k+1 may be in flight)
            const int4 e0 = pf.ent[k & 3][0], e1 = pf.ent[k & 3][1];
            sched_fetch(pf, sched, n_sched, k + 2);
            pf.W = reinterpret_cast<const wt_t *>(((unsigned long long) (unsigned) e0.y << 32) | (unsigned) e0.x);
            pf.K = e0.z;
            const unsigned pairs = (unsigned) e0.w;
            pf.ppc = e1.x;
            pf.q = (int) (pairs * blockIdx.x / gridDim.x);
            pf.p1 = (int) (pairs * (blockIdx.x + 1) / gridDim.x);
            continue;
        }
        const int nq = min(pf.ppc, pf.p1 - pf.q);
        const uint32_t bytes = (uint32_t) nq * 4u * (uint32_t) pf.K;
        const int slot = pf.slot;
        mbar_expect_tx(ring.full + slot, bytes);
        bulk_g2s(ring.base + (size_t) slot * ring.slot_bytes, pf.W + (size_t) 2 * pf.q * pf.K, bytes,
                 ring.full + slot, pf.policy);
        pf.q += nq;
        pf.seq++;
        pf.slot = (slot + 1 == ring.n_slots) ? 0 : slot + 1;
    }
}

// ------------------------------------------------------------------------------------------
// grid barrier (cooperative launch: all CTAs co-resident).  Between arriving and polling,
// thread 0 refills the weight ring, so the copy engine works through the barrier latency.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void grid_barrier(GridBarrier * bar, unsigned int & epoch,
                                             const WSched * sched, int n_sched, int n_frames, Prefetch & pf,
                                             uint32_t cons_seq, const Ring & ring, unsigned long long * prof_sync = nullptr) {
    const long long ts = prof_sync ? clock64() : 0;
    __syncthreads();
    if (threadIdx.x == 0) {
        if (prof_sync) *prof_sync += (unsigned long long) (clock64() - ts);
        const unsigned int target = epoch + gridDim.x;
        epoch = target;
        __threadfence();
        atomicAdd(&bar->count, 1u);
        // relaxed polling (an acquire load invalidates L1 on every iteration), one acquire at the end;
        // activations are read with ld.cg (L2), constants with ld.nc, so nothing stale can be hit
        unsigned int v;
        do {
            asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(&bar->count) : "memory");
        } while ((int) (v - target) < 0);
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(&bar->count) : "memory");
    } else if (threadIdx.x == kPfThread) {
        prefetch_issue(sched, n_sched, n_frames, pf, cons_seq, ring);
    }
    __syncthreads();
}

// ------------------------------------------------------------------------------------------
// OP_GEMV
// ------------------------------------------------------------------------------------------
constexpr int kMaxPerThread = 12;     // input elements per thread: K <= kThreads * 12 = 6144

// Flag-carrying activations: a single 64-bit relaxed store / load is single-copy atomic, so a
// consumer that sees the producer stage's tag in the upper half has the value in the lower half.
__device__ __forceinline__ void ll_store(unsigned long long * p, float v, uint32_t tag) {
    const unsigned long long w = ((unsigned long long) tag << 32) | (unsigned long long) __float_as_uint(v);
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" :: "l"(p), "l"(w) : "memory");
}
__device__ __forceinline__ unsigned long long ll_load(const unsigned long long * p) {
    unsigned long long w;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
    return w;
}
// a protocol error must not hang the device: ~2^22 polls (seconds) then trap
__device__ __forceinline__ void ll_spin_guard(uint32_t & spins) {
    if (++spins > (1u << 22)) __trap();
}
__device__ __forceinline__ float ll_wait(const unsigned long long * p, uint32_t tag) {
    unsigned long long w = ll_load(p);
    uint32_t spins = 0;
    while ((uint32_t) (w >> 32) != tag) { ll_spin_guard(spins); w = ll_load(p); }
    return __uint_as_float((uint32_t) w);
}

// raw input element i of token t (before the norm)
__device__ __forceinline__ float gemv_input_elem(const Op & op, const LocalState * ls, const __half * row,
                                                 const float * trow, int t, int i) {
    const int K = op.K;
    switch (op.in_kind) {
        case IN_VEC:       return __ldcg(op.in + (size_t) t * op.in_stride + i);
        case IN_EMBED_ROW: return __half2float(row[i]);
        case IN_CP_FIRST:  return (t == 0) ? __ldcg(op.in + i) : __half2float(row[i]);
        default: {
            // next Talker input (src/tts_transformer.cpp:2529-2553): sequential f32 sum, same order
            const EmbedSumArgs * es = op.es;
            float acc = __half2float(es->tables[0][(size_t) ls->feed_codes[0] * K + i]);
            for (int c = 1; c < es->n_books; ++c)
                acc += __half2float(es->tables[c][(size_t) ls->feed_codes[c] * K + i]);
            return acc + __ldcg(trow + i);
        }
    }
}

// 8 weights (F16, one uint4) x 8 activations (F16, one uint4), F32 accumulate in element order: 8 FHFMA, no unpacking
__device__ __forceinline__ float dot8h(const uint4 w, const uint4 x) {
    float s = fhfma2(w.x, x.x, 0.f);
    s = fhfma2(w.y, x.y, s);
    s = fhfma2(w.z, x.z, s);
    s = fhfma2(w.w, x.w, s);
    return s;
}

// rows are handled as pairs (2p, 2p+1) — EPI_SWIGLU needs exactly that pairing; the pair's
// weights are read from the shared-memory ring (chunk = cons_seq + q / pairs_per_chunk), the
// F16-rounded activations from xh.  Residual / bias values are requested before the dot products.
// Narrow projections (O, down, lm_head: <= 8 row pairs per CTA) left 8 - 13 of the 16 warps without rows while 3 - 4 warps
// walked the whole K (14 % of the kernel's stall samples sat at the next stage-top barrier, profiles/r02s_frame_ncu.md).
// Here S = 2 or 4 adjacent warps share a row pair: warp w takes pair w / S and the K slice w % S, the partial sums meet
// in shared memory (`part`, kWarps x 2 floats) and slice 0's warp adds them in fixed order and runs the epilogue.
// One-token stages only.  All threads of the CTA must call (one CTA barrier inside).
__device__ __forceinline__ void gemv_rows_split(const Op & op, const __half * xh, const Ring & ring, uint32_t cons_seq, int p0,
                                                int q_begin, int q_end, int S, float * part, uint32_t tag, bool ll) {
    const int K = op.K, kv = K >> 3;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ppc = ring.slot_bytes / (4 * K);
    const uint4 * x4 = reinterpret_cast<const uint4 *>(xh);
    const int qi = warp / S, ks = warp - qi * S;
    const int q = q_begin + qi, p = p0 + q;
    const bool active = q < q_end;
    float side = 0.f;
    const unsigned long long * side_p = nullptr;
    unsigned long long side_w = 0;
    if (active && ks == 0 && lane < 2) {
        const int r = 2 * p + lane;
        if (op.epi == EPI_ADD) {
            if (ll && op.res_ll) { side_p = op.res_ll + r; side_w = ll_load(side_p); }
            else side = __ldcg(op.res + r);
        }
        else if (op.epi == EPI_BIAS || op.epi == EPI_BIAS_SILU) side = __ldg(op.bias + r);
    }
    if (active) {
        const int j = q / ppc;
        const uint32_t seq = cons_seq + (uint32_t) j;
        const int slot = (int) (seq % (uint32_t) ring.n_slots);
        mbar_wait(ring.full + slot, (seq / (uint32_t) ring.n_slots) & 1u);
        const uint4 * w0 = reinterpret_cast<const uint4 *>(ring.base + (size_t) slot * ring.slot_bytes +
                                                            (size_t) (q - j * ppc) * 4 * K);
        const uint4 * w1 = w0 + kv;
        const int per = (kv + S - 1) / S, lo = ks * per, hi = min(kv, lo + per);
        float a0 = 0.f, a1 = 0.f;
        // a K slice is <= 3 uint4 per lane and the weights are already in shared memory: no load batching needed
#pragma unroll 1
        for (int i = lo + lane; i < hi; i += 32) {
            const uint4 xv = x4[i];
            a0 += dot8h(w0[i], xv);
            a1 += dot8h(w1[i], xv);
        }
        a0 = warp_sum(a0); a1 = warp_sum(a1);
        if (lane == 0) { part[2 * warp] = a0; part[2 * warp + 1] = a1; }
    }
    __syncthreads();
    if (active && ks == 0) {
        if (op.epi == EPI_SWIGLU) {
            if (lane == 0) {
                float g = 0.f, u = 0.f;
                for (int i = 0; i < S; ++i) { g += part[2 * (warp + i)]; u += part[2 * (warp + i) + 1]; }
                const float v = silu_f(g) * u;
                op.out[p] = v;
                if (ll && op.out_ll) ll_store(op.out_ll + p, v, tag);
            }
        } else if (lane < 2) {
            const int r = 2 * p + lane;
            float acc = 0.f;
            for (int i = 0; i < S; ++i) acc += part[2 * (warp + i) + lane];
            if (side_p) {
                const uint32_t want = tag - (uint32_t) op.res_dist;
                uint32_t spins = 0;
                while ((uint32_t) (side_w >> 32) != want) { ll_spin_guard(spins); side_w = ll_load(side_p); }
                side = __uint_as_float((uint32_t) side_w);
            }
            acc += side;
            if (op.epi == EPI_BIAS_SILU) acc = silu_f(acc);
            op.out[r] = acc;
            if (ll && op.out_ll) ll_store(op.out_ll + r, acc, tag);
        }
    }
}

template <int NT>
__device__ __forceinline__ void gemv_rows(const Op & op, const __half * xh, const Ring ring, uint32_t cons_seq, int p0,
                          int q_begin, int q_end, unsigned long long * prof_wait, uint32_t tag, bool ll) {
    const int K = op.K, kv = K >> 3;      // uint4 per row
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ppc = ring.slot_bytes / (4 * K);
    const uint4 * x4 = reinterpret_cast<const uint4 *>(xh);
    for (int q = q_begin + warp; q < q_end; q += kWarps) {
        const int p = p0 + q;
        float side = 0.f;       // residual or bias of (token lane>>1, row 2p + (lane&1))
        const unsigned long long * side_p = nullptr;     // flag-carrying residual: requested now, checked after the dots
        unsigned long long side_w = 0;
        if (lane < 2 * NT) {
            const int t = lane >> 1, r = 2 * p + (lane & 1);
            if (op.epi == EPI_ADD) {
#if QTTS_OPT_RES
                if (ll && op.res_ll) { side_p = op.res_ll + (size_t) t * op.out_stride + r; side_w = ll_load(side_p); }
#else
                if (ll && op.res_ll) side = ll_wait(op.res_ll + (size_t) t * op.out_stride + r, tag - (uint32_t) op.res_dist);
#endif
                else side = __ldcg(op.res + (size_t) t * op.out_stride + r);
            }
            else if (op.epi == EPI_BIAS || op.epi == EPI_BIAS_SILU) side = __ldg(op.bias + r);
        }
        const int j = q / ppc;
        const uint32_t seq = cons_seq + (uint32_t) j;
        const int slot = (int) (seq % (uint32_t) ring.n_slots);
        if (prof_wait && threadIdx.x == 0) {
            const long long tw = clock64();
            mbar_wait(ring.full + slot, (seq / (uint32_t) ring.n_slots) & 1u);
            *prof_wait += (unsigned long long) (clock64() - tw);
        } else {
            mbar_wait(ring.full + slot, (seq / (uint32_t) ring.n_slots) & 1u);
        }
        const uint4 * w0 = reinterpret_cast<const uint4 *>(ring.base + (size_t) slot * ring.slot_bytes +
                                                            (size_t) (q - j * ppc) * 4 * K);
        const uint4 * w1 = w0 + kv;
        float a0[NT], a1[NT];
#pragma unroll
        for (int t = 0; t < NT; ++t) { a0[t] = 0.f; a1[t] = 0.f; }
        for (int j0 = 0; j0 < kv; j0 += 128) {
            uint4 A[4], B[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = j0 + u * 32 + lane;
                if (i < kv) { A[u] = w0[i]; B[u] = w1[i]; }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = j0 + u * 32 + lane;
                if (i < kv) {
#pragma unroll
                    for (int t = 0; t < NT; ++t) {
                        const uint4 xv = x4[t * kv + i];
                        a0[t] += dot8h(A[u], xv);
                        a1[t] += dot8h(B[u], xv);
                    }
                }
            }
        }
#pragma unroll
        for (int t = 0; t < NT; ++t) { a0[t] = warp_sum(a0[t]); a1[t] = warp_sum(a1[t]); }
        if (op.epi == EPI_SWIGLU) {
            if (lane < NT) {
                const float g = (lane == 0) ? a0[0] : a0[NT - 1], u = (lane == 0) ? a1[0] : a1[NT - 1];
                const float v = silu_f(g) * u;
                op.out[(size_t) lane * op.out_stride + p] = v;
                if (ll && op.out_ll) ll_store(op.out_ll + (size_t) lane * op.out_stride + p, v, tag);
            }
        } else if (lane < 2 * NT) {
            const int t = lane >> 1, r = 2 * p + (lane & 1);
            float acc = (lane & 1) ? ((t == 0) ? a1[0] : a1[NT - 1]) : ((t == 0) ? a0[0] : a0[NT - 1]);
            if (side_p) {
                const uint32_t want = tag - (uint32_t) op.res_dist;
                uint32_t spins = 0;
                while ((uint32_t) (side_w >> 32) != want) { ll_spin_guard(spins); side_w = ll_load(side_p); }
                side = __uint_as_float((uint32_t) side_w);
            }
            acc += side;
            if (op.epi == EPI_BIAS_SILU) acc = silu_f(acc);
            op.out[(size_t) t * op.out_stride + r] = acc;
            if (ll && op.out_ll) ll_store(op.out_ll + (size_t) t * op.out_stride + r, acc, tag);
        }
    }
}

// input vector of token t -> F16-rounded activations in shared memory.  Every thread owns elements
// tid + j*kThreads (j < NPER); the input and the norm weights are requested together (one L2
// round trip) and the RMSNorm runs out of registers.
template <int NPER, bool LL>
__device__ __forceinline__ void gemv_prepare(const Op & op, const LocalState * ls, const __half * row, const float * trow,
                                             int t, __half * xh, double * red, uint32_t tag_in, uint32_t tag_x, uint32_t * tr) {
    const int K = op.K;
    float xv[NPER], nw[NPER];
    if (LL) {
        // poll the producer stage's flag-carrying output: all of this thread's words are requested
        // together, the round is repeated until every tag matches
        const unsigned long long * src = op.in_ll + (size_t) t * op.in_stride;
#pragma unroll
        for (int j = 0; j < NPER; ++j) {
            const int i = threadIdx.x + j * kThreads;
            nw[j] = (op.norm_w && i < K) ? __ldg(op.norm_w + i) : 1.f;
            xv[j] = 0.f;
        }
        uint32_t spins = 0;
        constexpr int kB = NPER < 6 ? NPER : 6;     // words in flight per polling round
#pragma unroll
        for (int j0 = 0; j0 < NPER; j0 += kB) {
            bool ok;
            do {
                unsigned long long w[kB];
#pragma unroll
                for (int j = 0; j < kB; ++j) {
                    const int i = threadIdx.x + (j0 + j) * kThreads;
                    w[j] = (i < K) ? ll_load(src + i) : ((unsigned long long) tag_in << 32);
                }
                ok = true;
#pragma unroll
                for (int j = 0; j < kB; ++j) {
                    ok = ok && ((uint32_t) (w[j] >> 32) == tag_in);
                    xv[j0 + j] = __uint_as_float((uint32_t) w[j]);
                }
                if (!ok) ll_spin_guard(spins);
            } while (!ok);
        }
    } else {
#pragma unroll
        for (int j = 0; j < NPER; ++j) {
            const int i = threadIdx.x + j * kThreads;
            xv[j] = 0.f; nw[j] = 1.f;
            if (i < K) {
                xv[j] = gemv_input_elem(op, ls, row, trow, t, i);
                if (op.norm_w) nw[j] = __ldg(op.norm_w + i);
            }
        }
    }
    if (tr) tr[2] = (uint32_t) clock64();
    if (op.x_out && blockIdx.x == 0) {
#pragma unroll
        for (int j = 0; j < NPER; ++j) {
            const int i = threadIdx.x + j * kThreads;
            if (i < K) {
                op.x_out[t * K + i] = xv[j];
                if (tag_x && op.x_out_ll) ll_store(op.x_out_ll + t * K + i, xv[j], tag_x);
            }
        }
    }
    float scale = 1.0f;
    if (op.norm_w) {
        // ggml_compute_forward_rms_norm_f32: double sum of float squares (ops.cpp:3738-3747)
        double ss = 0.0;
#pragma unroll
        for (int j = 0; j < NPER; ++j) ss += (double) (xv[j] * xv[j]);
        ss = block_sum_d(ss, red);
        const float mean = (float) (ss / (double) K);
        scale = 1.0f / sqrtf(mean + op.eps);
    }
#pragma unroll
    for (int j = 0; j < NPER; ++j) {
        const int i = threadIdx.x + j * kThreads;
        if (i < K) {
            float v = xv[j];
            if (op.norm_w) {
                v = (v * scale) * nw[j];
                if (op.hid_out && blockIdx.x == 0) op.hid_out[t * K + i] = v;
            }
            xh[t * K + i] = __float2half_rn(v);     // ggml_mul_mat rounds the activation to F16
        }
    }
}

// returns the number of ring chunks this CTA consumed
__device__ __forceinline__ uint32_t stage_gemv(const Op & op, StreamState * st, LocalState * ls, float * xs, double * red,
                               const Ring & ring, uint32_t cons_seq,
                               const WSched * sched, int n_sched, int n_frames, Prefetch & pf, unsigned long long * prof,
                               uint32_t tag, bool ll, bool pf_pending, uint32_t * tr) {
    const int K = op.K;
    const bool ll_in = ll && op.in_ll != nullptr && op.in_kind == IN_VEC;
    const uint32_t tag_in = tag - (uint32_t) op.in_dist;
    const bool prof_on = prof && threadIdx.x == 0;
    long long tA = prof_on ? clock64() : 0;
    __half * xh = reinterpret_cast<__half *>(xs);

    // input row pointers (token ids are CTA-local state or device memory written before the launch)
    const __half * row = nullptr;
    const float * trow = nullptr;
    if (op.in_kind == IN_EMBED_ROW) {
        const int tok = (op.tok_slot >= 0) ? ls->feed_codes[op.tok_slot] : __ldcg(op.tok_ptr);
        row = op.table + (size_t) tok * K;
    } else if (op.in_kind == IN_CP_FIRST) {
        row = op.table + (size_t) ls->feed_codes[0] * K;
    } else if (op.in_kind == IN_EMBED_SUM) {
        const int frame = ls->frame - 1;            // frame already counted by OP_FRAME_END
        trow = (frame < st->trailing_len) ? op.es->trailing + (size_t) frame * K : op.es->pad;
    }
    const int nper = (K + kThreads - 1) / kThreads;
    for (int t = 0; t < op.n_tok; ++t) {
        if (ll_in) {
            if (nper <= 2)      gemv_prepare<2, true>(op, ls, row, trow, t, xh, red, tag_in, ll ? tag : 0u, tr);
            else if (nper <= 4) gemv_prepare<4, true>(op, ls, row, trow, t, xh, red, tag_in, ll ? tag : 0u, tr);
            else if (nper <= 6) gemv_prepare<6, true>(op, ls, row, trow, t, xh, red, tag_in, ll ? tag : 0u, tr);
            else                gemv_prepare<kMaxPerThread, true>(op, ls, row, trow, t, xh, red, tag_in, ll ? tag : 0u, tr);
        } else {
            if (nper <= 2)      gemv_prepare<2, false>(op, ls, row, trow, t, xh, red, tag_in, ll ? tag : 0u, tr);
            else if (nper <= 4) gemv_prepare<4, false>(op, ls, row, trow, t, xh, red, tag_in, ll ? tag : 0u, tr);
            else if (nper <= 6) gemv_prepare<6, false>(op, ls, row, trow, t, xh, red, tag_in, ll ? tag : 0u, tr);
            else                gemv_prepare<kMaxPerThread, false>(op, ls, row, trow, t, xh, red, tag_in, ll ? tag : 0u, tr);
        }
    }
    __syncthreads();
    if (tr) tr[3] = (uint32_t) clock64();
    if (pf_pending && threadIdx.x == kPfThread) prefetch_issue(sched, n_sched, n_frames, pf, cons_seq, ring);
    if (prof_on) { const long long t = clock64(); prof[OP_GEMV * 8 + 4] += (unsigned long long) (t - tA); tA = t; }

    int p0, p1;
    cta_pairs(op, p0, p1);
    const int ppc = ring.slot_bytes / (4 * K);
    const int nch = (p1 - p0 + ppc - 1) / ppc;
    // a slice longer than the ring is consumed in rounds of n_slots chunks; between rounds the
    // freed slots are refilled
    for (int c0 = 0; c0 < nch; c0 += ring.n_slots) {
        const int c1 = min(c0 + ring.n_slots, nch);
        const int qb = c0 * ppc, qe = min(c1 * ppc, p1 - p0);
        int split = 1;
#if QTTS_OPT_SPLITK
        if (op.n_tok == 1 && nch <= ring.n_slots && (K & 255) == 0) {
            if (4 * (qe - qb) <= kWarps) split = 4; else if (2 * (qe - qb) <= kWarps) split = 2;
        }
#endif
        if (split > 1)          gemv_rows_split(op, xh, ring, cons_seq, p0, qb, qe, split, reinterpret_cast<float *>(red), tag, ll);
        else if (op.n_tok == 2) gemv_rows<2>(op, xh, ring, cons_seq, p0, qb, qe, prof_on ? prof + OP_GEMV * 8 + 6 : nullptr, tag, ll);
        else                    gemv_rows<1>(op, xh, ring, cons_seq, p0, qb, qe, prof_on ? prof + OP_GEMV * 8 + 6 : nullptr, tag, ll);
        if (c1 < nch) {
            __syncthreads();
            if (threadIdx.x == kPfThread) prefetch_issue(sched, n_sched, n_frames, pf, cons_seq + (uint32_t) c1, ring);
        }
    }
    if (op.flags & F_INC_NPAST) {
        if (threadIdx.x == 0) {
            ls->n_past += 1;
            if (blockIdx.x == 0) st->n_past = ls->n_past;
        }
    }
    if (prof_on) { const long long t = clock64(); prof[OP_GEMV * 8 + 5] += (unsigned long long) (t - tA); }
    if (tr) tr[4] = (uint32_t) clock64();
    return (uint32_t) nch;
}

// ------------------------------------------------------------------------------------------
// OP_ATTN
// ------------------------------------------------------------------------------------------
constexpr int kMaxRep = 4;      // query heads per KV head handled by one CTA in ATTN_FAST
constexpr int kTile   = 512;    // keys per tile of the sequential flash emulation

struct AttnSmem {
    float q[kMaxRep][kHeadDim];   // queries after norm+rope (rounded to F16 unless ATTN_F32)
    float kf[kHeadDim];           // current key, unrounded
    float vf[kHeadDim];           // current value, unrounded
    float k0f[kHeadDim];          // ATTN_F32: key / value of position 0, unrounded
    float v0f[kHeadDim];
    __half k16[kHeadDim];
    __half v16[kHeadDim];
    float tmp[(kMaxRep + 2) * kHeadDim];   // raw q heads, k, v staging
    double red[kWarps];
    float redf[kWarps];
    float part[4][kMaxRep][kHeadDim];      // partial P.V sums per key partition
    float chunkM[64], chunkS[64];
    float hm[kMaxRep], hl[kMaxRep];
    int   flag;
};

// q (nq heads from h_first), k, v of one token: per-head RMSNorm (double sum), NEOX RoPE at `pos`.
// A thread owns elements tid and tid + kThreads of the (nq + 2) x 128 staging block; its QK-norm
// weights and RoPE factors are requested BEFORE the projections are read (polled), so the stage
// has one L2 round trip on its critical path instead of three.
__device__ void attn_prepare(const Op & op, const float * qkv, int h_first, int nq, int g, int pos, AttnSmem * sm,
                             const unsigned long long * qkv_ll = nullptr, uint32_t tag_in = 0) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int rows = nq + 2;
    static_assert((kMaxRep + 2) * kHeadDim <= 2 * kThreads, "two staging elements per thread");
    float nw[2], rc[2], rs[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
        const int idx = tid + u * kThreads;
        const int row = idx / kHeadDim, d = idx % kHeadDim;
        nw[u] = 1.f; rc[u] = 1.f; rs[u] = 0.f;
        if (QTTS_OPT_ATTN && row <= nq) {
            // QK-norm only when the layer has the weight (src/tts_transformer.cpp:1424-1432)
            const float * w = (row < nq) ? op.q_norm_w : op.k_norm_w;
            if (w) nw[u] = __ldg(w + d);
            const float2 cs = __ldg(reinterpret_cast<const float2 *>(op.rope) + (size_t) pos * 64 + (d & 63));
            rc[u] = cs.x; rs[u] = cs.y;
        }
    }
    __syncthreads();
#pragma unroll
    for (int u = 0; u < 2; ++u) {
        const int idx = tid + u * kThreads;
        if (idx < rows * kHeadDim) {
            const int row = idx / kHeadDim, d = idx % kHeadDim;
            const float * src = (row < nq)  ? qkv + (size_t) (h_first + row) * kHeadDim
                              : (row == nq) ? qkv + (size_t) (op.n_heads + g) * kHeadDim
                                            : qkv + (size_t) (op.n_heads + op.n_kv_heads + g) * kHeadDim;
            sm->tmp[idx] = qkv_ll ? ll_wait(qkv_ll + (src - qkv) + d, tag_in) : __ldcg(src + d);
        }
    }
    __syncthreads();
    if (warp <= nq) {       // rows 0..nq (queries and the key) are normalised
        double ss = 0.0;
#pragma unroll
        for (int u = 0; u < 4; ++u) { const float v = sm->tmp[warp * kHeadDim + u * 32 + lane]; ss += (double) (v * v); }
        ss = warp_sum_d(ss);
        if (lane == 0) sm->red[warp] = ss;
    }
    __syncthreads();
#pragma unroll
    for (int u = 0; u < 2; ++u) {
        const int idx = tid + u * kThreads;
        if (idx < (nq + 1) * kHeadDim) {
            const int row = idx / kHeadDim;
            const float * w = (row < nq) ? op.q_norm_w : op.k_norm_w;
            if (w) {
                const float mean = (float) (sm->red[row] / (double) kHeadDim);
                const float scale = 1.0f / sqrtf(mean + op.eps);
                sm->tmp[idx] = (sm->tmp[idx] * scale) * (QTTS_OPT_ATTN ? nw[u] : w[idx % kHeadDim]);
            }
        }
    }
    __syncthreads();
#pragma unroll
    for (int u = 0; u < 2; ++u) {
        const int idx = tid + u * kThreads;
        if (idx < rows * kHeadDim) {
            const int row = idx / kHeadDim, d = idx % kHeadDim;
            if (row <= nq) {
                const int i = d & 63;
                const float c = QTTS_OPT_ATTN ? rc[u] : op.rope[((size_t) pos * 64 + i) * 2];
                const float sn = QTTS_OPT_ATTN ? rs[u] : op.rope[((size_t) pos * 64 + i) * 2 + 1];
                const float x0 = sm->tmp[row * kHeadDim + i], x1 = sm->tmp[row * kHeadDim + i + 64];
                const float r = (d < 64) ? (x0 * c - x1 * sn) : (x0 * sn + x1 * c);
                if (row < nq) sm->q[row][d] = (op.mode == ATTN_F32) ? r : f16r(r);
                else { sm->kf[d] = r; sm->k16[d] = __float2half_rn(r); }
            } else {
                const float v = sm->tmp[idx];
                sm->vf[d] = v;
                sm->v16[d] = __float2half_rn(v);
            }
        }
    }
    __syncthreads();
}

__device__ __forceinline__ void attn_append(const Op & op, int g, int pos, int kvw, const AttnSmem * sm) {
    if (threadIdx.x < kHeadDim) {
        op.kcache[(size_t) pos * kvw + g * kHeadDim + threadIdx.x] = sm->k16[threadIdx.x];
        op.vcache[(size_t) pos * kvw + g * kHeadDim + threadIdx.x] = sm->v16[threadIdx.x];
    }
}

// dot of 16 consecutive dims (2 x uint4 of halves) with q registers
__device__ __forceinline__ float dot16h(const uint4 a, const uint4 b, const float * q) {
    const __half2 * ha = reinterpret_cast<const __half2 *>(&a);
    const __half2 * hb = reinterpret_cast<const __half2 *>(&b);
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float2 f = __half22float2(ha[i]);
        s = fmaf(f.x, q[2 * i], s);
        s = fmaf(f.y, q[2 * i + 1], s);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float2 f = __half22float2(hb[i]);
        s = fmaf(f.x, q[8 + 2 * i], s);
        s = fmaf(f.y, q[8 + 2 * i + 1], s);
    }
    return s;
}

// scores sc[r][j - j0] = scale * <K16_j, q_r> for keys j0..j1 (8 lanes per key, 16 dims per lane);
// key `pos` (the token being appended) is read from shared memory
template <int REP>
__device__ void attn_scores(const Op & op, const AttnSmem * sm, int g, int kvw, int pos, int j0, int j1,
                            float * sc, int sc_stride) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = lane & 7, kin = lane >> 3;
    float qreg[REP][16];
#pragma unroll
    for (int r = 0; r < REP; ++r)
#pragma unroll
        for (int i = 0; i < 16; ++i) qreg[r][i] = sm->q[r][sub * 16 + i];
    for (int jb = j0 + warp * 4; jb < j1; jb += kWarps * 4) {
        const int j = jb + kin;
        uint4 ka = make_uint4(0, 0, 0, 0), kb = ka;
        if (j < j1) {
            if (j < pos) {
                const uint4 * k4 = reinterpret_cast<const uint4 *>(op.kcache + (size_t) j * kvw + g * kHeadDim + sub * 16);
                ka = __ldcg(k4); kb = __ldcg(k4 + 1);
            } else {
                const uint4 * k4 = reinterpret_cast<const uint4 *>(sm->k16 + sub * 16);
                ka = k4[0]; kb = k4[1];
            }
        }
#pragma unroll
        for (int r = 0; r < REP; ++r) {
            float s = dot16h(ka, kb, qreg[r]);
            s += __shfl_xor_sync(0xffffffffu, s, 1);
            s += __shfl_xor_sync(0xffffffffu, s, 2);
            s += __shfl_xor_sync(0xffffffffu, s, 4);
            if (sub == 0 && j < j1) sc[r * sc_stride + (j - j0)] = s * op.kq_scale;
        }
    }
}

// ---- ATTN_FAST: split-KV over (KV group, key range) CTAs, REP query heads per CTA; the last CTA
// of a group to finish merges the partials (fixed order -> deterministic)
template <int REP>
__device__ __noinline__ void attn_fast(const Op & op, const float * qkv, float * out, int pos, AttnSmem * sm, float * sc,
                                       const unsigned long long * qkv_ll, uint32_t tag_in) {
    const int tid = threadIdx.x;
    const int n_kv = pos + 1, kvw = op.n_kv_heads * kHeadDim;
    int S = (n_kv + kSplitKeys - 1) / kSplitKeys;
    S = max(1, min(S, min(kMaxSplits, (int) gridDim.x / op.n_kv_heads)));
    const int b = blockIdx.x;
    if (b >= op.n_kv_heads * S) return;
    const int g = b / S, s = b % S;
    const int range = (n_kv + S - 1) / S;
    const int j0 = s * range, j1 = min(j0 + range, n_kv);
    attn_prepare(op, qkv, g * REP, REP, g, pos, sm, qkv_ll, tag_in);
    if (s == S - 1) attn_append(op, g, pos, kvw, sm);
    const int cnt = max(j1 - j0, 0);
    attn_scores<REP>(op, sm, g, kvw, pos, j0, j1, sc, range);
    __syncthreads();
    float m[REP], l[REP];
#pragma unroll
    for (int r = 0; r < REP; ++r) {
        float mx = -CUDART_INF_F;
        for (int j = tid; j < cnt; j += kThreads) mx = fmaxf(mx, sc[r * range + j]);
        mx = block_max_f(mx, sm->redf);
        double sum = 0.0;
        for (int j = tid; j < cnt; j += kThreads) {
            const float e = __expf(sc[r * range + j] - mx);
            sc[r * range + j] = e;
            sum += (double) e;
        }
        sum = block_sum_d(sum, sm->red);
        m[r] = mx; l[r] = (float) sum;
    }
    __syncthreads();
    {   // P.V : 4 key partitions x 128 dims
        const int d = tid & (kHeadDim - 1), part = tid >> 7;
        float acc[REP];
#pragma unroll
        for (int r = 0; r < REP; ++r) acc[r] = 0.f;
        for (int j = part; j < cnt; j += 4) {
            const int jj = j0 + j;
            const float v = __half2float((jj < pos) ? __ldcg(op.vcache + (size_t) jj * kvw + g * kHeadDim + d) : sm->v16[d]);
#pragma unroll
            for (int r = 0; r < REP; ++r) acc[r] = fmaf(v, sc[r * range + j], acc[r]);
        }
#pragma unroll
        for (int r = 0; r < REP; ++r) sm->part[part][r][d] = acc[r];
    }
    __syncthreads();
    for (int idx = tid; idx < REP * kHeadDim; idx += kThreads) {
        const int r = idx / kHeadDim, d = idx % kHeadDim;
        const float o = (sm->part[0][r][d] + sm->part[1][r][d]) + (sm->part[2][r][d] + sm->part[3][r][d]);
        if (S == 1) {
            out[(size_t) (g * REP + r) * kHeadDim + d] = o / l[r];
        } else {
            float * p = op.attn_part + ((size_t) (g * REP + r) * kMaxSplits + s) * kPartStride;
            p[4 + d] = o;
            if (d == 0) { p[0] = m[r]; p[1] = l[r]; }
        }
    }
    if (S == 1) return;
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        const unsigned int old = atomicAdd(op.attn_cnt + g, 1u);
        sm->flag = (old == (unsigned int) (S - 1));
    }
    __syncthreads();
    if (!sm->flag) return;
    __threadfence();
    if (tid == 0) op.attn_cnt[g] = 0;
    for (int idx = tid; idx < REP * kHeadDim; idx += kThreads) {
        const int r = idx / kHeadDim, d = idx % kHeadDim;
        const float * base = op.attn_part + (size_t) (g * REP + r) * kMaxSplits * kPartStride;
        float M = -CUDART_INF_F;
        for (int i = 0; i < S; ++i) M = fmaxf(M, __ldcg(base + (size_t) i * kPartStride));
        float L = 0.f, o = 0.f;
        for (int i = 0; i < S; ++i) {
            const float w = __expf(__ldcg(base + (size_t) i * kPartStride) - M);
            L = fmaf(__ldcg(base + (size_t) i * kPartStride + 1), w, L);
            o = fmaf(__ldcg(base + (size_t) i * kPartStride + 4 + d), w, o);
        }
        out[(size_t) (g * REP + r) * kHeadDim + d] = o / L;
    }
}

// ---- ATTN_F32: code-predictor positions 0 and 1 in one stage (n_tok = 2), everything F32 with the
// un-cached K/V, as the reference's 2-token prefill graph (src/tts_transformer.cpp:1613-1626)
__device__ __noinline__ void attn_f32_pair(const Op & op, AttnSmem * sm, float * sc, const unsigned long long * qkv_ll,
                                           uint32_t tag_in, unsigned long long * out_ll_base, uint32_t tag) {
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (b >= 2 * op.n_heads) return;
    const int t = b / op.n_heads, h = b % op.n_heads;
    const int rep = op.n_heads / op.n_kv_heads, g = h / rep;
    const int kvw = op.n_kv_heads * kHeadDim;
    const int pos = op.pos_off + t;
    float * out = op.out + (size_t) t * op.out_stride + (size_t) h * kHeadDim;
    unsigned long long * out_ll = out_ll_base ? out_ll_base + (size_t) t * op.out_stride + (size_t) h * kHeadDim : nullptr;
    if (t == 1) {   // position 0's key / value, recomputed from token 0's projections
        attn_prepare(op, op.qkv, h, 0, g, op.pos_off, sm, qkv_ll, tag_in);
        if (tid < kHeadDim) { sm->k0f[tid] = sm->kf[tid]; sm->v0f[tid] = sm->vf[tid]; }
    }
    attn_prepare(op, op.qkv + (size_t) t * op.qkv_stride, h, 1, g, pos, sm, qkv_ll ? qkv_ll + (size_t) t * op.qkv_stride : nullptr, tag_in);
    if (h % rep == 0) attn_append(op, g, pos, kvw, sm);
    if (t == 0) {   // softmax over one key = 1.0 exactly
        if (tid < kHeadDim) {
            out[tid] = sm->vf[tid];
            if (out_ll) ll_store(out_ll + tid, sm->vf[tid], tag);
        }
        return;
    }
    if (warp < 2) {
        float s = 0.f;
        for (int d = lane; d < kHeadDim; d += 32) {
            const float kd = (warp == 0) ? sm->k0f[d] : sm->kf[d];
            s = fmaf(kd, sm->q[0][d], s);
        }
        s = warp_sum(s);
        if (lane == 0) sc[warp] = s * op.kq_scale;
    }
    __syncthreads();
    if (tid < kHeadDim) {
        const float mx = fmaxf(sc[0], sc[1]);
        const float e0 = expf(sc[0] - mx), e1 = expf(sc[1] - mx);
        const float inv = (float) (1.0 / ((double) e0 + (double) e1));
        const float p0 = e0 * inv, p1 = e1 * inv;
        const float o = fmaf(sm->vf[tid], p1, sm->v0f[tid] * p0);
        out[tid] = o;
        if (out_ll) ll_store(out_ll + tid, o, tag);
    }
}

// ---- ATTN_SOFTMAX_P16 / ATTN_FLASH_F16SEQ: one CTA per query head, the reference's exact
// rounding points (parity modes; the sequential F16 accumulator is inherently serial per dim)
__device__ __noinline__ void attn_reference(const Op & op, const LocalState * ls, AttnSmem * sm, float * sc,
                                            const unsigned long long * qkv_ll, uint32_t tag_in,
                                            unsigned long long * out_ll_base, uint32_t tag) {
    const int h = blockIdx.x;
    if (h >= op.n_heads) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = h / (op.n_heads / op.n_kv_heads);
    const int kvw = op.n_kv_heads * kHeadDim;
    const int pos = (op.use_n_past ? ls->n_past : 0) + op.pos_off;
    const int n_kv = pos + 1;
    float * ms = sc + kMaxScores;
    float * vs = ms + kTile;
    float * out = op.out + (size_t) h * kHeadDim;
    unsigned long long * out_ll = out_ll_base ? out_ll_base + (size_t) h * kHeadDim : nullptr;

    attn_prepare(op, op.qkv, h, 1, g, pos, sm, qkv_ll, tag_in);
    if (h % (op.n_heads / op.n_kv_heads) == 0) attn_append(op, g, pos, kvw, sm);
    attn_scores<1>(op, sm, g, kvw, pos, 0, n_kv, sc, 0);
    __syncthreads();

    if (op.mode == ATTN_SOFTMAX_P16) {
        float mx = -CUDART_INF_F;
        for (int j = tid; j < n_kv; j += kThreads) mx = fmaxf(mx, sc[j]);
        mx = block_max_f(mx, sm->redf);
        double sum = 0.0;
        for (int j = tid; j < n_kv; j += kThreads) {
            const float e = expf(sc[j] - mx);
            sc[j] = e;
            sum += (double) e;
        }
        sum = block_sum_d(sum, sm->red);
        const float inv = (float) (1.0 / sum);
        // soft_max scales by 1/sum (ops.cpp:5322-5323); the following mul_mat rounds P to F16
        for (int j = tid; j < n_kv; j += kThreads) sc[j] = f16r(sc[j] * inv);
        __syncthreads();
        const int d = tid & (kHeadDim - 1), part = tid >> 7;
        float acc = 0.f;
        for (int j = part; j < n_kv; j += 4) {
            const __half hv = (j < pos) ? __ldcg(op.vcache + (size_t) j * kvw + g * kHeadDim + d) : sm->v16[d];
            acc = fmaf(__half2float(hv), sc[j], acc);
        }
        sm->part[part][0][d] = acc;
        __syncthreads();
        if (tid < kHeadDim) {
            const float o = (sm->part[0][0][tid] + sm->part[1][0][tid]) + (sm->part[2][0][tid] + sm->part[3][0][tid]);
            out[tid] = o;
            if (out_ll) ll_store(out_ll + tid, o, tag);
        }
        return;
    }

    // ggml_compute_forward_flash_attn_ext_f16_one_chunk; split-KV only when n_kv >= 512
    // (ops.cpp:8740): chunk = ceil(n_kv / n_threads)
    int n_chunks = 1, chunk = n_kv;
    if (n_kv >= 512 && op.ref_threads > 1) {
        n_chunks = op.ref_threads;
        chunk = (n_kv + n_chunks - 1) / n_chunks;
    }
    float outv = 0.f;   // valid for tid < 128
    float Mf = -CUDART_INF_F, Sf = 0.f;
    for (int c = 0; c < n_chunks; ++c) {
        const int c0 = c * chunk, c1 = min(c0 + chunk, n_kv);
        if (c0 >= n_kv) break;
        float run = -CUDART_INF_F;      // warp 0: running max of the chunk's earlier scores
        float acc = 0.f, S = 0.f;
        for (int t0 = c0; t0 < c1; t0 += kTile) {
            const int t1 = min(t0 + kTile, c1);
            if (warp == 0) {
                for (int j0 = t0; j0 < t1; j0 += 32) {
                    const int j = j0 + lane;
                    const float s = (j < t1) ? sc[j] : -CUDART_INF_F;
                    float incl = s;             // inclusive prefix max within the group
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const float tt = __shfl_up_sync(0xffffffffu, incl, o);
                        if (lane >= o) incl = fmaxf(incl, tt);
                    }
                    float excl = __shfl_up_sync(0xffffffffu, incl, 1);
                    if (lane == 0) excl = -CUDART_INF_F;
                    if (j < t1) ms[j - t0] = fmaxf(run, excl);   // M_{j-1}
                    run = fmaxf(run, __shfl_sync(0xffffffffu, incl, 31));
                }
            }
            __syncthreads();
            for (int j = t0 + tid; j < t1; j += kThreads) {
                const float mprev = ms[j - t0], s = sc[j];
                if (s > mprev) { ms[j - t0] = expf_cr(mprev - s); vs[j - t0] = 1.0f; }
                else           { ms[j - t0] = 1.0f;               vs[j - t0] = expf_cr(s - mprev); }
            }
            __syncthreads();
            if (tid < kHeadDim) {
                const __half * vbase = op.vcache + g * kHeadDim + tid;
                for (int j0 = t0; j0 < t1; j0 += 8) {
                    float vv[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const int j = j0 + u;
                        vv[u] = (j < t1) ? __half2float((j < pos) ? __ldcg(vbase + (size_t) j * kvw) : sm->v16[tid]) : 0.f;
                    }
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const int j = j0 + u;
                        if (j < t1) {
                            const float m = ms[j - t0];
                            if (m != 1.0f) acc = f16r(acc * m);               // ggml_vec_scale_f16
                            acc = f16r(fmaf(vv[u], vs[j - t0], acc));         // ggml_vec_mad_f16
                        }
                    }
                }
            } else if (tid == kHeadDim) {
                for (int j = t0; j < t1; ++j) S = fmaf(S, ms[j - t0], vs[j - t0]);   // S = S*ms + vs
            }
            __syncthreads();
        }
        if (warp == 0 && lane == 0) sm->chunkM[c] = run;
        if (tid == kHeadDim) sm->chunkS[c] = S;
        __syncthreads();
        if (n_chunks == 1) {
            const float Sc = sm->chunkS[0];
            outv = acc * ((Sc == 0.0f) ? 0.0f : 1.0f / Sc);
        } else {
            // ggml_flash_attn_ext_reduce_partials (ops.cpp:8652-8688)
            const float Mc = sm->chunkM[c], Sc = sm->chunkS[c];
            if (Sc != 0.0f) {
                const float Mn = fmaxf(Mf, Mc);
                const float so = expf_cr(Mf - Mn), sn = expf_cr(Mc - Mn);
                outv = outv * so + acc * sn;
                Sf = Sf * so + Sc * sn;
                Mf = Mn;
            }
        }
    }
    if (n_chunks > 1 && Sf != 0.0f) outv = outv * (1.0f / Sf);
    if (tid < kHeadDim) {
        out[tid] = outv;
        if (out_ll) ll_store(out_ll + tid, outv, tag);
    }
}

// ---- ATTN_SOFTMAX_P16 with at most 16 keys (the code predictor's single-token passes, 70 stages per frame, and
// the first prefill tokens): ONE WARP per query head, no CTA barrier at all.  The CTA-wide path above spends ~12
// __syncthreads and three block reductions on 16 keys (measured 4.85 us per stage); here lane l holds dims
// l + 32 j (j < 4) of q / k / v - the NEOX pair (i, i + 64) sits in one lane - and for the key loop half-warp `hw`
// takes the keys of its parity, lane l16 of it 8 consecutive dims (one 16-byte load per cached key).
// Same arithmetic and rounding points as attn_reference / the reference's step graph (src/tts_transformer.cpp:
// 1666-1827): F16 Q, F16 cache, F32 scores, softmax with a double sum, P rounded to F16, P.V partial sums per key
// class j mod 4 combined as (p0 + p1) + (p2 + p3).
struct AttnWarpSmem {
    __half kc[16][kHeadDim], vc[16][kHeadDim];     // the KV group's cache rows [0, pos] (cp.async) + the current token's
    float q[kHeadDim];
    float wq[kHeadDim], wk[kHeadDim];
    float2 cs[64];
    float sc[16];
};

__device__ __noinline__ void attn_p16_warp(const Op & op, int pos, float * scratch, const unsigned long long * qkv_ll,
                                           uint32_t tag_in, unsigned long long * out_ll_base, uint32_t tag) {
    const int h = blockIdx.x;
    if (h >= op.n_heads || threadIdx.x >= 32) return;
    const int lane = threadIdx.x, hw = lane >> 4, l16 = lane & 15;
    const int rep = op.n_heads / op.n_kv_heads, g = h / rep;
    const int kvw = op.n_kv_heads * kHeadDim, n_kv = pos + 1;
    AttnWarpSmem * sm = reinterpret_cast<AttnWarpSmem *>(scratch);
    // Everything that does not depend on this token is requested BEFORE the projections are polled and waits in
    // shared memory, not in registers (the kernel is capped at 128 registers per thread, and the IEEE sqrt /
    // divisions below are calls that would spill whatever is live across them): the cached key / value rows of the
    // earlier positions by cp.async (16 bytes per lane and row), the norm weights and the RoPE factors.
    {
        const __half * kbase = op.kcache + g * kHeadDim + 8 * l16;
        const __half * vbase = op.vcache + g * kHeadDim + 8 * l16;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int jk = 2 * i + hw;
            if (jk < pos) {
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(smem_u32(&sm->kc[jk][8 * l16])), "l"(kbase + (size_t) jk * kvw) : "memory");
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(smem_u32(&sm->vc[jk][8 * l16])), "l"(vbase + (size_t) jk * kvw) : "memory");
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        sm->wq[lane + 32 * j] = op.q_norm_w ? __ldg(op.q_norm_w + lane + 32 * j) : 1.f;
        sm->wk[lane + 32 * j] = op.k_norm_w ? __ldg(op.k_norm_w + lane + 32 * j) : 1.f;
    }
#pragma unroll
    for (int j = 0; j < 2; ++j)
        sm->cs[lane + 32 * j] = __ldg(reinterpret_cast<const float2 *>(op.rope) + (size_t) pos * 64 + lane + 32 * j);
    float q[4], k[4], v[4];
    {
        const size_t oq = (size_t) h * kHeadDim, ok = (size_t) (op.n_heads + g) * kHeadDim,
                     ov = (size_t) (op.n_heads + op.n_kv_heads + g) * kHeadDim;
        if (qkv_ll) {
            unsigned long long w[12];
            uint32_t spins = 0;
            bool good;
            do {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    w[j] = ll_load(qkv_ll + oq + lane + 32 * j);
                    w[4 + j] = ll_load(qkv_ll + ok + lane + 32 * j);
                    w[8 + j] = ll_load(qkv_ll + ov + lane + 32 * j);
                }
                good = true;
#pragma unroll
                for (int j = 0; j < 12; ++j) good = good && ((uint32_t) (w[j] >> 32) == tag_in);
                if (!good) ll_spin_guard(spins);
            } while (!good);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                q[j] = __uint_as_float((uint32_t) w[j]); k[j] = __uint_as_float((uint32_t) w[4 + j]); v[j] = __uint_as_float((uint32_t) w[8 + j]);
            }
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                q[j] = __ldcg(op.qkv + oq + lane + 32 * j); k[j] = __ldcg(op.qkv + ok + lane + 32 * j); v[j] = __ldcg(op.qkv + ov + lane + 32 * j);
            }
        }
    }
    // per-head RMSNorm (double sum of float squares, ops.cpp:3738-3747) + NEOX RoPE
    auto norm_rope = [&](float * x, const float * w, bool do_norm) {
        if (do_norm) {
            double ss = 0.0;
#pragma unroll
            for (int j = 0; j < 4; ++j) ss += (double) (x[j] * x[j]);
            ss = warp_sum_d(ss);
            const float mean = (float) (ss * (1.0 / (double) kHeadDim));      // exact: a power of two
            const float scale = 1.0f / sqrtf(mean + op.eps);
#pragma unroll
            for (int j = 0; j < 4; ++j) x[j] = (x[j] * scale) * w[lane + 32 * j];
        }
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const float2 cs = sm->cs[lane + 32 * j];
            const float x0 = x[j], x1 = x[j + 2];
            x[j] = x0 * cs.x - x1 * cs.y;
            x[j + 2] = x0 * cs.y + x1 * cs.x;
        }
    };
    norm_rope(k, sm->wk, op.k_norm_w != nullptr);
    norm_rope(q, sm->wq, op.q_norm_w != nullptr);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const __half kh = __float2half_rn(k[j]), vh = __float2half_rn(v[j]);
        sm->kc[pos][lane + 32 * j] = kh; sm->vc[pos][lane + 32 * j] = vh;
        sm->q[lane + 32 * j] = f16r(q[j]);
        if (h % rep == 0) {     // one head of the group appends to the cache; nobody reads row `pos` from it in this stage
            op.kcache[(size_t) pos * kvw + g * kHeadDim + lane + 32 * j] = kh;
            op.vcache[(size_t) pos * kvw + g * kHeadDim + lane + 32 * j] = vh;
        }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncwarp();
    float qa[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) qa[i] = sm->q[8 * l16 + i];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int jk = 2 * i + hw;
        const uint4 kr = (jk < n_kv) ? *reinterpret_cast<const uint4 *>(&sm->kc[jk][8 * l16]) : make_uint4(0u, 0u, 0u, 0u);
        const __half2 * hk = reinterpret_cast<const __half2 *>(&kr);
        float sdot = 0.f;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const float2 f = __half22float2(hk[u]);
            sdot = fmaf(f.x, qa[2 * u], sdot);
            sdot = fmaf(f.y, qa[2 * u + 1], sdot);
        }
#pragma unroll
        for (int o2 = 8; o2 > 0; o2 >>= 1) sdot += __shfl_xor_sync(0xffffffffu, sdot, o2);
        if (l16 == 0 && jk < n_kv) sm->sc[jk] = sdot * op.kq_scale;
    }
    __syncwarp();
    float mx = -CUDART_INF_F;
    for (int jk = 0; jk < n_kv; ++jk) mx = fmaxf(mx, sm->sc[jk]);
    double sum = 0.0;
    for (int jk = 0; jk < n_kv; ++jk) sum += (double) expf(sm->sc[jk] - mx);
    const float inv = (float) (1.0 / sum);
    float accA[8], accB[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) { accA[u] = 0.f; accB[u] = 0.f; }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int jk = 2 * i + hw;
        if (jk < n_kv) {
            const float pk = f16r(expf(sm->sc[jk] - mx) * inv);        // soft_max scales by 1/sum; the P.V mul_mat rounds P to F16
            const uint4 vr = *reinterpret_cast<const uint4 *>(&sm->vc[jk][8 * l16]);
            const __half2 * hv = reinterpret_cast<const __half2 *>(&vr);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const float2 f = __half22float2(hv[u]);
                if (i & 1) { accB[2 * u] = fmaf(f.x, pk, accB[2 * u]); accB[2 * u + 1] = fmaf(f.y, pk, accB[2 * u + 1]); }
                else       { accA[2 * u] = fmaf(f.x, pk, accA[2 * u]); accA[2 * u + 1] = fmaf(f.y, pk, accA[2 * u + 1]); }
            }
        }
    }
    float * out = op.out + (size_t) h * kHeadDim + 8 * l16;
    unsigned long long * out_ll = out_ll_base ? out_ll_base + (size_t) h * kHeadDim + 8 * l16 : nullptr;
#pragma unroll
    for (int u = 0; u < 8; ++u) {
        // half-warp 0 holds key classes 0 (A) and 2 (B), half-warp 1 classes 1 and 3: (p0 + p1) + (p2 + p3)
        const float a = accA[u] + __shfl_xor_sync(0xffffffffu, accA[u], 16);
        const float b = accB[u] + __shfl_xor_sync(0xffffffffu, accB[u], 16);
        if (hw == 0) {
            const float o = a + b;
            out[u] = o;
            if (out_ll) ll_store(out_ll + u, o, tag);
        }
    }
}

__device__ __noinline__ void stage_attn(const Op & op, const LocalState * ls, float * scratch, uint32_t tag, bool ll) {
    AttnSmem * sm = reinterpret_cast<AttnSmem *>(scratch);
    float * sc = scratch + (sizeof(AttnSmem) + 15) / 16 * 4;
    // flag-carrying input (the QKV stage's out_ll) / output (polled by the O projection)
    const unsigned long long * qkv_ll = ll ? op.in_ll : nullptr;
    unsigned long long * out_ll = ll ? op.out_ll : nullptr;
    const uint32_t tag_in = tag - (uint32_t) op.in_dist;
    if (op.mode == ATTN_F32) { attn_f32_pair(op, sm, sc, qkv_ll, tag_in, out_ll, tag); return; }
    if (op.mode == ATTN_SOFTMAX_P16 && QTTS_OPT_ATTN_WARP) {
        const int p16_pos = (op.use_n_past ? ls->n_past : 0) + op.pos_off;
        if (p16_pos < 16) { attn_p16_warp(op, p16_pos, scratch, qkv_ll, tag_in, out_ll, tag); return; }
    }
    if (op.mode != ATTN_FAST) { attn_reference(op, ls, sm, sc, qkv_ll, tag_in, out_ll, tag); return; }
    const int pos = (op.use_n_past ? ls->n_past : 0) + op.pos_off;
    const int rep = op.n_heads / op.n_kv_heads;
    if (rep == 1)      attn_fast<1>(op, op.qkv, op.out, pos, sm, sc, qkv_ll, tag_in);
    else if (rep == 2) attn_fast<2>(op, op.qkv, op.out, pos, sm, sc, qkv_ll, tag_in);
    else               attn_fast<4>(op, op.qkv, op.out, pos, sm, sc, qkv_ll, tag_in);
}

// ------------------------------------------------------------------------------------------
// sampling (executed by every CTA; CTA 0 mirrors results to global memory)
// ------------------------------------------------------------------------------------------
struct SampleSmem {
    float l[kMaxVocab];
    unsigned short idx[kMaxVocab];
    unsigned int hist[256];
    float redf[kWarps];
    int   redi[kWarps];
    double red[kWarps];
    unsigned int sel_prefix, sel_k;
    int   count;
    int   token;
    int   wcount[kWarps];
};

__device__ __forceinline__ unsigned int ord_key(float f) {
    const unsigned int u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// first index of the maximum (argmax, src/tts_transformer.cpp:2051-2061)
__device__ int block_argmax(const float * l, int n, SampleSmem * sm) {
    float bv = -CUDART_INF_F; int bi = 0x7fffffff;
    for (int i = threadIdx.x; i < n; i += kThreads) {
        const float v = l[i];
        if (v > bv || (v == bv && i < bi)) { bv = v; bi = i; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
    }
    __syncthreads();
    if ((threadIdx.x & 31) == 0) { sm->redf[threadIdx.x >> 5] = bv; sm->redi[threadIdx.x >> 5] = bi; }
    __syncthreads();
    bv = sm->redf[0]; bi = sm->redi[0];
    for (int w = 1; w < kWarps; ++w) {
        const float ov = sm->redf[w]; const int oi = sm->redi[w];
        if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
    }
    if (bi == 0x7fffffff) bi = 0;   // all -inf / NaN: reference returns index 0
    return bi;
}

// k-th largest value of l[0..n) (radix select on order-preserving keys), 1 <= k <= n
__device__ __noinline__ float block_kth_largest(const float * l, int n, int k, SampleSmem * sm) {
    unsigned int prefix = 0, mask = 0;
    int kk = k;
    for (int shift = 24; shift >= 0; shift -= 8) {
        for (int i = threadIdx.x; i < 256; i += kThreads) sm->hist[i] = 0;
        __syncthreads();
        for (int i = threadIdx.x; i < n; i += kThreads) {
            const unsigned int key = ord_key(l[i]);
            if ((key & mask) == prefix) atomicAdd(&sm->hist[(key >> shift) & 255u], 1u);
        }
        __syncthreads();
        if (threadIdx.x < 32) {
            // lane t owns bins 255-8t .. 248-8t (descending)
            const int lane = threadIdx.x;
            unsigned int c[8], tot = 0;
#pragma unroll
            for (int b = 0; b < 8; ++b) { c[b] = sm->hist[255 - 8 * lane - b]; tot += c[b]; }
            unsigned int incl = tot;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            const unsigned int excl = incl - tot;
            if (excl < (unsigned) kk && (unsigned) kk <= incl) {
                unsigned int run = excl;
#pragma unroll
                for (int b = 0; b < 8; ++b) {
                    if (run < (unsigned) kk && (unsigned) kk <= run + c[b]) {
                        sm->sel_prefix = (unsigned) (255 - 8 * lane - b);
                        sm->sel_k = (unsigned) kk - run;
                    }
                    run += c[b];
                }
            }
        }
        __syncthreads();
        prefix |= sm->sel_prefix << shift;
        mask |= 255u << shift;
        kk = (int) sm->sel_k;
        __syncthreads();
    }
    const unsigned int u = (prefix & 0x80000000u) ? (prefix & 0x7fffffffu) : ~prefix;   // invert ord_key
    return __uint_as_float(u);
}

// Philox4x32-10 -> one uniform in [0,1)
__device__ float philox_uniform(uint64_t seed, uint32_t c0, uint32_t c1, uint32_t c2) {
    uint32_t k0 = (uint32_t) seed, k1 = (uint32_t) (seed >> 32);
    uint32_t x0 = c0, x1 = c1, x2 = c2, x3 = 0x51757433u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, x0), lo0 = 0xD2511F53u * x0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, x2), lo1 = 0xCD9E8D57u * x2;
        const uint32_t n0 = hi1 ^ x1 ^ k0, n1 = lo1, n2 = hi0 ^ x3 ^ k1, n3 = lo0;
        x0 = n0; x1 = n1; x2 = n2; x3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return (float) (x0 >> 8) * (1.0f / 16777216.0f);
}

// temperature / top-k / softmax / CDF scan on sm->l[0..n); p = scratch for n probabilities.
// eos_id >= 0: EOS survives top-k (src/tts_transformer.cpp:2451-2494); eos_id < 0: code-predictor
// rule (:2202-2232) with the CDF scan of the TRT sampler (src/trt_cuda_kernels.cu:160-180) in
// place of std::discrete_distribution.
__device__ __noinline__ int block_sample(int n, float temperature, int top_k, float u, int eos_id, SampleSmem * sm, float * p,
                                         int ref_nth) {
    float * l = sm->l;
    for (int i = threadIdx.x; i < n; i += kThreads) l[i] = l[i] / temperature;
    __syncthreads();
    const float eos_logit = (eos_id >= 0) ? l[eos_id] : 0.f;
    if (top_k > 0 && top_k < n) {
        float thr;
        if (ref_nth) {
            // the reference's codebook-0 threshold: scored_vals[scored_idx[top_k - 1]] after std::nth_element
            // (src/tts_transformer.cpp:2464-2466) - SOME member of the top-k set, whichever libstdc++'s introselect
            // leaves there; reproduced by running the same algorithm (qtts_nth.cuh) on one thread
            if (threadIdx.x == 0) {
                unsigned short * ix = reinterpret_cast<unsigned short *>(p);       // p is rewritten below
                for (int i = 0; i < n; ++i) ix[i] = (unsigned short) i;
                nth::nth_element_desc(ix, top_k, n, l);
                sm->redf[0] = l[ix[top_k - 1]];
            }
            __syncthreads();
            thr = sm->redf[0];
            __syncthreads();
        } else {
            thr = block_kth_largest(l, n, top_k, sm);
        }
        for (int i = threadIdx.x; i < n; i += kThreads) if (l[i] < thr) l[i] = -CUDART_INF_F;
        __syncthreads();
    }
    if (eos_id >= 0 && threadIdx.x == 0) l[eos_id] = eos_logit;
    __syncthreads();
    float mx = -CUDART_INF_F;
    for (int i = threadIdx.x; i < n; i += kThreads) mx = fmaxf(mx, l[i]);
    mx = block_max_f(mx, sm->redf);
    double sum = 0.0;
    for (int i = threadIdx.x; i < n; i += kThreads) {
        const float e = expf(l[i] - mx);
        p[i] = e;
        sum += (double) e;
    }
    sum = block_sum_d(sum, sm->red);
    // ordered compaction of the non-zero probabilities (zeros do not move the running sum)
    const int per = (n + kThreads - 1) / kThreads;
    const int b0 = threadIdx.x * per, b1 = min(b0 + per, n);
    int cnt = 0;
    for (int i = b0; i < b1; ++i) {
        const float pr = (float) ((double) p[i] / sum);
        p[i] = pr;
        cnt += (pr != 0.0f);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) sm->wcount[warp] = incl;
    __syncthreads();
    int base = 0;
    for (int w = 0; w < warp; ++w) base += sm->wcount[w];
    int pos = base + incl - cnt;
    // candidates in index order: id in idx[], probability in l[] (the logits are not needed any more), so the scan
    // below reads two dense arrays instead of chasing p[idx[t]]
    for (int i = b0; i < b1; ++i) if (p[i] != 0.0f) { sm->idx[pos] = (unsigned short) i; l[pos] = p[i]; ++pos; }
    if (threadIdx.x == kThreads - 1) sm->count = base + incl;
    __syncthreads();
    if (threadIdx.x == 0) {
        // src/tts_transformer.cpp:2487-2494: sequential float running sum, first index with cumsum >= u; the scan
        // starts at index 0 with cumsum = probs[0], so a draw of exactly 0 selects index 0 even when its probability
        // is 0 (zeros do not move the sum and are skipped).  No early exit: the adds form one dependent chain of
        // ~4 cycles each while the loads run ahead; `hit` keeps the first match.
        float cum = 0.f;
        int hit = -1;
        const int m = (u <= 0.0f) ? 0 : sm->count;
#pragma unroll 4
        for (int t = 0; t < m; ++t) {
            cum += l[t];
            if (hit < 0 && cum >= u) hit = t;
        }
        sm->token = (u <= 0.0f) ? 0 : (hit >= 0 ? (int) sm->idx[hit] : n - 1);
    }
    __syncthreads();
    return sm->token;
}

__device__ __noinline__ void stage_sample_cb0(const Op & op, StreamState * st, LocalState * ls, float * scratch, float * xs) {
    SampleSmem * sm = reinterpret_cast<SampleSmem *>(scratch);
    const bool lead = (blockIdx.x == 0);
    const int n = op.n_vocab;
    const int frame = ls->frame;
    const float rep = st->rep_pen;
    const float temperature = (op.temperature_override < -1e29f) ? st->temperature : op.temperature_override;
    const int top_k = (op.temperature_override < -1e29f) ? st->top_k : op.top_k_override;
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += kThreads) {
        float v = __ldcg(op.logits + i);
        if (i >= op.suppress_start && i != op.eos_id) v = -CUDART_INF_F;      // :2418-2422
        if (rep != 1.0f && ((ls->seen[i >> 5] >> (i & 31)) & 1u)) {           // :2425-2435
            v = (v > 0.0f) ? v / rep : v * rep;
        }
        sm->l[i] = v;
    }
    __syncthreads();
    const int expected = max(20, st->n_text_tokens * 4);                      // :2439-2445
    if (frame >= expected) {
        float mx = -CUDART_INF_F;
        for (int i = threadIdx.x; i < n; i += kThreads) mx = fmaxf(mx, sm->l[i]);
        mx = block_max_f(mx, sm->redf);
        if (threadIdx.x == 0) {
            const float ramp = fminf(1.0f, (float) (frame - expected) / (float) expected);
            const float target = mx + 5.0f;
            // one FMA, like the reference as compiled (GCC contracts `l += ramp * (target - l)` wherever the ISA has FMA)
            sm->l[op.eos_id] = __fmaf_rn(ramp, target - sm->l[op.eos_id], sm->l[op.eos_id]);
        }
        __syncthreads();
    }
    if (op.dump && lead) for (int i = threadIdx.x; i < n; i += kThreads) op.dump[i] = sm->l[i];
    const int ignore_eos = st->ignore_eos;
    if (ignore_eos) {   // benchmark switch (not reference behaviour): fixed-length utterances
        if (threadIdx.x == 0) sm->l[op.eos_id] = -CUDART_INF_F;
        __syncthreads();
    }
    int tok;
    if (temperature <= 0.0f) {
        tok = block_argmax(sm->l, n, sm);
    } else {
        const float * su = st->uniforms;
        const float u = op.uniforms ? __ldcg(op.uniforms + frame)
                      : (su && frame < st->uniform_frames) ? __ldcg(su + (size_t) frame * op.n_books)
                      : philox_uniform(st->seed, (uint32_t) frame, 0u, 0x0cb0u);
        tok = block_sample(n, temperature, top_k, u, ignore_eos ? -1 : op.eos_id, sm, xs, st->sample_mode);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        // teacher forcing (parity hook): the engine's own choice is still what the frame emits
        const int32_t * fc = st->forced;
        const bool forcing = fc != nullptr && frame < st->forced_frames;
        if (tok == op.eos_id && !forcing) {
            ls->done = 1; ls->eos = 1;
            if (lead) { st->done = 1; st->eos = 1; }
        } else {
            const int feed = forcing ? __ldcg(fc + (size_t) frame * op.n_books) : tok;
            ls->cur_codes[0] = tok;
            ls->feed_codes[0] = feed;
            ls->seen[feed >> 5] |= 1u << (feed & 31);
            if (lead) { st->cur_codes[0] = tok; st->feed_codes[0] = feed; st->seen[feed >> 5] = ls->seen[feed >> 5]; }
        }
    }
    __syncthreads();
}

__device__ __noinline__ void stage_sample_cp(const Op & op, StreamState * st, LocalState * ls, float * scratch, float * xs) {
    SampleSmem * sm = reinterpret_cast<SampleSmem *>(scratch);
    const bool lead = (blockIdx.x == 0);
    const int n = op.n_vocab;
    const float temperature = (op.temperature_override < -1e29f) ? st->temperature : op.temperature_override;
    const int top_k = (op.temperature_override < -1e29f) ? st->top_k : op.top_k_override;
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += kThreads) {
        const float v = __ldcg(op.logits + i);
        sm->l[i] = v;
        if (op.dump && lead) op.dump[i] = v;
    }
    __syncthreads();
    int tok;
    if (temperature <= 0.0f) {
        tok = block_argmax(sm->l, n, sm);
    } else {
        const float * su = (op.temperature_override < -1e29f) ? st->uniforms : nullptr;   // (not in qtts_predict_codes)
        const float u = op.uniforms ? __ldcg(op.uniforms + (op.book - 1))
                      : (su && ls->frame < st->uniform_frames) ? __ldcg(su + (size_t) ls->frame * op.n_books + op.book)
                      : philox_uniform(st->seed, (uint32_t) ls->frame, (uint32_t) op.book, 0x0c9du);
        tok = block_sample(n, temperature, top_k, u, -1, sm, xs, 0);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const int32_t * fc = (op.temperature_override < -1e29f) ? st->forced : nullptr;
        const int feed = op.forced ? __ldcg(op.forced + (op.book - 1))
                       : (fc && ls->frame < st->forced_frames) ? __ldcg(fc + (size_t) ls->frame * op.n_books + op.book) : tok;
        ls->cur_codes[op.book] = tok;
        ls->feed_codes[op.book] = feed;
        if (lead) { st->cur_codes[op.book] = tok; st->feed_codes[op.book] = feed; }
    }
    __syncthreads();
}

__device__ __noinline__ void stage_frame_end(const Op & op, StreamState * st, LocalState * ls) {
    __syncthreads();
    if (threadIdx.x == 0) {
        const int f = ls->frame;
        if (blockIdx.x == 0) {
            for (int c = 0; c < op.n_books; ++c) op.codes_out[(size_t) f * op.n_books + c] = ls->cur_codes[c];
            st->frame = f + 1;
        }
        ls->frame = f + 1;
        if (f + 1 >= st->max_len) {             // src/tts_transformer.cpp:2525-2527
            ls->done = 1;
            if (blockIdx.x == 0) st->done = 1;
        }
    }
    __syncthreads();
}

__device__ __noinline__ void stage_copy(const Op & op) {
    const int i0 = blockIdx.x * kThreads + threadIdx.x;
    const __half * row = op.table ? op.table + (size_t) __ldcg(op.tok_ptr) * op.n : nullptr;
    for (int i = i0; i < op.n; i += gridDim.x * kThreads) {
        const float a = row ? __half2float(row[i]) : __ldcg(op.a + i);
        op.out[i] = op.b ? (a + __ldcg(op.b + i)) : a;
    }
}

// ------------------------------------------------------------------------------------------
// the program runner
// ------------------------------------------------------------------------------------------
// persistent = 1: cooperative launch, grid barrier after every all-to-all stage, loops over
//                 n_frames frames (stops early once the stream is done when check_done is set).
// persistent = 0: debug path, the host launches the same code one op at a time (n_ops = 1).
// Dynamic shared memory: [ring: n_slots x slot_bytes][xs: xs_elems floats][scratch].
constexpr int kScratchBytes = 40 * 1024;

extern "C" __global__ void __launch_bounds__(kThreads, 1)
qtts_run_program(const Op * __restrict__ ops, int n_ops, StreamState * st, GridBarrier * bar,
                 int n_frames, RunCfg cfg) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[kNumSlots];
    __shared__ double red[kWarps];
    __shared__ LocalState ls;
    __shared__ Op sop2[2];         // stage descriptors, double buffered: [n & 1] is stage n's
    static_assert(sizeof(Op) % 4 == 0 && sizeof(Op) / 4 <= kThreads, "one descriptor word per thread");
    uint32_t next_word = 0;

    Ring ring;
    ring.base = smem_raw;
    ring.full = full_bar;
    ring.slot_bytes = cfg.slot_bytes;
    ring.n_slots = cfg.n_slots;
    float * xs = reinterpret_cast<float *>(smem_raw + (size_t) cfg.n_slots * cfg.slot_bytes);
    float * scratch = xs + cfg.xs_elems;

    // CTA-local state from the stream's global state (the previous launch has completed)
    for (int i = threadIdx.x; i < (int) (sizeof(LocalState) / 4); i += kThreads) {
        int32_t * dst = reinterpret_cast<int32_t *>(&ls);
        const int nb = kMaxBooks, o_cur = 4, o_feed = 4 + nb, o_seen = 4 + 2 * nb;
        int32_t v;
        if (i == 0) v = st->n_past;
        else if (i == 1) v = st->frame;
        else if (i == 2) v = st->done;
        else if (i == 3) v = st->eos;
        else if (i < o_feed) v = st->cur_codes[i - o_cur];
        else if (i < o_seen) v = st->feed_codes[i - o_feed];
        else v = (int32_t) st->seen[i - o_seen];
        dst[i] = v;
    }
    if (threadIdx.x == 0) {
        for (int i = 0; i < kNumSlots; ++i) mbar_init(full_bar + i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    __shared__ Prefetch pf;
    uint32_t cons_seq = 0;          // ring chunks consumed so far (uniform across the CTA)
    __shared__ unsigned int epoch;  // thread 0 only; bar->count is zeroed by the host before every persistent launch
    if (threadIdx.x == 0) epoch = 0;
    if (threadIdx.x == kPfThread) {
        pf.idx = -1; pf.frame = (cfg.n_sched > 0) ? 0 : n_frames; pf.q = 0; pf.p1 = 0; pf.K = 0; pf.ppc = 1; pf.slot = 0;
        pf.W = nullptr; pf.seq = 0;
        pf.policy = policy_evict_first();
        if (cfg.n_sched > 0) {
            sched_fetch(pf, cfg.sched, cfg.n_sched, 0);
            sched_fetch(pf, cfg.sched, cfg.n_sched, 1);
        }
        prefetch_issue(cfg.sched, cfg.n_sched, n_frames, pf, cons_seq, ring);
    }
    int cur = 0;
    if (threadIdx.x < (int) (sizeof(Op) / 4)) {
        reinterpret_cast<uint32_t *>(&sop2[0])[threadIdx.x] = __ldg(reinterpret_cast<const uint32_t *>(ops) + threadIdx.x);
        next_word = __ldg(reinterpret_cast<const uint32_t *>(ops + (1 % n_ops)) + threadIdx.x);
    }
    bool stop = false;
    bool pf_pending = false;        // a stage without a grid barrier leaves the ring refill to the next stage's entry
    const bool ll = cfg.persistent != 0;
    for (int f = 0; f < n_frames && !stop; ++f) {
        if (cfg.check_done && ls.done) break;
        for (int i = 0; i < n_ops && !stop; ++i) {
            // stage descriptors live in shared memory (with ~225 KB of shared memory per CTA there is no
            // L1 left: every field read from the global descriptor was an L2 round trip).  Stage n+1's
            // descriptor is written to the other buffer now, from the register it was requested into
            // one stage ago, and stage n+2's is requested: one __syncthreads per stage boundary.
            __syncthreads();
            if (threadIdx.x < (int) (sizeof(Op) / 4)) {
                reinterpret_cast<uint32_t *>(&sop2[cur ^ 1])[threadIdx.x] = next_word;
                const int ni = (i + 2) % n_ops;
                next_word = __ldg(reinterpret_cast<const uint32_t *>(ops + ni) + threadIdx.x);
            }
            const Op & op = sop2[cur];
            cur ^= 1;
            // ring refill left over by a stage that ended without a grid barrier: a GEMV stage issues it
            // after its input is staged (the prefetcher's warp is the one with the fewest rows), any
            // other stage right here
            if (pf_pending && op.type != OP_GEMV && threadIdx.x == kPfThread)
                prefetch_issue(cfg.sched, cfg.n_sched, n_frames, pf, cons_seq, ring);
            const bool pf_in_stage = pf_pending && op.type == OP_GEMV;
            pf_pending = false;
            bool all_to_all = true;
            uint32_t * tr = nullptr;
            if (cfg.trace && blockIdx.x == cfg.prof_cta && threadIdx.x == 0 && f * n_ops + i < cfg.trace_cap) {
                tr = cfg.trace + (size_t) (f * n_ops + i) * 8;
                tr[0] = (uint32_t) op.type | ((uint32_t) i << 8);
                tr[6] = (uint32_t) f;
                tr[1] = (uint32_t) clock64();
            }
            const bool prof_on = cfg.prof && blockIdx.x == cfg.prof_cta && threadIdx.x == 0;
            const long long t0 = prof_on ? clock64() : 0;
            const long long t0p = (cfg.prof && threadIdx.x == 0) ? clock64() : 0;
            switch (op.type) {
                case OP_GEMV:       cons_seq += stage_gemv(op, st, &ls, xs, red, ring, cons_seq, cfg.sched, cfg.n_sched, n_frames, pf, prof_on ? cfg.prof : nullptr,
                                                           cfg.tag_base + 1u + (uint32_t) (f * n_ops + i), ll, pf_in_stage, tr); break;
                case OP_ATTN:       stage_attn(op, &ls, scratch, cfg.tag_base + 1u + (uint32_t) (f * n_ops + i), ll); break;
                case OP_COPY:       stage_copy(op); break;
                case OP_SAMPLE_CB0: stage_sample_cb0(op, st, &ls, scratch, xs); all_to_all = false; break;
                case OP_SAMPLE_CP:  stage_sample_cp(op, st, &ls, scratch, xs); all_to_all = false; break;
                case OP_FRAME_END:  stage_frame_end(op, st, &ls); all_to_all = false; break;
            }
            const long long t1 = prof_on ? clock64() : 0;
            const long long t1p = ((cfg.prof || tr) && threadIdx.x == 0) ? clock64() : 0;
            if (!all_to_all) {
                if (cfg.check_done && op.type != OP_SAMPLE_CP && ls.done) stop = true;
            } else if (ll && (op.flags & F_LL_OUT)) {
                pf_pending = true;      // consumers poll the flag-carrying output; no barrier
            } else if (cfg.persistent) {
                grid_barrier(bar, epoch, cfg.sched, cfg.n_sched, n_frames, pf, cons_seq, ring, prof_on ? cfg.prof + op.type * 8 + 7 : nullptr);
            } else {
                __syncthreads();
            }
            if (tr) { if (op.type != OP_GEMV) tr[4] = (uint32_t) t1p; tr[5] = (uint32_t) clock64(); }
            if (cfg.prof && threadIdx.x == 0 && op.type <= OP_ATTN) {   // per-CTA view: prof[64 + cta*4 + ...]
                const long long t2 = clock64();
                cfg.prof[64 + blockIdx.x * 4 + op.type * 2 + 0] += (unsigned long long) (t1p - t0p);
                cfg.prof[64 + blockIdx.x * 4 + op.type * 2 + 1] += (unsigned long long) (t2 - t1p);
            }
            if (prof_on) {      // [type][0] = count, [1] = stage cycles, [2] = barrier cycles (CTA 0's view)
                const long long t2 = clock64();
                cfg.prof[op.type * 8 + 0] += 1;
                cfg.prof[op.type * 8 + 1] += (unsigned long long) (t1 - t0);
                cfg.prof[op.type * 8 + 2] += (unsigned long long) (t2 - t1);
            }
        }
    }
    // drain: every bulk copy issued into this CTA's shared memory must land before it exits
    __syncthreads();
    if (threadIdx.x == kPfThread) {
        asm volatile("cp.async.wait_all;" ::: "memory");
        for (uint32_t s = cons_seq; s != pf.seq; ++s)
            mbar_wait(full_bar + (s % (uint32_t) cfg.n_slots), (s / (uint32_t) cfg.n_slots) & 1u);
    }
    __syncthreads();
}

// ring slot size: >= two rows of the widest projection input, multiple of 1024, at least 16 KiB
size_t slot_bytes_for(int k_max) {
    size_t b = (size_t) 4 * k_max;
    if (b < 16384) b = 16384;
    return (b + 1023) & ~(size_t) 1023;
}
size_t xs_elems_for(int k_max, int k2_max) {
    size_t e = (size_t) k_max;
    if ((size_t) 2 * k2_max > e) e = (size_t) 2 * k2_max;
    if (e < (size_t) kMaxVocab) e = kMaxVocab;
    return (e + 31) & ~(size_t) 31;
}
int n_slots_for(int k_max, int k2_max, size_t smem_limit) {
    const size_t fixed = xs_elems_for(k_max, k2_max) * sizeof(float) + kScratchBytes + 2048 /* static */;
    if (smem_limit < fixed + 2 * slot_bytes_for(k_max)) return 0;
    size_t n = (smem_limit - fixed) / slot_bytes_for(k_max);
    return (int) (n > (size_t) kNumSlots ? kNumSlots : n);
}
size_t smem_required(int k_max, int k2_max, int n_slots) {
    return (size_t) n_slots * slot_bytes_for(k_max) + xs_elems_for(k_max, k2_max) * sizeof(float) + kScratchBytes;
}
static_assert(sizeof(AttnSmem) + (kMaxScores + 2 * kTile) * sizeof(float) + 64 <= kScratchBytes, "attention scratch");
static_assert(sizeof(SampleSmem) <= kScratchBytes, "sampling scratch");

}  // namespace qtts
