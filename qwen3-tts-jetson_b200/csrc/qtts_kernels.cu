// qtts_kernels.cu — sm_100a device stages of the Qwen3-TTS codec-frame loop, batch-1 /
// small-batch regime (HBM-bound GEMV path).  See DESIGN.md for the stage list and rooflines.
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
#include <math_constants.h>

namespace qtts {

// ------------------------------------------------------------------------------------------
// small helpers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float f16r(float v) { return __half2float(__float2half_rn(v)); }

__device__ __forceinline__ uint4 ld_stream(const uint4 * p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ void prefetch_l2(const void * p) {
    asm volatile("prefetch.global.L2 [%0];" :: "l"(p));
}
__device__ __forceinline__ float bf_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }

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
// grid barrier (cooperative launch: all CTAs co-resident)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void grid_barrier(GridBarrier * bar, unsigned int & epoch) {
    __syncthreads();
    if (threadIdx.x == 0) {
        epoch += gridDim.x;
        __threadfence();
        atomicAdd(&bar->count, 1u);
        unsigned int v;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(&bar->count));
        } while ((int) (v - epoch) < 0);
        __threadfence();
    }
    __syncthreads();
}

// ------------------------------------------------------------------------------------------
// OP_GEMV
// ------------------------------------------------------------------------------------------
__device__ void gemv_load_input(const Op & op, const StreamState * st, float * xs) {
    const int K = op.K;
    if (op.in_kind == IN_VEC) {
        for (int i = threadIdx.x; i < K; i += kThreads) xs[i] = __ldcg(op.in + i);
    } else if (op.in_kind == IN_EMBED_ROW) {
        const int tok = __ldcg(op.tok_ptr);
        const __half * row = op.table + (size_t) tok * K;
        for (int i = threadIdx.x; i < K; i += kThreads) xs[i] = __half2float(row[i]);
    } else {
        // next Talker input (src/tts_transformer.cpp:2529-2553): sequential f32 sum, same order
        const EmbedSumArgs * es = op.es;
        const int frame = __ldcg(&st->frame) - 1;   // frame already counted by OP_FRAME_END
        const int tl = __ldcg(&st->trailing_len);
        const float * trow = (frame < tl) ? es->trailing + (size_t) frame * K : es->pad;
        for (int i = threadIdx.x; i < K; i += kThreads) {
            float acc = __half2float(es->tables[0][(size_t) __ldcg(&st->cur_codes[0]) * K + i]);
            for (int c = 1; c < es->n_books; ++c)
                acc += __half2float(es->tables[c][(size_t) __ldcg(&st->cur_codes[c]) * K + i]);
            acc += __ldcg(trow + i);
            xs[i] = acc;
        }
    }
}

__device__ __forceinline__ float dot8(const uint4 w, const float4 a, const float4 b) {
    float s = bf_lo(w.x) * a.x;
    s = fmaf(bf_hi(w.x), a.y, s);
    s = fmaf(bf_lo(w.y), a.z, s);
    s = fmaf(bf_hi(w.y), a.w, s);
    s = fmaf(bf_lo(w.z), b.x, s);
    s = fmaf(bf_hi(w.z), b.y, s);
    s = fmaf(bf_lo(w.w), b.z, s);
    s = fmaf(bf_hi(w.w), b.w, s);
    return s;
}

__device__ void stage_gemv(const Op & op, StreamState * st, float * smem, double * red) {
    const int K = op.K;
    float * xs = smem;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    gemv_load_input(op, st, xs);
    __syncthreads();
    if (op.x_out && blockIdx.x == 0)
        for (int i = threadIdx.x; i < K; i += kThreads) op.x_out[i] = xs[i];

    if (op.norm_w) {
        // ggml_compute_forward_rms_norm_f32: double sum of float squares (ops.cpp:3738-3747)
        double ss = 0.0;
        for (int i = threadIdx.x; i < K; i += kThreads) ss += (double) (xs[i] * xs[i]);
        ss = block_sum_d(ss, red);
        const float mean = (float) (ss / (double) K);
        const float scale = 1.0f / sqrtf(mean + op.eps);
        for (int i = threadIdx.x; i < K; i += kThreads) {
            const float v = (xs[i] * scale) * op.norm_w[i];
            if (op.hid_out && blockIdx.x == 0) op.hid_out[i] = v;
            xs[i] = f16r(v);
        }
    } else {
        for (int i = threadIdx.x; i < K; i += kThreads) xs[i] = f16r(xs[i]);
    }
    __syncthreads();

    // rows are handled as pairs (2p, 2p+1); EPI_SWIGLU needs exactly that pairing
    const int pairs = op.rows >> 1;
    const int p0 = (int) ((long long) pairs * blockIdx.x / gridDim.x);
    const int p1 = (int) ((long long) pairs * (blockIdx.x + 1) / gridDim.x);
    const int kv = K >> 3;   // uint4 per row
    const float4 * xs4 = reinterpret_cast<const float4 *>(xs);

    for (int p = p0 + warp; p < p1; p += kWarps) {
        const uint4 * w0 = reinterpret_cast<const uint4 *>(op.W + (size_t) (2 * p) * K);
        const uint4 * w1 = w0 + kv;
        float a0 = 0.f, a1 = 0.f;
        for (int j0 = 0; j0 < kv; j0 += 128) {
            uint4 A[4], B[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int j = j0 + u * 32 + lane;
                if (j < kv) { A[u] = ld_stream(w0 + j); B[u] = ld_stream(w1 + j); }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int j = j0 + u * 32 + lane;
                if (j < kv) {
                    const float4 xa = xs4[2 * j], xb = xs4[2 * j + 1];
                    a0 += dot8(A[u], xa, xb);
                    a1 += dot8(B[u], xa, xb);
                }
            }
        }
        a0 = warp_sum(a0);
        a1 = warp_sum(a1);
        if (lane == 0) {
            const int r0 = 2 * p, r1 = 2 * p + 1;
            switch (op.epi) {
                case EPI_STORE: op.out[r0] = a0; op.out[r1] = a1; break;
                case EPI_ADD:
                    op.out[r0] = a0 + __ldcg(op.res + r0);
                    op.out[r1] = a1 + __ldcg(op.res + r1);
                    break;
                case EPI_SWIGLU: op.out[p] = silu_f(a0) * a1; break;
                case EPI_BIAS:
                    op.out[r0] = a0 + op.bias[r0];
                    op.out[r1] = a1 + op.bias[r1];
                    break;
                case EPI_BIAS_SILU:
                    op.out[r0] = silu_f(a0 + op.bias[r0]);
                    op.out[r1] = silu_f(a1 + op.bias[r1]);
                    break;
            }
        }
    }
    if ((op.flags & F_INC_NPAST) && blockIdx.x == 0 && threadIdx.x == 0) st->n_past = st->n_past + 1;
}

// issue L2 prefetches for the weight rows this CTA will read in a following GEMV stage
__device__ void gemv_prefetch(const Op & op) {
    if (op.type != OP_GEMV) return;
    const int pairs = op.rows >> 1;
    const int p0 = (int) ((long long) pairs * blockIdx.x / gridDim.x);
    const int p1 = (int) ((long long) pairs * (blockIdx.x + 1) / gridDim.x);
    const char * base = reinterpret_cast<const char *>(op.W + (size_t) (2 * p0) * op.K);
    const size_t bytes = (size_t) (p1 - p0) * 2 * op.K * sizeof(__nv_bfloat16);
    for (size_t off = (size_t) threadIdx.x * 128; off < bytes; off += (size_t) kThreads * 128)
        prefetch_l2(base + off);
}

// ------------------------------------------------------------------------------------------
// OP_ATTN — one CTA per query head
// ------------------------------------------------------------------------------------------
struct AttnSmem {
    float q[kHeadDim];        // query after norm+rope (rounded to F16 unless ATTN_F32)
    float kf[kHeadDim];       // current key, unrounded
    float vf[kHeadDim];       // current value, unrounded
    __half k16[kHeadDim];
    __half v16[kHeadDim];
    float tmp[3 * kHeadDim];  // raw q,k,v staging
    double red[kWarps];
    float redf[kWarps];
    float part[4][kHeadDim];  // partial P.V sums
    float chunkM[64], chunkS[64];
};

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

__device__ void stage_attn(const Op & op, StreamState * st, float * smem_f) {
    const int h = blockIdx.x;
    if (h >= op.n_heads) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = h / (op.n_heads / op.n_kv_heads);
    const int kvw = op.n_kv_heads * kHeadDim;            // cache row width (elements)
    const int pos = (op.use_n_past ? __ldcg(&st->n_past) : 0) + op.pos_off;
    const int n_kv = pos + 1;
    const int mode = op.mode;

    AttnSmem * sm = reinterpret_cast<AttnSmem *>(smem_f);
    float * sc = smem_f + (sizeof(AttnSmem) + 3) / 4;    // scores [n_kv]
    float * ms = sc + ((n_kv + 3) & ~3);                 // per-key rescale factors
    float * vs = ms + ((n_kv + 3) & ~3);

    // ---- load raw q,k,v; per-head RMSNorm (double sum) ; NEOX RoPE -------------------------
    if (tid < 3 * kHeadDim) {
        const int which = tid / kHeadDim, d = tid % kHeadDim;
        const float * src = (which == 0) ? op.qkv + (size_t) h * kHeadDim
                          : (which == 1) ? op.qkv + (size_t) (op.n_heads + g) * kHeadDim
                                         : op.qkv + (size_t) (op.n_heads + op.n_kv_heads + g) * kHeadDim;
        sm->tmp[tid] = __ldcg(src + d);
    }
    __syncthreads();
    {
        // warps 0-3: q, warps 4-7: k  -> sums in red[0..7]
        double ss = 0.0;
        if (tid < 2 * kHeadDim) ss = (double) (sm->tmp[tid] * sm->tmp[tid]);
        ss = warp_sum_d(ss);
        if (lane == 0) sm->red[warp] = ss;
    }
    __syncthreads();
    if (tid < 2 * kHeadDim) {
        const int which = tid / kHeadDim, d = tid % kHeadDim;
        const double tot = sm->red[which * 4] + sm->red[which * 4 + 1] + sm->red[which * 4 + 2] + sm->red[which * 4 + 3];
        const float mean = (float) (tot / (double) kHeadDim);
        const float scale = 1.0f / sqrtf(mean + op.eps);
        // QK-norm only when the layer has the weight (src/tts_transformer.cpp:1424-1432)
        const float * w = which ? op.k_norm_w : op.q_norm_w;
        if (w) sm->tmp[tid] = (sm->tmp[tid] * scale) * w[d];
    }
    __syncthreads();
    if (tid < 2 * kHeadDim) {
        const int which = tid / kHeadDim, d = tid % kHeadDim;
        const int i = d & 63;
        const float c = op.rope[((size_t) pos * 64 + i) * 2], s = op.rope[((size_t) pos * 64 + i) * 2 + 1];
        const float x0 = sm->tmp[which * kHeadDim + i], x1 = sm->tmp[which * kHeadDim + i + 64];
        const float r = (d < 64) ? (x0 * c - x1 * s) : (x0 * s + x1 * c);
        if (which == 0) {
            sm->q[d] = (mode == ATTN_F32) ? r : f16r(r);
        } else {
            sm->kf[d] = r;
            sm->k16[d] = __float2half_rn(r);
        }
    } else if (tid < 3 * kHeadDim) {
        const int d = tid - 2 * kHeadDim;
        const float v = sm->tmp[tid];
        sm->vf[d] = v;
        sm->v16[d] = __float2half_rn(v);
    }
    __syncthreads();
    // append to the cache (first query head of each KV group writes)
    if (h % (op.n_heads / op.n_kv_heads) == 0 && tid < kHeadDim) {
        op.kcache[(size_t) pos * kvw + g * kHeadDim + tid] = sm->k16[tid];
        op.vcache[(size_t) pos * kvw + g * kHeadDim + tid] = sm->v16[tid];
        if (mode == ATTN_F32 && pos == 0) {
            op.kv0_f32[g * kHeadDim + tid] = sm->kf[tid];
            op.kv0_f32[kvw + g * kHeadDim + tid] = sm->vf[tid];
        }
    }
    float * out = const_cast<float *>(op.out) + (size_t) h * kHeadDim;

    // ---- ATTN_F32: code-predictor positions 0 and 1, all-F32 (src/tts_transformer.cpp:1613-1626)
    if (mode == ATTN_F32) {
        if (pos == 0) {   // softmax over one key = 1.0 exactly
            if (tid < kHeadDim) out[tid] = sm->vf[tid];
            return;
        }
        // pos == 1: keys {0 (from kv0_f32), 1 (own)}
        if (warp < 2) {
            float s = 0.f;
            for (int d = lane; d < kHeadDim; d += 32) {
                const float kd = (warp == 0) ? __ldcg(op.kv0_f32 + g * kHeadDim + d) : sm->kf[d];
                s = fmaf(kd, sm->q[d], s);
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
            const float v0 = __ldcg(op.kv0_f32 + kvw + g * kHeadDim + tid);
            out[tid] = fmaf(sm->vf[tid], p1, v0 * p0);
        }
        return;
    }

    // ---- scores s_j = scale * <K16_j, q16>  (8 lanes per key, 16 dims per lane) -------------
    {
        float qreg[16];
        const int sub = lane & 7, kin = lane >> 3;
#pragma unroll
        for (int i = 0; i < 16; ++i) qreg[i] = sm->q[sub * 16 + i];
        for (int j0 = warp * 4; j0 < n_kv; j0 += kWarps * 4) {
            const int j = j0 + kin;
            float s = 0.f;
            if (j < n_kv) {
                uint4 ka, kb;
                if (j < pos) {
                    const uint4 * k4 = reinterpret_cast<const uint4 *>(op.kcache + (size_t) j * kvw + g * kHeadDim + sub * 16);
                    ka = __ldcg(k4); kb = __ldcg(k4 + 1);
                } else {
                    const uint4 * k4 = reinterpret_cast<const uint4 *>(sm->k16 + sub * 16);
                    ka = k4[0]; kb = k4[1];
                }
                s = dot16h(ka, kb, qreg);
            }
            s += __shfl_xor_sync(0xffffffffu, s, 1);
            s += __shfl_xor_sync(0xffffffffu, s, 2);
            s += __shfl_xor_sync(0xffffffffu, s, 4);
            if (sub == 0 && j < n_kv) sc[j] = s * op.kq_scale;
        }
    }
    __syncthreads();

    if (mode == ATTN_SOFTMAX_P16 || mode == ATTN_FAST) {
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
        if (mode == ATTN_SOFTMAX_P16) {
            // soft_max scales by 1/sum (ops.cpp:5322-5323); the following mul_mat rounds P to F16
            for (int j = tid; j < n_kv; j += kThreads) sc[j] = f16r(sc[j] * inv);
        } else {
            for (int j = tid; j < n_kv; j += kThreads) sc[j] = sc[j] * inv;
        }
        __syncthreads();
        // P.V : 4 key partitions x 128 dims
        const int d = tid & (kHeadDim - 1), part = tid >> 7;
        float acc = 0.f;
        for (int j = part; j < n_kv; j += 4) {
            const __half hv = (j < pos) ? __ldcg(op.vcache + (size_t) j * kvw + g * kHeadDim + d) : sm->v16[d];
            acc = fmaf(__half2float(hv), sc[j], acc);
        }
        sm->part[part][d] = acc;
        __syncthreads();
        if (tid < kHeadDim)
            out[tid] = (sm->part[0][tid] + sm->part[1][tid]) + (sm->part[2][tid] + sm->part[3][tid]);
        return;
    }

    // ---- ATTN_FLASH_F16SEQ: ggml_compute_forward_flash_attn_ext_f16_one_chunk ---------------
    // split-KV only when n_kv >= 512 (ops.cpp:8740): chunk = ceil(n_kv / n_threads)
    int n_chunks = 1, chunk = n_kv;
    if (n_kv >= 512 && op.ref_threads > 1) {
        n_chunks = op.ref_threads;
        chunk = (n_kv + n_chunks - 1) / n_chunks;
    }
    // per-key (ms, vs): depends only on the running max of the chunk's earlier scores
    if (warp == 0) {
        for (int c = 0; c < n_chunks; ++c) {
            const int c0 = c * chunk, c1 = min(c0 + chunk, n_kv);
            float run = -CUDART_INF_F;      // max of scores before the current 32-group
            for (int j0 = c0; j0 < c1; j0 += 32) {
                const int j = j0 + lane;
                const float s = (j < c1) ? sc[j] : -CUDART_INF_F;
                float incl = s;             // inclusive prefix max within the group
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const float t = __shfl_up_sync(0xffffffffu, incl, o);
                    if (lane >= o) incl = fmaxf(incl, t);
                }
                float excl = __shfl_up_sync(0xffffffffu, incl, 1);
                if (lane == 0) excl = -CUDART_INF_F;
                const float mprev = fmaxf(run, excl);
                if (j < c1) ms[j] = mprev;  // stash M_{j-1}; turned into factors below
                run = fmaxf(run, __shfl_sync(0xffffffffu, incl, 31));
            }
            if (lane == 0) sm->chunkM[c] = run;
        }
    }
    __syncthreads();
    for (int j = tid; j < n_kv; j += kThreads) {
        const float mprev = ms[j], s = sc[j];
        if (s > mprev) { ms[j] = expf_cr(mprev - s); vs[j] = 1.0f; }
        else           { ms[j] = 1.0f;               vs[j] = expf_cr(s - mprev); }
    }
    __syncthreads();

    float outv = 0.f;   // valid for tid < 128
    float Mf = -CUDART_INF_F, Sf = 0.f;
    for (int c = 0; c < n_chunks; ++c) {
        const int c0 = c * chunk, c1 = min(c0 + chunk, n_kv);
        if (c0 >= n_kv) break;
        float acc = 0.f;
        if (tid < kHeadDim) {
            const __half * vbase = op.vcache + g * kHeadDim + tid;
            for (int j0 = c0; j0 < c1; j0 += 8) {
                float vv[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int j = j0 + u;
                    vv[u] = (j < c1) ? __half2float((j < pos) ? __ldcg(vbase + (size_t) j * kvw) : sm->v16[tid]) : 0.f;
                }
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int j = j0 + u;
                    if (j < c1) {
                        const float m = ms[j];
                        if (m != 1.0f) acc = f16r(acc * m);          // ggml_vec_scale_f16
                        acc = f16r(fmaf(vv[u], vs[j], acc));          // ggml_vec_mad_f16
                    }
                }
            }
        } else if (tid == kHeadDim) {
            float S = 0.f;
            for (int j = c0; j < c1; ++j) S = fmaf(S, ms[j], vs[j]);  // S = S*ms + vs
            sm->chunkS[c] = S;
        }
        __syncthreads();
        if (n_chunks == 1) {
            const float S = sm->chunkS[0];
            const float S_inv = (S == 0.0f) ? 0.0f : 1.0f / S;
            outv = acc * S_inv;
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
    if (tid < kHeadDim) out[tid] = outv;
}

// ------------------------------------------------------------------------------------------
// sampling
// ------------------------------------------------------------------------------------------
struct SampleSmem {
    float l[kMaxVocab];
    float p[kMaxVocab];
    int   idx[kMaxVocab];
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
__device__ float block_kth_largest(const float * l, int n, int k, SampleSmem * sm) {
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
    // invert ord_key
    const unsigned int u = (prefix & 0x80000000u) ? (prefix & 0x7fffffffu) : ~prefix;
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

// temperature / top-k / softmax / CDF scan on sm->l[0..n).  eos_id >= 0: EOS survives top-k
// (src/tts_transformer.cpp:2451-2494); eos_id < 0: code-predictor rule (:2202-2232) with the
// CDF scan of the TRT sampler (src/trt_cuda_kernels.cu:160-180) in place of
// std::discrete_distribution.
__device__ int block_sample(int n, float temperature, int top_k, float u, int eos_id, SampleSmem * sm) {
    float * l = sm->l;
    for (int i = threadIdx.x; i < n; i += kThreads) l[i] = l[i] / temperature;
    __syncthreads();
    const float eos_logit = (eos_id >= 0) ? l[eos_id] : 0.f;
    if (top_k > 0 && top_k < n) {
        const float thr = block_kth_largest(l, n, top_k, sm);
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
        sm->p[i] = e;
        sum += (double) e;
    }
    sum = block_sum_d(sum, sm->red);
    // ordered compaction of the non-zero probabilities (zeros do not move the running sum)
    const int per = (n + kThreads - 1) / kThreads;
    const int b0 = threadIdx.x * per, b1 = min(b0 + per, n);
    int cnt = 0;
    for (int i = b0; i < b1; ++i) {
        const float pr = (float) ((double) sm->p[i] / sum);
        sm->p[i] = pr;
        cnt += (pr != 0.0f);
    }
    // block exclusive scan of cnt
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
    for (int i = b0; i < b1; ++i) if (sm->p[i] != 0.0f) { sm->idx[pos++] = i; }
    if (threadIdx.x == kThreads - 1) sm->count = base + incl;
    __syncthreads();
    if (threadIdx.x == 0) {
        float cum = 0.f;
        int tok = n - 1;
        const int m = sm->count;
        for (int t = 0; t < m; ++t) {
            const int i = sm->idx[t];
            cum += sm->p[i];
            if (cum >= u) { tok = i; break; }
        }
        sm->token = tok;
    }
    __syncthreads();
    return sm->token;
}

__device__ void stage_sample_cb0(const Op & op, StreamState * st, float * smem_f) {
    if (blockIdx.x != 0) return;
    SampleSmem * sm = reinterpret_cast<SampleSmem *>(smem_f);
    const int n = op.n_vocab;
    const int frame = st->frame;
    const float rep = st->rep_pen;
    const float temperature = (op.temperature_override < -1e29f) ? st->temperature : op.temperature_override;
    const int top_k = (op.temperature_override < -1e29f) ? st->top_k : op.top_k_override;
    for (int i = threadIdx.x; i < n; i += kThreads) {
        float v = __ldcg(op.logits + i);
        if (i >= op.suppress_start && i != op.eos_id) v = -CUDART_INF_F;      // :2418-2422
        if (rep != 1.0f && ((st->seen[i >> 5] >> (i & 31)) & 1u)) {           // :2425-2435
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
            sm->l[op.eos_id] += ramp * (target - sm->l[op.eos_id]);
        }
        __syncthreads();
    }
    if (op.dump) for (int i = threadIdx.x; i < n; i += kThreads) op.dump[i] = sm->l[i];
    const int ignore_eos = st->ignore_eos;
    if (ignore_eos) {   // benchmark switch (not reference behaviour): fixed-length utterances
        if (threadIdx.x == 0) sm->l[op.eos_id] = -CUDART_INF_F;
        __syncthreads();
    }
    int tok;
    if (temperature <= 0.0f) {
        tok = block_argmax(sm->l, n, sm);
    } else {
        const float u = op.uniforms ? __ldcg(op.uniforms + frame)
                                    : philox_uniform(st->seed, (uint32_t) frame, 0u, 0x0cb0u);
        tok = block_sample(n, temperature, top_k, u, ignore_eos ? -1 : op.eos_id, sm);
    }
    if (threadIdx.x == 0) {
        if (tok == op.eos_id) {
            st->done = 1; st->eos = 1;
        } else {
            st->cur_codes[0] = tok;
            st->feed_codes[0] = tok;
            st->seen[tok >> 5] |= 1u << (tok & 31);
        }
    }
}

__device__ void stage_sample_cp(const Op & op, StreamState * st, float * smem_f) {
    if (blockIdx.x != 0) return;
    SampleSmem * sm = reinterpret_cast<SampleSmem *>(smem_f);
    const int n = op.n_vocab;
    const float temperature = (op.temperature_override < -1e29f) ? st->temperature : op.temperature_override;
    const int top_k = (op.temperature_override < -1e29f) ? st->top_k : op.top_k_override;
    for (int i = threadIdx.x; i < n; i += kThreads) {
        const float v = __ldcg(op.logits + i);
        sm->l[i] = v;
        if (op.dump) op.dump[i] = v;
    }
    __syncthreads();
    int tok;
    if (temperature <= 0.0f) {
        tok = block_argmax(sm->l, n, sm);
    } else {
        const float u = op.uniforms ? __ldcg(op.uniforms + (op.book - 1))
                                    : philox_uniform(st->seed, (uint32_t) st->frame, (uint32_t) op.book, 0x0c9du);
        tok = block_sample(n, temperature, top_k, u, -1, sm);
    }
    if (threadIdx.x == 0) {
        st->cur_codes[op.book] = tok;
        st->feed_codes[op.book] = op.forced ? __ldcg(op.forced + (op.book - 1)) : tok;
    }
}

__device__ void stage_frame_end(const Op & op, StreamState * st) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    const int f = st->frame;
    for (int c = 0; c < op.n_books; ++c) op.codes_out[(size_t) f * op.n_books + c] = st->cur_codes[c];
    st->frame = f + 1;
    if (f + 1 >= st->max_len) st->done = 1;     // src/tts_transformer.cpp:2525-2527
}

__device__ void stage_copy(const Op & op) {
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
// persistent = 1: cooperative launch, grid barrier between stages, loops over n_frames frames.
// persistent = 0: runs exactly ops[0] (host launches stage by stage).
extern "C" __global__ void __launch_bounds__(kThreads, 1)
qtts_run_program(const Op * __restrict__ ops, int n_ops, StreamState * st, GridBarrier * bar,
                 int n_frames, int persistent, int check_done) {
    extern __shared__ __align__(16) float smem[];
    __shared__ double red[kWarps];
    unsigned int epoch = 0;
    // bar->count is zeroed by the host (stream-ordered memset) before every persistent launch
    for (int f = 0; f < n_frames; ++f) {
        if (check_done && __ldcg(&st->done)) break;
        bool stop = false;
        for (int i = 0; i < n_ops && !stop; ++i) {
            const Op & op = ops[i];
            switch (op.type) {
                case OP_GEMV:       stage_gemv(op, st, smem, red); break;
                case OP_ATTN:       stage_attn(op, st, smem); break;
                case OP_SAMPLE_CB0: stage_sample_cb0(op, st, smem); break;
                case OP_SAMPLE_CP:  stage_sample_cp(op, st, smem); break;
                case OP_FRAME_END:  stage_frame_end(op, st); break;
                case OP_COPY:       stage_copy(op); break;
            }
            if (persistent) {
                if (i + 1 < n_ops) gemv_prefetch(ops[i + 1]);
                grid_barrier(bar, epoch);
                if (check_done && (op.type == OP_SAMPLE_CB0 || op.type == OP_FRAME_END) && __ldcg(&st->done))
                    stop = true;
            }
        }
        if (stop) break;
    }
}

size_t smem_required(int k_max, int max_ctx) {
    size_t a = (size_t) k_max * sizeof(float);
    size_t b = sizeof(SampleSmem);
    size_t c = ((sizeof(AttnSmem) + 3) / 4) * 4 + 3 * (size_t) ((max_ctx + 8) & ~3) * sizeof(float);
    size_t m = a > b ? a : b;
    m = m > c ? m : c;
    return (m + 127) & ~(size_t) 127;
}

}  // namespace qtts
