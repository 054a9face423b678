"""TEST INFRASTRUCTURE — numpy restatement of the reference's codec-frame loop.

This is the CPU oracle of the hot path: a plain restatement of what
`qwen3_tts::TTSTransformer` computes when it runs on the ggml CPU backend, INCLUDING the
rounding points of that execution (F16 rounding of every mat-mul activation, F16 KV cache,
sequential F16 accumulator of the flash-attention kernel, F16 softmax probabilities in the
non-flash attention).  Each function cites the reference file:line it follows
(paths relative to the reference tree; `ops.cpp` = ggml/src/ggml-cpu/ops.cpp).

Pinning: tests/test_oracle.py checks this module against tests/golden/*.npz, which were
produced by the UNMODIFIED reference compiled from /root/reference (oracle/Makefile.ref ->
oracle/_ref/libqtts_ref.so) with tests/golden/make_golden.py.  The reference itself ships no
golden vectors for this path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module;
the product (qwen3-tts-jetson_b200/) never does.
"""
from __future__ import annotations

import numpy as np

F32 = np.float32


def f16r(x):
    """round-to-nearest-even F32 -> F16 -> F32 (ggml_cpu_fp32_to_fp16, F16C `vcvtps2ph`)."""
    return np.asarray(x, dtype=F32).astype(np.float16).astype(F32)


# ----------------------------------------------------------------------------------------------
# weights
# ----------------------------------------------------------------------------------------------
class Model:
    """Tensors of one GGUF as float32 arrays (F16 -> F32 is exact) + config dict.

    Tensor names: scripts/convert_tts_to_gguf.py:42-109; consumed by create_tensors
    src/tts_transformer.cpp:444-731."""

    def __init__(self, cfg: dict, tensors: dict):
        self.cfg = cfg
        self.t = tensors

    @classmethod
    def from_gguf(cls, path: str, parse_config):
        import gguf
        rd = gguf.GGUFReader(path)
        fields = {}
        for key, f in rd.fields.items():
            if len(f.types) == 1 and f.types[0] not in (gguf.GGUFValueType.STRING, gguf.GGUFValueType.ARRAY):
                fields[key] = f.parts[f.data[0]][0]
        cfg = parse_config(fields)
        tensors = {}
        for t in rd.tensors:
            if t.name.startswith("talker.") or t.name.startswith("code_pred."):
                tensors[t.name] = np.asarray(t.data).astype(F32)
        return cls(cfg, tensors)

    def layer(self, prefix: str, i: int) -> dict:
        p = f"{prefix}.blk.{i}."
        return {k: self.t.get(p + k + ".weight") for k in
                ("attn_norm", "attn_q", "attn_k", "attn_v", "attn_output", "attn_q_norm", "attn_k_norm",
                 "ffn_norm", "ffn_gate", "ffn_up", "ffn_down")}


# ----------------------------------------------------------------------------------------------
# ggml-cpu operators, as the reference's graphs use them
# ----------------------------------------------------------------------------------------------
def mul_mat(W, x):
    """ggml_mul_mat, F16 weight x F32 activation: the activation is converted to F16
    (vec_dot_type of F16, ggml/src/ggml-cpu/ggml-cpu.c:214-219,1241-1243), products are
    accumulated in F32 (ggml_vec_dot_f16).  W: [out, in] float32 holding F16 values."""
    return (W @ f16r(x).T).T.astype(F32) if np.ndim(x) == 2 else (W @ f16r(x)).astype(F32)


def rms_norm(x, eps):
    """ggml_compute_forward_rms_norm_f32 (ops.cpp:3709-3758): double sum of float squares,
    mean and scale in float, y = x*scale."""
    x = np.asarray(x, F32)
    ss = (x * x).astype(np.float64).sum(axis=-1)
    mean = (ss / x.shape[-1]).astype(F32)
    scale = (F32(1.0) / np.sqrt(mean + F32(eps), dtype=F32)).astype(F32)
    return (x * scale[..., None]).astype(F32)


def silu(x):
    """ggml_silu: x / (1 + exp(-x)) (ggml/src/ggml-cpu/vec.h ggml_v_silu / ggml_silu_f32)."""
    x = np.asarray(x, F32)
    return (x / (F32(1.0) + np.exp(-x, dtype=F32))).astype(F32)


def rope_table(max_pos: int, theta_base: float, n_dims: int = 128):
    """ggml_rope_cache_init (ops.cpp:5612-5627): theta = pos, multiplied by
    theta_scale = powf(base, -2/n_dims) pair after pair in float; cosf/sinf."""
    theta_scale = F32(np.power(F32(theta_base), F32(-2.0) / F32(n_dims), dtype=F32))
    cos = np.zeros((max_pos, n_dims // 2), F32)
    sin = np.zeros((max_pos, n_dims // 2), F32)
    for p in range(max_pos):
        theta = F32(p)
        for i in range(n_dims // 2):
            cos[p, i] = np.cos(theta, dtype=F32)
            sin[p, i] = np.sin(theta, dtype=F32)
            theta = F32(theta * theta_scale)
    return cos, sin


def rope_neox(x, cos, sin):
    """rotate_pairs with n_offset = n_dims/2 (ops.cpp:5700-5716): x [..., 128]."""
    h = x.shape[-1] // 2
    x0, x1 = x[..., :h], x[..., h:]
    return np.concatenate([x0 * cos - x1 * sin, x0 * sin + x1 * cos], axis=-1).astype(F32)


def softmax_rows(s):
    """ggml_compute_forward_soft_max_f32 (ops.cpp:5224-5330): exp(x - max), double sum,
    scale by (float)(1/sum)."""
    mx = s.max(axis=-1, keepdims=True)
    e = np.exp((s - mx).astype(F32), dtype=F32)
    inv = (1.0 / e.astype(np.float64).sum(axis=-1, keepdims=True)).astype(F32)
    return (e * inv).astype(F32)


def attn_softmax_p16(q, K16, V16, scale):
    """Non-flash attention of the prefill and code-predictor step graphs
    (src/tts_transformer.cpp:1328-1337, 1783-1792): KQ = mul_mat(K_f16, Q -> F16), scale,
    soft_max, KQV = mul_mat(V^T_f16, P -> F16).  q [nh,128]; K16/V16 [n_kv, nkv, 128] (values
    already F16-representable)."""
    nh, nkv = q.shape[0], K16.shape[1]
    g = nh // nkv
    out = np.zeros_like(q)
    q16 = f16r(q)
    for h in range(nh):
        k = K16[:, h // g, :]
        v = V16[:, h // g, :]
        s = ((k @ q16[h]).astype(F32) * F32(scale)).astype(F32)
        p = f16r(softmax_rows(s[None, :])[0])
        out[h] = (v.T @ p).astype(F32)
    return out


def attn_f32(q, K, V, scale):
    """Code-predictor 2-token prefill (src/tts_transformer.cpp:1613-1626): K, V, Q, P all F32
    (the graph multiplies the un-cached Kcur/Vcur tensors)."""
    nh, nkv = q.shape[0], K.shape[1]
    g = nh // nkv
    out = np.zeros_like(q)
    for h in range(nh):
        k = K[:, h // g, :]
        v = V[:, h // g, :]
        s = ((k @ q[h]).astype(F32) * F32(scale)).astype(F32)
        p = softmax_rows(s[None, :])[0]
        out[h] = (v.T @ p).astype(F32)
    return out


def _flash_chunk(q16, k, v, scale):
    """ggml_compute_forward_flash_attn_ext_f16_one_chunk (ops.cpp:8108-8344) for one query row
    of every head, V type F16: online softmax with an F16 accumulator.
    q16 [nh,128] F16 values; k, v [n, nh, 128] (already expanded per query head).
    Returns (M [nh], S [nh], VKQ32 [nh,128])."""
    nh = q16.shape[0]
    M = np.full(nh, -np.inf, F32)
    S = np.zeros(nh, F32)
    acc = np.zeros((nh, 128), F32)     # holds F16-representable values
    for j in range(k.shape[0]):
        s = (np.einsum("hd,hd->h", k[j], q16).astype(F32) * F32(scale)).astype(F32)
        new_max = s > M
        with np.errstate(invalid="ignore"):
            ms = np.where(new_max, np.exp((M - s).astype(F32), dtype=F32), F32(1.0)).astype(F32)
            vs = np.where(new_max, F32(1.0), np.exp((s - M).astype(F32), dtype=F32)).astype(F32)
        ms = np.where(np.isnan(ms), F32(0.0), ms)          # exp(-inf - s) = 0 on the first key
        M = np.where(new_max, s, M)
        # ggml_vec_scale_f16 only when a new maximum arrived, then ggml_vec_mad_f16 (FMA, F16 store)
        scaled = f16r(acc * ms[:, None])
        acc = np.where(new_max[:, None], scaled, acc)
        acc = f16r(np.float64(v[j]) * np.float64(vs[:, None]) + np.float64(acc))   # single rounding == FMA -> F16
        S = (S * ms + vs).astype(F32)
    return M, S, acc


def attn_flash(q, K16, V16, scale, n_threads=4):
    """ggml_flash_attn_ext as the Talker step graph calls it (src/tts_transformer.cpp:1473):
    Q -> F16, F16 K/V.  n_kv >= 512 takes the split-KV path (ops.cpp:8740-8772): one chunk per
    thread, partials merged in order by ggml_flash_attn_ext_reduce_partials (ops.cpp:8652-8688)."""
    nh, nkv = q.shape[0], K16.shape[1]
    g = nh // nkv
    idx = np.arange(nh) // g
    q16 = f16r(q)
    k = K16[:, idx, :]
    v = V16[:, idx, :]
    n_kv = k.shape[0]
    if n_kv >= 512 and n_threads > 1:
        chunk = (n_kv + n_threads - 1) // n_threads
        Mf = np.full(nh, -np.inf, F32)
        Sf = np.zeros(nh, F32)
        out = np.zeros((nh, 128), F32)
        for c in range(n_threads):
            c0, c1 = c * chunk, min((c + 1) * chunk, n_kv)
            if c0 >= n_kv:
                continue
            Mc, Sc, acc = _flash_chunk(q16, k[c0:c1], v[c0:c1], scale)
            Mn = np.maximum(Mf, Mc)
            with np.errstate(invalid="ignore"):
                so = np.exp((Mf - Mn).astype(F32), dtype=F32)
            so = np.where(np.isnan(so), F32(0.0), so)
            sn = np.exp((Mc - Mn).astype(F32), dtype=F32)
            out = (out * so[:, None] + acc * sn[:, None]).astype(F32)
            Sf = (Sf * so + Sc * sn).astype(F32)
            Mf = Mn
        return (out * (F32(1.0) / Sf)[:, None]).astype(F32)
    M, S, acc = _flash_chunk(q16, k, v, scale)
    return (acc * (F32(1.0) / S)[:, None]).astype(F32)


# ----------------------------------------------------------------------------------------------
# the transformer
# ----------------------------------------------------------------------------------------------
class KVCache:
    """tts_kv_cache (src/tts_transformer.h:137-149): F16, token-major."""

    def __init__(self, n_layers, n_ctx, n_kv_heads):
        self.k = np.zeros((n_layers, n_ctx, n_kv_heads, 128), F32)   # F16-representable values
        self.v = np.zeros((n_layers, n_ctx, n_kv_heads, 128), F32)


class Restatement:
    def __init__(self, model: Model, n_ctx: int = 1024, n_threads: int = 4):
        self.m = model
        c = model.cfg
        self.c = c
        self.n_threads = n_threads
        self.cos, self.sin = rope_table(max(n_ctx, 16), c["rope_theta"])
        self.cache = KVCache(c["n_layers"], n_ctx, c["n_kv_heads"])
        self.cp_cache = KVCache(c["cp_layers"], 16, c["cp_n_kv_heads"])
        self.talker_layers = [model.layer("talker", i) for i in range(c["n_layers"])]
        self.cp_layers = [model.layer("code_pred", i) for i in range(c["cp_layers"])]
        self.last_hidden = None

    # -- one decoder layer on one token ----------------------------------------------------------
    def _qkv(self, L, x, pos, nh, nkv):
        eps = self.c["rms_norm_eps"]
        cur = rms_norm(x, eps) * L["attn_norm"]
        q = mul_mat(L["attn_q"], cur).reshape(nh, 128)
        k = mul_mat(L["attn_k"], cur).reshape(nkv, 128)
        v = mul_mat(L["attn_v"], cur).reshape(nkv, 128)
        if L["attn_q_norm"] is not None:
            q = rms_norm(q, eps) * L["attn_q_norm"]
        if L["attn_k_norm"] is not None:
            k = rms_norm(k, eps) * L["attn_k_norm"]
        q = rope_neox(q, self.cos[pos], self.sin[pos])
        k = rope_neox(k, self.cos[pos], self.sin[pos])
        return q, k, v

    def _mlp(self, L, x, attn_out):
        eps = self.c["rms_norm_eps"]
        x = (mul_mat(L["attn_output"], attn_out.reshape(-1)) + x).astype(F32)
        cur = rms_norm(x, eps) * L["ffn_norm"]
        act = silu(mul_mat(L["ffn_gate"], cur)) * mul_mat(L["ffn_up"], cur)
        return (mul_mat(L["ffn_down"], act) + x).astype(F32)

    # -- Talker: build_step_graph / build_prefill_forward_graph (src/tts_transformer.cpp:1233-1512)
    def talker_token(self, x, pos, flash: bool):
        c = self.c
        scale = F32(1.0) / np.sqrt(F32(128.0))
        for il, L in enumerate(self.talker_layers):
            q, k, v = self._qkv(L, x, pos, c["n_heads"], c["n_kv_heads"])
            self.cache.k[il, pos] = f16r(k)
            self.cache.v[il, pos] = f16r(v)
            K, V = self.cache.k[il, :pos + 1], self.cache.v[il, :pos + 1]
            a = attn_flash(q, K, V, scale, self.n_threads) if flash else attn_softmax_p16(q, K, V, scale)
            x = self._mlp(L, x, a)
        hidden = (rms_norm(x, c["rms_norm_eps"]) * self.m.t["talker.output_norm.weight"]).astype(F32)
        logits = mul_mat(self.m.t["talker.codec_head.weight"], hidden)
        return hidden, logits

    def forward_prefill(self, embd, n_past=0):
        """forward_prefill (src/tts_transformer.cpp:1829-1920).  The causal graph is evaluated
        token by token: every key/value it reads is the F16 cache copy, so rows are independent."""
        for t in range(embd.shape[0]):
            hidden, logits = self.talker_token(np.asarray(embd[t], F32), n_past + t, flash=False)
        self.last_hidden = hidden
        return hidden, logits

    def forward_step(self, step_embd, n_past):
        """forward_step (src/tts_transformer.cpp:1952-2028)."""
        hidden, logits = self.talker_token(np.asarray(step_embd, F32), n_past, flash=True)
        self.last_hidden = hidden
        return logits, hidden

    # -- code predictor (src/tts_transformer.cpp:1514-1827, 2153-2340) -------------------------------
    def _cp_input(self, vec):
        t = self.m.t
        if "code_pred.mtp_proj.weight" in t:
            vec = mul_mat(t["code_pred.mtp_proj.weight"], vec)
            if "code_pred.mtp_proj.bias" in t:
                vec = (vec + t["code_pred.mtp_proj.bias"]).astype(F32)
        return vec

    def predict_codes(self, hidden, cb0, temperature=0.0, top_k=50, uniforms=None, forced=None):
        """Returns (codes[15], logits[15, cp_vocab]).  Positions 0,1 = the 2-token prefill graph
        (all-F32 attention over the un-cached K/V); positions >= 2 = step graphs."""
        c, t = self.c, self.m.t
        nh, nkv = c["cp_n_heads"], c["cp_n_kv_heads"]
        scale = F32(1.0) / np.sqrt(F32(128.0))
        n_codes = c["n_codebooks"] - 1
        codes = np.zeros(n_codes, np.int32)
        logits_all = np.zeros((n_codes, c["cp_vocab_size"]), F32)
        kv_f32 = [[None, None] for _ in self.cp_layers]   # unrounded K,V of positions 0,1
        feed = int(cb0)
        for p in range(n_codes + 1):
            if p == 0:
                x = self._cp_input(np.asarray(hidden, F32))
            elif p == 1:
                x = self._cp_input(t["talker.codec_embd.weight"][feed])
            else:
                x = self._cp_input(t[f"code_pred.codec_embd.{p - 2}.weight"][feed])
            for il, L in enumerate(self.cp_layers):
                q, k, v = self._qkv(L, x, p, nh, nkv)
                self.cp_cache.k[il, p] = f16r(k)
                self.cp_cache.v[il, p] = f16r(v)
                if p < 2:
                    kv_f32[il][p] = (k, v)
                    K = np.stack([kv_f32[il][j][0] for j in range(p + 1)])
                    V = np.stack([kv_f32[il][j][1] for j in range(p + 1)])
                    a = attn_f32(q, K, V, scale)
                else:
                    a = attn_softmax_p16(q, self.cp_cache.k[il, :p + 1], self.cp_cache.v[il, :p + 1], scale)
                x = self._mlp(L, x, a)
            if p >= 1:
                h = (rms_norm(x, c["rms_norm_eps"]) * t["code_pred.output_norm.weight"]).astype(F32)
                lg = mul_mat(t[f"code_pred.lm_head.{p - 1}.weight"], h)
                logits_all[p - 1] = lg
                if temperature <= 0:
                    tok = argmax_first(lg)
                else:
                    tok = sample_topk(lg, temperature, top_k, float(uniforms[p - 1]), eos_id=-1)
                codes[p - 1] = tok
                feed = int(forced[p - 1]) if forced is not None else tok
        return codes, logits_all

    # -- host-side embedding assembly (src/tts_transformer.cpp:1026-1231) ------------------------------
    def project_text_tokens(self, ids):
        t = self.m.t
        rows = t["talker.text_embd.weight"][np.asarray(ids)]
        cur = (mul_mat(t["talker.text_proj.fc1.weight"], rows) + t["talker.text_proj.fc1.bias"]).astype(F32)
        cur = silu(cur)
        return (mul_mat(t["talker.text_proj.fc2.weight"], cur) + t["talker.text_proj.fc2.bias"]).astype(F32)

    def build_prefill(self, tokens, speaker=None, language_id=2050):
        c, t = self.c, self.m.t
        tokens = list(int(x) for x in tokens)
        if len(tokens) < 4:
            raise ValueError("Need at least 4 text tokens for prefill")
        bos, eos, pad = self.project_text_tokens([c["tts_bos_id"], c["tts_eos_id"], c["tts_pad_id"]])
        role = self.project_text_tokens(tokens[:3])
        if language_id < 0:
            ids = [c["codec_nothink_id"], c["codec_think_bos_id"], c["codec_think_eos_id"]]
        else:
            ids = [c["codec_think_id"], c["codec_think_bos_id"], language_id, c["codec_think_eos_id"]]
        ce = t["talker.codec_embd.weight"]
        rows = [ce[i] for i in ids]
        if speaker is not None:
            rows.append(np.asarray(speaker, F32))
        rows += [ce[c["codec_pad_id"]], ce[c["codec_bos_id"]]]
        codec_in = np.stack(rows).astype(F32)
        n_ov = codec_in.shape[0] - 1
        overlay = np.stack([pad] * (n_ov - 1) + [bos])
        plus = (overlay + codec_in[:n_ov]).astype(F32)
        first = (self.project_text_tokens(tokens[3:4])[0] + codec_in[-1]).astype(F32)
        prefill = np.concatenate([role, plus, first[None, :]]).astype(F32)
        n_trail = max(0, len(tokens) - 9)
        trail = [self.project_text_tokens(tokens[4:4 + n_trail])] if n_trail else []
        trailing = np.concatenate(trail + [eos[None, :]]).astype(F32)
        return prefill, trailing, pad

    # -- generate (src/tts_transformer.cpp:2342-2574) ------------------------------------------------
    def generate(self, tokens, speaker=None, max_len=64, language_id=2050, rep_pen=1.05,
                 temperature=0.0, top_k=50, uniforms_cb0=None, uniforms_cp=None):
        c, t = self.c, self.m.t
        prefill, trailing, pad = self.build_prefill(tokens, speaker, language_id)
        hidden, logits = self.forward_prefill(prefill, 0)
        n_past = prefill.shape[0]
        seen = set()
        out = []
        for frame in range(max_len):
            lg = cb0_postprocess(logits, seen, frame, len(tokens), rep_pen, c["codec_eos_id"], c["codec_vocab_size"])
            if temperature <= 0:
                tok = argmax_first(lg)
            else:
                tok = sample_topk(lg, temperature, top_k, float(uniforms_cb0[frame]), eos_id=c["codec_eos_id"])
            if tok == c["codec_eos_id"]:
                break
            seen.add(tok)
            u = None if uniforms_cp is None else uniforms_cp[frame]
            codes, _ = self.predict_codes(self.last_hidden, tok, temperature, top_k, u)
            out.append([tok] + codes.tolist())
            if frame + 1 >= max_len:
                break
            step = t["talker.codec_embd.weight"][tok].copy()
            for cb in range(1, c["n_codebooks"]):
                step = (step + t[f"code_pred.codec_embd.{cb - 1}.weight"][codes[cb - 1]]).astype(F32)
            step = (step + (trailing[frame] if frame < trailing.shape[0] else pad)).astype(F32)
            logits, hidden = self.forward_step(step, n_past)
            n_past += 1
        return np.asarray(out, np.int32).reshape(-1, c["n_codebooks"])


# ----------------------------------------------------------------------------------------------
# sampling (src/tts_transformer.cpp:2051-2061, 2198-2236, 2417-2495)
# ----------------------------------------------------------------------------------------------
def argmax_first(x) -> int:
    """argmax with strict '>' (first maximum wins), src/tts_transformer.cpp:2051-2061."""
    return int(np.argmax(x))


def cb0_postprocess(logits, seen, frame, n_text_tokens, rep_pen, eos_id, vocab):
    """src/tts_transformer.cpp:2417-2445: suppress the last 1024 ids except EOS, repetition penalty
    over the SET of emitted tokens, EOS boost after max(20, 4*n_tokens) frames."""
    lg = np.array(logits, F32)
    keep = lg[eos_id]
    lg[vocab - 1024:] = -np.inf
    lg[eos_id] = keep
    if rep_pen != 1.0:
        for tok in seen:
            lg[tok] = lg[tok] / F32(rep_pen) if lg[tok] > 0 else lg[tok] * F32(rep_pen)
    expected = max(20, n_text_tokens * 4)
    if frame >= expected:
        ramp = min(F32(1.0), F32(frame - expected) / F32(expected))
        target = F32(lg.max() + F32(5.0))
        # `logits[eos] += ramp * (target - logits[eos])`: GCC contracts this into one FMA wherever the ISA has it
        # (-ffp-contract=fast is its default; the reference builds with -march=native, the oracle with -mfma), so the
        # product is not rounded.  float32 x float32 is exact in float64; the sum is rounded once.
        lg[eos_id] = F32(np.float64(F32(ramp)) * np.float64(F32(target - lg[eos_id])) + np.float64(lg[eos_id]))
    return lg


def sample_topk(logits, temperature, top_k, u, eos_id=-1) -> int:
    """Temperature, top-k threshold keeping ties (mask `< threshold`), EOS re-inserted for
    codebook 0 (src/tts_transformer.cpp:2451-2475), softmax with float exp and double sum
    (:2476-2484), sequential float CDF scan against one uniform (:2487-2494; the TRT sampler
    does the same for the code predictor, src/trt_cuda_kernels.cu:160-180)."""
    lg = (np.array(logits, F32) / F32(temperature)).astype(F32)
    n = lg.shape[0]
    eos_logit = lg[eos_id] if eos_id >= 0 else None
    if 0 < top_k < n:
        thr = np.partition(lg, n - top_k)[n - top_k]
        lg[lg < thr] = -np.inf
    if eos_id >= 0:
        lg[eos_id] = eos_logit
    e = np.exp((lg - lg.max()).astype(F32), dtype=F32)
    probs = (e.astype(np.float64) / e.astype(np.float64).sum()).astype(F32)
    if u <= 0.0:
        return 0      # the scan starts with cumsum = probs[0] >= 0: a draw of exactly 0 selects index 0 (:2490-2494)
    cum = F32(0.0)
    for i in np.nonzero(probs)[0]:
        cum = F32(cum + probs[i])
        if cum >= F32(u):
            return int(i)
    return n - 1
