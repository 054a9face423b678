"""Protocol stress / determinism tests (round-1 verdict, "what's weak" #11): both frame kernels hand data between
stages through hand-rolled memory ordering — flag-carrying 64-bit words polled with relaxed loads (batch-1 kernel),
row counters and relaxed-poll grid barriers, tcgen05.commit / mbarrier chains and async-proxy fences (batched
kernel).  A protocol slip shows up as a rare wrong value, so these tests run LONG (thousands of frames) and require
bit-equal codes (1) across three repetitions and (2) against the same device code with the protocol replaced by
grid barriers (QTTS_LL=0: barrier after every stage; QTTS_BROWS=0: barrier instead of the row counter)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def make_engine(pkg, path, env, **kw):
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        T = pkg.TTSTransformer(device=0, **kw)
        assert T.load_model(path), T.get_error()
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    T.set_numerics(pkg.NUMERICS_FAST, 4)
    T.set_ignore_eos(True)
    return T


def test_persistent_kernel_2000_frames_vs_barrier_path(pkg, make_gguf):
    """0.6B, batch 1, temperature 0.9: 2000 frames (50 launches of 40 frames, KV up to 2010 keys: the cross-CTA
    split-KV merge included) three times with the flag-carrying hand-off, once with grid barriers only."""
    path, cfg = make_gguf("0.6b", 1234)
    n = 2000
    toks = pkg.synth.synth_tokens(cfg, n_text=300, seed=3)
    spk = pkg.synth.synth_speaker(cfg, seed=3)
    T = make_engine(pkg, path, {"QTTS_LL": "1"}, max_streams=1, max_ctx=n + 32)
    runs = []
    for rep in range(3):
        out = T.generate(toks, spk, max_len=n, temperature=0.9, top_k=50, seed=11)
        assert out is not None and len(out) == n, T.get_error()
        runs.append(out)
    ms_ll = T.last_loop_ms() / n
    T.unload_model()
    T = make_engine(pkg, path, {"QTTS_LL": "0"}, max_streams=1, max_ctx=n + 32)
    ref = T.generate(toks, spk, max_len=n, temperature=0.9, top_k=50, seed=11)
    ms_bar = T.last_loop_ms() / n
    T.unload_model()
    assert ref is not None and len(ref) == n
    for rep, out in enumerate(runs):
        bad = np.nonzero((out != ref).any(axis=1))[0]
        assert bad.size == 0, f"repetition {rep}: codes differ from the barrier path from frame {bad[0]} on"
    print(f"2000 frames x 3 bit-equal to the barrier path; {ms_ll:.3f} ms/frame flag-carrying vs {ms_bar:.3f} ms/frame barriers")


@pytest.mark.parametrize("streams", [64, 140])
def test_batched_kernel_500_steps_vs_barrier_path(pkg, make_gguf, streams):
    """0.6B, lock-step group of 64 (M = 64 tiles, deferred split-K) / 140 streams (two M tiles, split-M), ragged texts,
    temperature 0.9: 500 frame-steps three times, and once with a grid barrier after every norm stage."""
    path, cfg = make_gguf("0.6b", 1234)
    n = 500
    r = np.random.default_rng(5)
    toks = [pkg.synth.synth_tokens(cfg, n_text=int(r.integers(20, 200)), seed=i) for i in range(streams)]
    spks = [None if i % 5 == 4 else pkg.synth.synth_speaker(cfg, seed=i) for i in range(streams)]

    def run(T):
        for i in range(streams):
            assert T.stream_prepare(i, toks[i], spks[i], max_len=n, temperature=0.9, top_k=50, seed=100 + i), T.get_error()
        got = [[] for _ in range(streams)]
        while True:
            out = T.step_frames(50)
            assert out is not None, T.get_error()
            codes, produced, active = out
            for i in range(streams):
                got[i].append(codes[i, :produced[i]])
            if not active.any():
                break
        return np.stack([np.concatenate(g) for g in got])

    T = make_engine(pkg, path, {"QTTS_BROWS": "1"}, max_streams=streams, max_ctx=n + 32)
    runs = [run(T) for _ in range(3)]
    T.unload_model()
    T = make_engine(pkg, path, {"QTTS_BROWS": "0"}, max_streams=streams, max_ctx=n + 32)
    ref = run(T)
    T.unload_model()
    assert ref.shape == (streams, n, 16)
    for rep, out in enumerate(runs):
        bad = np.nonzero((out != ref).any(axis=(1, 2)))[0]
        assert bad.size == 0, f"repetition {rep}: {bad.size} streams differ from the barrier path (first: stream {bad[0]})"


def test_batched_cluster_multicast_equals_plain_launch(pkg, make_gguf):
    """The opt-in cluster mode of qtts_run_batched_ext (QTTS_BCLUSTER=1: CTA pairs, every activation piece copied half by
    each CTA and multicast to both, buffers recycled by multicast tcgen05.commit) only changes how operands travel: codes
    must be bit-equal to the plain launch, for a group with two M tiles (140 streams, split-M re-dealt for the pairs) and
    for a group of <= 64 rows."""
    path, cfg = make_gguf("0.6b", 1234)
    n = 60
    for streams in (140, 48):
        r = np.random.default_rng(7)
        toks = [pkg.synth.synth_tokens(cfg, n_text=int(r.integers(20, 120)), seed=i) for i in range(streams)]
        spks = [pkg.synth.synth_speaker(cfg, seed=i) for i in range(streams)]

        def run(T):
            for i in range(streams):
                assert T.stream_prepare(i, toks[i], spks[i], max_len=n, temperature=0.9, top_k=50, seed=300 + i), T.get_error()
            got = [[] for _ in range(streams)]
            while True:
                out = T.step_frames(30)
                assert out is not None, T.get_error()
                codes, produced, active = out
                for i in range(streams):
                    got[i].append(codes[i, :produced[i]])
                if not active.any():
                    break
            return np.stack([np.concatenate(g) for g in got])

        T = make_engine(pkg, path, {"QTTS_BCLUSTER": "0"}, max_streams=streams, max_ctx=n + 160)
        plain = run(T)
        T.unload_model()
        T = make_engine(pkg, path, {"QTTS_BCLUSTER": "1"}, max_streams=streams, max_ctx=n + 160)
        clustered = [run(T) for _ in range(2)]
        T.unload_model()
        assert plain.shape == (streams, n, 16)
        for out in clustered:
            assert np.array_equal(out, plain), f"{streams} streams: cluster launch differs from the plain launch"
