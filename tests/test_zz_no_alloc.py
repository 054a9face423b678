"""qtts_codes_chunk_noalloc (include/qtts_b200.h): the allocation-free host-destination form of the vocoder hand-off.
No counterpart in the reference; checked against the existing entry point
qtts_codes_chunk through the C ABI, bit-exact (the chunk is integer data), on consecutive calls that share the library's
static device scratch.  The file name keeps it last in pytest's collection order."""
import numpy as np
import pytest

import qtts_pkg

pytestmark = pytest.mark.gpu


def _run_stream(pkg, T, cfg, n_frames, seed):
    toks = pkg.synth.synth_tokens(cfg, n_text=12, seed=seed)
    spk = pkg.synth.synth_speaker(cfg, seed=seed)
    assert T.stream_prepare(0, toks, spk, max_len=n_frames, temperature=0.9, seed=seed), T.get_error()
    got = []
    while True:
        codes, produced, active = T.step_frames(40)
        got.append(codes[0, :produced[0]].copy())
        if not active.any():
            break
    frames = np.concatenate(got)
    assert len(frames) == n_frames
    return frames


def test_codes_chunk_noalloc_matches_codes_chunk(make_gguf):
    pkg = qtts_pkg.load()
    path, cfg = make_gguf("tiny", 1234)
    T = pkg.TTSTransformer(device=0, max_streams=1, max_ctx=160)
    assert T.load_model(path), T.get_error()
    T.set_ignore_eos(True)
    frames = _run_stream(pkg, T, cfg, 95, seed=11)
    nb = T.config["n_codebooks"]
    limit = 8192                                     # QTTS_NOALLOC_CHUNK_WORDS
    # two calls in a row (and more): a full chunk, then SHORTER ones whose zero tail must not show the previous call's
    # contents of the shared scratch, then the largest request that fits
    calls = [(0, 40, 40), (80, 15, 40), (40, 40, 40), (90, 5, 8), (0, 95, limit // nb)]
    for rep in range(2):
        for first, n, fixed in calls:
            want = T.codes_chunk(0, first, n, fixed)                 # existing entry point (allocating staging buffer)
            assert want is not None, T.get_error()
            out = np.full((1, nb, fixed), -7, np.int64)              # poisoned: every word must be written
            got = T.codes_chunk_noalloc(0, first, n, fixed, out=out)
            assert got is not None, T.get_error()
            assert got.dtype == np.int64 and np.array_equal(got, want), (rep, first, n, fixed)
            ref = np.zeros((1, nb, fixed), np.int64)
            ref[0, :, :n] = frames[first:first + n].T
            assert np.array_equal(got, ref)
    # a second utterance in the same engine: the scratch holds the first one's codes, the result must be the new one's
    frames2 = _run_stream(pkg, T, cfg, 50, seed=12)
    got = T.codes_chunk_noalloc(0, 40, 10, 40)
    assert got is not None, T.get_error()
    assert np.array_equal(got, T.codes_chunk(0, 40, 10, 40)) and np.array_equal(got[0, :, :10], frames2[40:50].T)
    assert (got[0, :, 10:] == 0).all()
    # limits and errors: beyond the static scratch, frames not generated, bad range - the engine stays usable
    assert T.codes_chunk_noalloc(0, 0, 10, limit // nb + 1) is None
    msg = T.get_error()
    assert "exceeds the static scratch" in msg and str(limit) in msg
    assert T.codes_chunk_noalloc(0, 45, 10, 40) is None and "not generated" in T.get_error()
    assert T.codes_chunk_noalloc(0, 0, 41, 40) is None and "bad frame range" in T.get_error()
    assert np.array_equal(T.codes_chunk_noalloc(0, 0, 40, 40), T.codes_chunk(0, 0, 40, 40))
    T.unload_model()


def test_codes_chunk_noalloc_two_engines_share_the_scratch(make_gguf):
    """The array belongs to the library (per GPU), not to an engine: sequential calls from two engines give each its own
    frames."""
    pkg = qtts_pkg.load()
    path, cfg = make_gguf("tiny", 1234)
    eng = []
    for seed in (21, 22):
        T = pkg.TTSTransformer(device=0, max_streams=1, max_ctx=96)
        assert T.load_model(path), T.get_error()
        T.set_ignore_eos(True)
        eng.append((T, _run_stream(pkg, T, cfg, 45, seed=seed)))
    for rep in range(2):
        for T, frames in eng:
            got = T.codes_chunk_noalloc(0, 5, 40, 40)
            assert got is not None, T.get_error()
            assert np.array_equal(got[0], frames[5:45].T) and np.array_equal(got, T.codes_chunk(0, 5, 40, 40))
    for T, _ in eng:
        T.unload_model()
