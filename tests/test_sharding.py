"""Multi-GPU host logic (qwen3-tts-jetson_b200/sharding.py): utterance sharding, wave scheduling and
the job-level counters, run on CPU with world_size 2 over gloo against a deterministic stand-in
engine; on the GPU the same routine drives the real engine (world 1) and must return what
generate() returns per utterance."""
import os
import socket

import numpy as np
import pytest

import qtts_pkg

NB = 16


class FakeEngine:
    """TTSTransformer's stream interface with frames that are a pure function of (tokens, seed, frame)."""

    def __init__(self, max_streams, ms_per_call):
        self.max_streams, self.ms = max_streams, ms_per_call
        self.config = {"n_codebooks": NB}
        self.slots = {}

    @staticmethod
    def n_frames(tokens, max_len):
        return min(max_len, 3 + int(np.sum(tokens)) % 11)

    @staticmethod
    def frame(tokens, seed, f):
        return (np.arange(NB, dtype=np.int64) * 31 + int(np.sum(tokens)) * 7 + (int(seed) % 1000) * 13 + f * 101) % 2048

    def stream_begin(self, slot, tokens, speaker, max_len, seed, **kw):
        if slot >= self.max_streams:
            return False
        self.slots[slot] = [np.asarray(tokens), seed, 0, self.n_frames(tokens, max_len)]
        return True

    def step_frames(self, n):
        codes = np.zeros((self.max_streams, n, NB), np.int32)
        produced = np.zeros(self.max_streams, np.int32)
        active = np.zeros(self.max_streams, np.int32)
        for s, st in self.slots.items():
            k = min(n, st[3] - st[2])
            for i in range(k):
                codes[s, i] = self.frame(st[0], st[1], st[2] + i)
            st[2] += k
            produced[s] = k
            active[s] = 1 if st[2] < st[3] else 0
        return codes, produced, active

    def last_loop_ms(self):
        return self.ms

    def get_error(self):
        return "fake"


def make_utterances(n):
    r = np.random.default_rng(5)
    return [{"tokens": r.integers(0, 4000, size=int(r.integers(4, 40))).astype(np.int32), "seed": 100 + i} for i in range(n)]


def _worker(rank, world, port, n_utt, out_dir):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sh = qtts_pkg.load().sharding
    eng = FakeEngine(max_streams=3, ms_per_call=1.0 + rank)      # rank 1 is the slow one
    res = sh.serve_sharded(eng, make_utterances(n_utt), rank, world, sh.JobCounters(dist, "cpu"), max_len=9, frames_per_call=4)
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), frames=res["frames"], device_ms=res["device_ms"], value=res["value"],
             per_utt=np.array(res["frames_per_utterance"]), keys=np.array(sorted(res["codes"])),
             **{f"c{u}": c for u, c in res["codes"].items()})
    dist.destroy_process_group()


def test_shard_bounds_cover_and_balance():
    sh = qtts_pkg.load().sharding
    for n in (0, 1, 5, 8, 511, 512):
        for world in (1, 2, 3, 8):
            b = [sh.shard_bounds(n, r, world) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sh.shard_bounds(4, 2, 2)
    assert sh.slot_groups(300) == [(0, 256), (256, 300)] and sh.slot_groups(300, 128) == [(0, 128), (128, 256), (256, 300)]
    assert sh.slot_groups(64) == [(0, 64)]


def test_single_process_counters():
    sh = qtts_pkg.load().sharding
    red = sh.JobCounters(None).reduce({"ms": 4.0}, {"frames": 10})
    assert red == {"ms": 4.0, "frames": 10.0}
    assert sh.throughput(10, 4.0) == 2500.0


@pytest.mark.parametrize("n_utt", [7, 1])
def test_world2_gloo_matches_single_process(tmp_path, n_utt):
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(2, port, n_utt, str(tmp_path)), nprocs=2, join=True)
    sh = qtts_pkg.load().sharding
    utts = make_utterances(n_utt)
    solo = sh.serve_sharded(FakeEngine(3, 1.0), utts, 0, 1, sh.JobCounters(None), max_len=9, frames_per_call=4)
    r = [np.load(tmp_path / f"rank{i}.npz") for i in range(2)]
    # both ranks hold the same job-level numbers
    for k in ("frames", "device_ms", "value"):
        assert float(r[0][k]) == float(r[1][k])
    assert list(r[0]["per_utt"]) == list(r[1]["per_utt"]) == solo["frames_per_utterance"]
    assert float(r[0]["frames"]) == solo["frames"] == sum(FakeEngine.n_frames(u["tokens"], 9) for u in utts)
    # every utterance served exactly once, by the rank its shard says, with the frames a single engine produces
    keys = [list(r[i]["keys"]) for i in range(2)]
    assert sorted(keys[0] + keys[1]) == list(range(n_utt))
    assert keys[0] == list(range(*sh.shard_bounds(n_utt, 0, 2)))
    for i in range(2):
        for u in keys[i]:
            assert np.array_equal(r[i][f"c{u}"], solo["codes"][u])
    # time = the slowest rank's (rank 1 reports 2 ms per call), value = total frames over that time
    calls = [0, 0]
    for i in range(2):
        lo, hi = sh.shard_bounds(n_utt, i, 2)
        for w0 in range(lo, hi, 3):
            calls[i] += max(-(-FakeEngine.n_frames(utts[u]["tokens"], 9) // 4) for u in range(w0, min(w0 + 3, hi)))
    assert float(r[0]["device_ms"]) == max(calls[0] * 1.0, calls[1] * 2.0)
    assert float(r[0]["value"]) == pytest.approx(solo["frames"] / (float(r[0]["device_ms"]) / 1e3))


@pytest.mark.gpu
def test_serve_sharded_real_engine_matches_generate(make_gguf):
    pkg = qtts_pkg.load()
    sh = pkg.sharding
    path, cfg = make_gguf("tiny", 1234)
    utts = [{"tokens": pkg.synth.synth_tokens(cfg, n_text=8 + 3 * i, seed=i), "speaker": pkg.synth.synth_speaker(cfg, seed=i),
             "seed": 40 + i} for i in range(5)]
    T = pkg.TTSTransformer(device=0, max_streams=3, max_ctx=256)
    assert T.load_model(path), T.get_error()
    T.set_batched(False)      # batch-1 kernel stream after stream: bit-equal to generate()
    res = sh.serve_sharded(T, utts, 0, 1, sh.JobCounters(None), max_len=12, frames_per_call=5,
                           temperature=0.9, top_k=50, repetition_penalty=1.05)
    assert sorted(res["codes"]) == list(range(5))
    for u, utt in enumerate(utts):
        want = T.generate(utt["tokens"], utt["speaker"], max_len=12, temperature=0.9, top_k=50, seed=utt["seed"])
        assert np.array_equal(res["codes"][u], want), u
    assert res["frames"] == sum(res["frames_per_utterance"]) and res["value"] > 0
    # batched kernel: same frame counts and shapes (ids can part at logit near-ties, see test_gpu_batched)
    T.set_batched(True)
    res_b = sh.serve_sharded(T, utts, 0, 1, sh.JobCounters(None), max_len=12, frames_per_call=5,
                             temperature=0.0, top_k=0, repetition_penalty=1.05)
    assert all(c.shape[1] == 16 and 1 <= len(c) <= 12 for c in res_b["codes"].values())
    T.unload_model()
