"""Multi-GPU host logic: utterances are independent, so N GPUs = N engines, each taking a
contiguous shard of the utterance list; there is NO data-path collective.
torch.distributed carries only the launch barrier and the job-level counters (max of the device
time over ranks, sums of frames / launches), plus an optional gather of per-utterance frame counts
so rank 0 can check that every utterance was served exactly once.

Works on any backend: NCCL with CUDA tensors on the GPU box, gloo with CPU tensors in the tests."""
from __future__ import annotations

from typing import Sequence


def shard_bounds(n_items: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous balanced shard [lo, hi) of `n_items` for `rank`: the first n_items % world ranks get
    one extra item; ranks beyond n_items get an empty shard."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world: {rank}/{world}")
    if n_items < 0:
        raise ValueError("n_items must be >= 0")
    base, extra = divmod(n_items, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard(items: Sequence, rank: int, world: int):
    lo, hi = shard_bounds(len(items), rank, world)
    return items[lo:hi]


def slot_groups(n_streams: int, group: int = 256) -> list[tuple[int, int]]:
    """Slots of one GPU advance in lock-step groups of <= `group` streams (one batched launch each; the engine's
    group is 256 = Engine::bgroup, csrc/qtts_engine.h)."""
    return [(lo, min(lo + group, n_streams)) for lo in range(0, n_streams, group)]


class JobCounters:
    """max-over-ranks / sum-over-ranks reduction of a rank's counters.  `dist` is torch.distributed
    (initialised) or None for a single process."""

    def __init__(self, dist=None, device="cpu"):
        self.dist, self.device = dist, device

    def reduce(self, maxed: dict, summed: dict) -> dict:
        import torch
        out = {}
        if self.dist is None:
            out.update({k: float(v) for k, v in maxed.items()})
            out.update({k: float(v) for k, v in summed.items()})
            return out
        mk, sk = sorted(maxed), sorted(summed)
        mx = torch.tensor([float(maxed[k]) for k in mk], dtype=torch.float64, device=self.device)
        sm = torch.tensor([float(summed[k]) for k in sk], dtype=torch.float64, device=self.device)
        if mk:
            self.dist.all_reduce(mx, op=self.dist.ReduceOp.MAX)
        if sk:
            self.dist.all_reduce(sm, op=self.dist.ReduceOp.SUM)
        out.update({k: float(v) for k, v in zip(mk, mx.tolist())})
        out.update({k: float(v) for k, v in zip(sk, sm.tolist())})
        return out

    def gather_frames(self, n_items: int, lo: int, frames_per_item: Sequence[int]) -> list[int]:
        """Every rank contributes the frame counts of its shard [lo, lo+len); returns the full
        per-utterance list (-1 = served by nobody, which the caller treats as an error)."""
        import torch
        full = torch.full((n_items,), -1, dtype=torch.int64, device=self.device)
        if len(frames_per_item):
            full[lo:lo + len(frames_per_item)] = torch.as_tensor(list(frames_per_item), dtype=torch.int64, device=self.device)
        if self.dist is not None:
            # shards are disjoint and counts are >= 0, so MAX merges them without a gather buffer
            self.dist.all_reduce(full, op=self.dist.ReduceOp.MAX)
        return [int(v) for v in full.tolist()]


def throughput(frames_total: float, ms_max: float) -> float:
    """Whole-job frames/s: frames of all ranks over the slowest rank's time."""
    return frames_total / (ms_max / 1e3) if ms_max > 0 else 0.0


def serve_sharded(engine, utterances: Sequence[dict], rank: int, world: int, counters: JobCounters,
                  max_len: int, frames_per_call: int = 32, **sampling) -> dict:
    """Serve `utterances` (dicts with 'tokens' and optional 'speaker', 'seed') on `world` engines.
    This rank's shard is served in waves of at most engine.max_streams concurrent streams
    (stream_begin per slot, then step_frames until every stream of the wave is done).
    `engine` needs the TTSTransformer stream interface: max_streams, stream_begin, step_frames,
    get_error, last_loop_ms, config['n_codebooks'].
    Returns {'codes': {utterance index: int32 [frames, n_codebooks]} (this rank's shard only),
             'frames_per_utterance': full list over all ranks, 'frames', 'device_ms', 'value'}.
    device_ms (and therefore value) INCLUDES the batched Talker prefill of every wave: it runs inside the wave's
    first step_frames call, whose timed region (qtts_last_loop_ms) starts before it."""
    import numpy as np
    lo, hi = shard_bounds(len(utterances), rank, world)
    nb = engine.config["n_codebooks"]
    codes: dict[int, np.ndarray] = {}
    device_ms = 0.0
    for w0 in range(lo, hi, engine.max_streams):
        wave = list(range(w0, min(w0 + engine.max_streams, hi)))
        for slot, u in enumerate(wave):
            utt = utterances[u]
            begin = getattr(engine, "stream_prepare", None) or engine.stream_begin    # batched prefill at the first step
            if not begin(slot, utt["tokens"], utt.get("speaker"), max_len=max_len,
                                       seed=int(utt.get("seed", u)), **sampling):
                raise RuntimeError(f"stream_begin(utterance {u}): {engine.get_error()}")
        parts = {u: [] for u in wave}
        while True:
            out = engine.step_frames(frames_per_call)
            if out is None:
                raise RuntimeError("step_frames: " + engine.get_error())
            c, produced, active = out
            device_ms += engine.last_loop_ms()
            for slot, u in enumerate(wave):
                if produced[slot]:
                    parts[u].append(np.asarray(c[slot, :produced[slot]], np.int32).copy())
            if not np.asarray(active)[:len(wave)].any():
                break
        for u in wave:
            codes[u] = np.concatenate(parts[u]) if parts[u] else np.zeros((0, nb), np.int32)
    mine = [len(codes[u]) for u in range(lo, hi)]
    per_utt = counters.gather_frames(len(utterances), lo, mine)
    if any(v < 0 for v in per_utt):
        raise RuntimeError("sharding error: some utterance was served by no rank")
    red = counters.reduce({"device_ms": device_ms}, {"frames": float(sum(mine))})
    return {"codes": codes, "frames_per_utterance": per_utt, "frames": red["frames"],
            "device_ms": red["device_ms"], "value": throughput(red["frames"], red["device_ms"])}
