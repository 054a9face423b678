#!/usr/bin/env python
"""bench.py — codec frames/s of the Qwen3-TTS codec-frame loop on B200 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

N=1 workload = BASELINE.json configs[1]: 0.6B, one stream, temperature 0.9 / top-k 50, 512 frames.
One "step" = one utterance of 512 codec frames through the frame loop (Talker step + 16
code-predictor positions + sampling per frame).  `value` = frames/s over the frame loop timed
with CUDA events on the engine's stream (weights, KV and prefill resident in HBM); `e2e` = the
same metric through the reference-facing call (TTSTransformer.generate over the C ABI) with host
buffers: token ids / speaker H2D, prefill, frame loop, codes D2H inside the timed region.
N>1: one process per GPU, independent utterances per rank (weak scaling, no collective on the
path); torch.distributed is used only for the barrier and the max-over-ranks of the device time.

--impl reference times the UNMODIFIED reference (oracle/_ref/libqtts_ref.so, ggml CPU backend) on
the host cores on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FRAMES = 512            # frames per utterance (configs[1])
N_TEXT = 200            # synthetic text ids: EOS ramp would start at frame 4*208 > 512
PRESET, SEED = "0.6b", 1234
TEMP, TOP_K, REP_PEN = 0.9, 50, 1.05
REF_FRAMES = 24         # bounded CPU sample per step of the reference arm
B_FRAMES = 64           # frames per stream of the batched legs (configs[2] / configs[3])
B_TEXT = (32, 256)      # ragged text lengths of the batched legs
B_TOTAL = 512           # configs[3]: 512 concurrent streams sharded by utterance over the GPUs
METRIC, UNIT = "codec_frames_per_s", "frames/s"

# algorithmic bytes per frame-step, 0.6B, BF16 weights (SURVEY.md §8d / DESIGN.md):
#   talker 887,095,296 + code-pred layers 15 x 157,286,400 + lm_heads 62,914,560
WEIGHT_BYTES_PER_FRAME = 887_095_296 + 15 * 157_286_400 + 62_914_560
KV_BYTES_PER_TOKEN = 114_688      # 28 layers x 2 x 8 x 128 x 2 B, read per cached token per frame


def gguf_path() -> str:
    d = os.environ.get("QTTS_GGUF_CACHE") or tempfile.gettempdir()
    return os.path.join(d, f"synth-{PRESET}-{SEED}.gguf")


def ensure_gguf(pkg, rank: int, barrier) -> str:
    path = gguf_path()
    if rank == 0 and not os.path.exists(path):
        tmp = path + ".tmp"
        pkg.synth.write_gguf(tmp, PRESET, SEED)
        os.replace(tmp, path)
    barrier()
    return path


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.stop = index, [], threading.Event()
        self.t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            self.stop.wait(0.2)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.t.join(timeout=6)

    def summary(self) -> dict:
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v == "Active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def measured_peak_hbm():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the UNMODIFIED reference on the host cores (oracle/_ref)
# ---------------------------------------------------------------------------------------------------
def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def time_reference(path: str, toks, spk, n_frames: int, steps: int, warmup: int, threads: int):
    """Returns (frames_per_s over whole generate() calls, ms list, frames list)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ref
    if not ref.available():
        raise RuntimeError("oracle/_ref/libqtts_ref.so is missing (make -f oracle/Makefile.ref)")
    R = ref.Reference(path, threads=threads)
    ms, frames = [], []
    for i in range(warmup + steps):
        # greedy: the reference seeds its sampler from std::random_device (src/tts_transformer.h:298),
        # so a sampled run would stop at a random frame; per-frame arithmetic is identical
        codes, wall_ms = R.generate(toks, spk, max_len=n_frames, temperature=0.0, top_k=TOP_K, rep_pen=REF_PEN)
        if i >= warmup:
            ms.append(wall_ms)
            frames.append(len(codes))
    R.close()
    return sum(frames) / (sum(ms) / 1e3), ms, frames


REF_PEN = REP_PEN


def run_reference(args, pkg):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    path = ensure_gguf(pkg, 0, lambda: None)
    cfg = pkg.synth.PRESETS[PRESET]
    toks = pkg.synth.synth_tokens(cfg, n_text=N_TEXT, seed=0)
    spk = pkg.synth.synth_speaker(cfg)
    cores = host_cores()
    fps, ms, frames = time_reference(path, toks, spk, REF_FRAMES, args.steps, args.warmup, cores)
    sample = (f"{args.steps} x generate() of {REF_FRAMES} frames (prefill of {len(toks)} text tokens included), greedy, "
              f"same GGUF/text/speaker as the GPU arm, ggml CPU backend, {cores} threads")
    line = {"impl": "reference", "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": float(np.mean(ms)), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f16 weights / f32 accumulate (ggml-cpu)", "data": "synthetic",
            "config": workload_config(1),
            "cpu_baseline": {"value": fps, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
            "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


NCU_DRAM_BYTES_PER_LAUNCH = 129936908000 + 29512448


def workload_config(world: int) -> dict:
    return {"workload": "BASELINE configs[1]: Qwen3-TTS-12Hz-0.6B random-init, batch 1 per GPU, temperature 0.9 top-k 50, "
                        f"{FRAMES} frames per utterance",
            "streams_per_gpu": 1, "frames_per_step": FRAMES, "text_tokens": N_TEXT + 8,
            "eos": "suppressed by the benchmark switch qtts_set_ignore_eos so every utterance runs all frames",
            "numerics": "fast (F16-rounded activations/KV like the reference, FP32 P.V accumulation)",
            "l2": "inputs larger than L2 (3.3 GB of weights streamed per frame vs 126 MB L2)",
            "parallelism": f"{world} x independent utterance streams, no collective"}


# ---------------------------------------------------------------------------------------------------
# our arm: the CUDA engine through the C ABI
# ---------------------------------------------------------------------------------------------------
def run_ours(args, pkg):
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("no CUDA device: this benchmark has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    pkg.build.build()
    path = ensure_gguf(pkg, rank, barrier)
    cfg = pkg.synth.PRESETS[PRESET]
    toks = pkg.synth.synth_tokens(cfg, n_text=N_TEXT, seed=rank)      # a different utterance per rank
    spk = pkg.synth.synth_speaker(cfg, seed=7 + rank)
    T = pkg.TTSTransformer(device=local, max_streams=1, max_ctx=1024)
    if not T.load_model(path):
        raise RuntimeError("load_model: " + T.get_error())
    T.set_numerics(pkg.NUMERICS_FAST, 4)
    T.set_ignore_eos(True)

    def step(seed):
        t0 = time.perf_counter()
        out = T.generate(toks, spk, max_len=FRAMES, temperature=TEMP, top_k=TOP_K,
                         repetition_penalty=REP_PEN, seed=seed)
        wall = time.perf_counter() - t0
        if out is None:
            raise RuntimeError("generate: " + T.get_error())
        return len(out), T.last_loop_ms(), wall

    for i in range(args.warmup):
        step(100 + i)
    barrier()
    l0 = T.launch_count()
    with ClockSampler(local) as clocks:
        frames = loop_ms = 0.0
        t0 = time.perf_counter()
        for i in range(args.steps):
            n, ms, _ = step(i)
            frames += n
            loop_ms += ms
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
    barrier()
    launches = T.launch_count() - l0
    T.unload_model()
    batched = []
    if not args.no_batched:
        # 64 (configs[2]), 512 / world (configs[3]) and 256 = one full lock-step group per GPU
        for streams in sorted({64, 256, max(1, B_TOTAL // world)}):
            res = run_batched_leg(pkg, torch, dist, path, cfg, local, rank, world, streams, barrier)
            res["streams_per_gpu"] = streams
            batched.append(res)
    # job-level counters: device / wall time = max over ranks, frames / launches = sum over ranks
    red = pkg.sharding.JobCounters(dist, "cuda").reduce({"loop_ms": loop_ms, "wall_ms": wall * 1e3},
                                                        {"frames": frames, "launches": launches})
    loop_ms, wall_ms, frames, launches = red["loop_ms"], red["wall_ms"], red["frames"], int(red["launches"])
    value = frames / (loop_ms / 1e3)
    e2e = frames / (wall_ms / 1e3)

    if rank == 0:
        peak, peak_src = measured_peak_hbm()
        # dominant kernel = qtts_run_program (the persistent frame kernel: 40 frames per launch).
        # algorithmic bytes per frame = weights once + KV read of the cached tokens + KV append
        n_kv_mean = 10 + (FRAMES - 1) / 2.0 + 1
        bytes_per_frame = WEIGHT_BYTES_PER_FRAME + KV_BYTES_PER_TOKEN * (n_kv_mean + 1)
        per_gpu_frames_per_s = value / world
        achieved = bytes_per_frame * per_gpu_frames_per_s / 1e9
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": loop_ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": workload_config(world),
                "ms_per_frame": loop_ms / (frames / world), "rtf_per_stream": (loop_ms / 1e3) / ((frames / world) / 12.5),
                "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": int(toks.nbytes + spk.nbytes),
                        "d2h_bytes_per_step": FRAMES * 16 * 4},
                "gpu_launches": launches,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             # dram__bytes_read.sum + dram__bytes_write.sum of one 40-frame launch, from the
                             # ncu --set full capture summarised in profiles/r01h_frame_ncu.md
                             "traffic": NCU_DRAM_BYTES_PER_LAUNCH, "traffic_source": "profiles/r01h_frame_ncu.md",
                             "frames_per_launch": 40, "algorithmic_bytes_per_launch": bytes_per_frame * 40,
                             "kernel": "qtts_run_program", "peak_source": peak_src,
                             "algorithmic_bytes_per_frame": bytes_per_frame},
                "clocks": clocks.summary()}
        line["gpu_launches"] = launches + sum(b["launches"] for b in batched)
        line["batched"] = []
        for b in batched:
            S = b["streams_per_gpu"]
            fps = b["frames"] / (b["dev_ms"] / 1e3)
            step_ms = b["dev_ms"] / (b["frames"] / (S * world))
            # one frame-step of a group of <= 128 streams streams the weights once; KV per stream
            groups = (S + 127) // 128
            bytes_step = groups * WEIGHT_BYTES_PER_FRAME + S * KV_BYTES_PER_TOKEN * (b["n_kv_mean"] + 1)
            ach = bytes_step / (step_ms / 1e3) / 1e9
            line["batched"].append({
                "workload": ("BASELINE configs[2]: 64 concurrent utterances on one GPU" if S == 64 and world == 1 else
                             f"BASELINE configs[3]: {B_TOTAL} concurrent streams sharded by utterance over {world} GPU(s)"
                             if S * world == B_TOTAL else
                             f"one full lock-step group per GPU: {S} streams x {world} GPU(s), sharded by utterance") +
                            f", ragged texts {B_TEXT[0]}-{B_TEXT[1]} tokens, cached speaker embeddings, temperature {TEMP} top-k {TOP_K}, "
                            f"{B_FRAMES} frames per stream, batched tcgen05 kernel (groups of <= 256 streams in lock step)",
                "streams_per_gpu": S, "streams_total": S * world, "value": fps, "unit": UNIT,
                "ms_per_frame_step": step_ms, "rtf_per_stream": (step_ms / 1e3) * 12.5,
                "e2e": {"value": b["frames"] / (b["wall_ms"] / 1e3), "unit": UNIT,
                        "includes": "stream_prepare per stream (token/speaker H2D, embedding assembly) + one batched prefill_pending + step_frames with codes D2H",
                        "stream_begin_ms_total": b["begin_ms"]},
                "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                             "kernel": "qtts_run_batched", "algorithmic_bytes_per_frame_step": bytes_step}})
        if world == 1 and not args.no_cpu_baseline:
            try:
                cores = host_cores()
                fps, ms, fr = time_reference(path, toks, spk, 48, 1, 0, cores)
                line["cpu_baseline"] = {"value": fps, "unit": UNIT, "cores": cores, "kind": "reference",
                                        "sample": f"1 x generate() of {fr[0]} frames incl. prefill, greedy, same GGUF/text, "
                                                  f"unmodified reference on ggml CPU backend ({ms[0] / 1e3:.1f} s)"}
            except Exception as ex:   # the checker is missing: say so instead of inventing a number
                line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": f"unavailable: {ex}"}
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()
    return 0


def run_batched_leg(pkg, torch, dist, path, cfg, local, rank, world, streams, barrier):
    """`streams` concurrent utterances on this GPU advanced in lock step by the batched tcgen05 kernel.
    Returns (frames over all ranks, device ms max over ranks, wall ms max over ranks, launches, mean n_kv)."""
    n_text_max = B_TEXT[1]
    T = pkg.TTSTransformer(device=local, max_streams=streams, max_ctx=B_FRAMES + n_text_max + 32)
    if not T.load_model(path):
        raise RuntimeError("load_model (batched): " + T.get_error())
    T.set_numerics(pkg.NUMERICS_FAST, 4)
    T.set_ignore_eos(True)
    r = np.random.default_rng(1000 + rank)
    texts = [pkg.synth.synth_tokens(cfg, n_text=int(r.integers(B_TEXT[0], B_TEXT[1] + 1)), seed=rank * 4096 + i) for i in range(streams)]
    spks = [pkg.synth.synth_speaker(cfg, seed=rank * 4096 + i) for i in range(streams)]

    def one_pass(seed0):
        t0 = time.perf_counter()
        for i in range(streams):
            if not T.stream_prepare(i, texts[i], spks[i], max_len=B_FRAMES, temperature=TEMP, top_k=TOP_K,
                                    repetition_penalty=REP_PEN, seed=seed0 + i):
                raise RuntimeError("stream_prepare: " + T.get_error())
        if not T.prefill_pending():               # Talker prefill of all streams as lock-step passes of the batched kernel
            raise RuntimeError("prefill_pending: " + T.get_error())
        t1 = time.perf_counter()
        frames, dev_ms = 0, 0.0
        while True:
            out = T.step_frames(32)
            if out is None:
                raise RuntimeError("step_frames: " + T.get_error())
            codes, produced, active = out
            frames += int(produced.sum())
            dev_ms += T.last_loop_ms()
            if not active.any():
                break
        t2 = time.perf_counter()
        return frames, dev_ms, (t2 - t0) * 1e3, (t1 - t0) * 1e3

    one_pass(7)                                   # warm-up pass (also pages in the tiled weights)
    barrier()
    l0 = T.launch_count()
    frames, dev_ms, wall_ms, begin_ms = one_pass(99)
    barrier()
    launches = T.launch_count() - l0
    T.unload_model()
    red = pkg.sharding.JobCounters(dist, "cuda").reduce({"dev_ms": dev_ms, "wall_ms": wall_ms, "begin_ms": begin_ms},
                                                        {"frames": frames, "launches": launches})
    dev_ms, wall_ms, begin_ms, frames, launches = red["dev_ms"], red["wall_ms"], red["begin_ms"], red["frames"], int(red["launches"])
    # mean cached tokens per stream per frame: prefill (~10) + trailing grows one per frame
    n_kv_mean = 10 + (B_FRAMES - 1) / 2.0 + 1
    return {"frames": frames, "dev_ms": dev_ms, "wall_ms": wall_ms, "begin_ms": begin_ms, "launches": launches, "n_kv_mean": n_kv_mean}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-batched", action="store_true", help="skip the batched (configs[2]/[3]) legs")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    import qtts_pkg
    pkg = qtts_pkg.load()
    return run_reference(args, pkg) if args.impl == "reference" else run_ours(args, pkg)


if __name__ == "__main__":
    sys.exit(main())
