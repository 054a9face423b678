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
REF_FRAMES = (8, 136)   # reference arm: two generate() lengths per step; frame loop = their difference (128 frames)
B_FRAMES = 64           # frames per stream of the batched legs (configs[2] / configs[3])
B_TEXT = (32, 256)      # ragged text lengths of the batched legs
B_TOTAL = 512           # configs[3]: 512 concurrent streams sharded by utterance over the GPUs
METRIC, UNIT = "codec_frames_per_s", "frames/s"

# algorithmic bytes per frame-step, 0.6B, BF16 weights (SURVEY.md §8d / DESIGN.md):
#   talker 887,095,296 + code-pred layers 15 x 157,286,400 + lm_heads 62,914,560
WEIGHT_BYTES_PER_FRAME = 887_095_296 + 15 * 157_286_400 + 62_914_560
KV_BYTES_PER_TOKEN = 114_688      # 28 layers x 2 x 8 x 128 x 2 B, read per cached token per frame
FLOP_PER_FRAME_STREAM = 2 * (443_547_648 + 16 * 78_643_200 + 31_457_280)     # 3.467 GFLOP
ATTN_FLOP_PER_KEY = 229_376       # 2 x 2 x 16 heads x 128 x 28 layers per cached key
LOCKSTEP_GROUP = 256              # streams one launch of qtts_run_batched advances (qtts_engine.h: Engine::bgroup)


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


def measured_peak_tensor():
    """sustained dense BF16 TFLOP/s: the batched kernel is timed inside a long step"""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["bf16_tflops_sustained"]), "measured (MEASURED_PEAKS.json bf16_tflops_sustained)"
        except Exception:
            pass
    return 1400.0, "fallback (B200_PROFILING.md ~1.4 PFLOP/s sustained)"


# ---------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the UNMODIFIED reference on the host cores (oracle/_ref)
# ---------------------------------------------------------------------------------------------------
def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def time_reference(path: str, toks, spk, steps: int, warmup: int, threads: int, lengths=REF_FRAMES):
    """The UNMODIFIED reference's generate() (oracle/_ref, ggml CPU backend) timed at two lengths per step.
    The reference exposes no timer inside generate() (src/tts_transformer.cpp:2342-2574), so the frame loop is the
    difference: frames(long) - frames(short) over ms(long) - ms(short); the short call's time is prefill
    (text projection + 10-token pass) plus its few frames.  Returns a dict."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ref
    if not ref.available():
        raise RuntimeError("oracle/_ref/libqtts_ref.so is missing (make -f oracle/Makefile.ref)")
    R = ref.Reference(path, threads=threads)
    rows = []
    for i in range(warmup + steps):
        # greedy: the reference has no EOS-off switch and seeds its sampler from std::random_device
        # (src/tts_transformer.h:298), so a sampled run would stop at a random frame; the per-frame
        # arithmetic (Talker step + 16 code-predictor positions) is the same
        c0, ms0 = R.generate(toks, spk, max_len=lengths[0], temperature=0.0, top_k=TOP_K, rep_pen=REP_PEN)
        c1, ms1 = R.generate(toks, spk, max_len=lengths[1], temperature=0.0, top_k=TOP_K, rep_pen=REP_PEN)
        if i >= warmup:
            rows.append((len(c0), ms0, len(c1), ms1))
    R.close()
    f0 = sum(r[0] for r in rows); m0 = sum(r[1] for r in rows); f1 = sum(r[2] for r in rows); m1 = sum(r[3] for r in rows)
    loop_frames, loop_ms = f1 - f0, m1 - m0
    if loop_frames <= 0 or loop_ms <= 0:
        raise RuntimeError(f"reference stopped early (frames {f0}/{f1}): cannot separate the frame loop")
    ms_per_frame = loop_ms / loop_frames
    prefill_ms = (m0 - f0 * ms_per_frame) / len(rows)
    return {"loop_fps": loop_frames / (loop_ms / 1e3), "ms_per_frame": ms_per_frame, "prefill_ms": prefill_ms,
            "call_fps": f1 / (m1 / 1e3), "long_call_ms": m1 / len(rows), "frames_long": f1 // len(rows),
            "frames_short": f0 // len(rows), "threads": threads, "steps": len(rows)}


def run_reference(args, pkg):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    path = ensure_gguf(pkg, 0, lambda: None)
    cfg = pkg.synth.PRESETS[PRESET]
    toks = pkg.synth.synth_tokens(cfg, n_text=N_TEXT, seed=0)
    spk = pkg.synth.synth_speaker(cfg)
    cores = host_cores()
    r = time_reference(path, toks, spk, args.steps, args.warmup, cores)
    # the reference as shipped ignores -j and runs ggml's default of 4 threads (src/main.cpp:203-214,
    # ggml/include/ggml.h:232): BASELINE.md section 3 promises that row too
    r4 = time_reference(path, toks, spk, 1, 0, 4) if cores != 4 else r
    sample = (f"{r['steps']} x [generate({r['frames_short']} frames) + generate({r['frames_long']} frames)], greedy, same GGUF / text "
              f"({len(toks)} tokens) / speaker as the GPU arm, unmodified reference on the ggml CPU backend, {cores} threads; "
              f"value = frame loop only ({r['frames_long'] - r['frames_short']} frames by difference, {r['ms_per_frame']:.1f} ms/frame), "
              f"prefill {r['prefill_ms']:.0f} ms per call reported separately; e2e = whole generate({r['frames_long']}) calls incl. prefill")
    conf = workload_config(1)
    conf["workload"] = (f"BASELINE configs[1] workload on the reference's CPU path, BOUNDED sample: Qwen3-TTS-12Hz-0.6B random-init, batch 1, "
                        f"{r['frames_long']} frames per generate() (not {FRAMES}), GREEDY (temperature 0, not {TEMP}: the reference cannot "
                        f"suppress EOS and would stop at a random frame when sampling)")
    conf.update({"frames_per_step": r["frames_long"], "temperature": 0.0, "numerics": "ggml-cpu (F16 weights, F32 accumulate)",
                 "eos": "not suppressed (greedy run did not reach EOS)", "l2": "n/a (CPU)", "parallelism": f"{cores} host threads"})
    line = {"impl": "reference", "metric": METRIC, "value": r["loop_fps"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": r["long_call_ms"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f16 weights / f32 accumulate (ggml-cpu)", "data": "synthetic",
            "config": conf, "ms_per_frame": r["ms_per_frame"], "prefill_ms_per_call": r["prefill_ms"],
            "cpu_baseline": {"value": r["loop_fps"], "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample,
                             "default_4_threads": {"value": r4["loop_fps"], "ms_per_frame": r4["ms_per_frame"],
                                                   "prefill_ms_per_call": r4["prefill_ms"], "cores": 4}},
            "e2e": {"value": r["call_fps"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


NCU_DRAM_BYTES_PER_LAUNCH = 129986763000 + 33807360      # dram__bytes_read.sum + dram__bytes_write.sum, profiles/r02s_frame_ncu.md


def workload_config(world: int) -> dict:
    return {"workload": "BASELINE configs[1]: Qwen3-TTS-12Hz-0.6B random-init, batch 1 per GPU, temperature 0.9 top-k 50, "
                        f"{FRAMES} frames per utterance",
            "streams_per_gpu": 1, "frames_per_step": FRAMES, "text_tokens": N_TEXT + 8,
            "eos": "suppressed by the benchmark switch qtts_set_ignore_eos so every utterance runs all frames",
            "numerics": "fast: F16 weights as stored in the GGUF, F16-rounded activations / KV like ggml-cpu, F32 accumulate "
                        "(parity vs the reference per frame: profiles/r02_summary.md section 1)",
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
                "scaling": "weak", "vs_baseline": None, "dtype": "f16", "data": "synthetic",
                "config": workload_config(world),
                "ms_per_frame": loop_ms / (frames / world), "rtf_per_stream": (loop_ms / 1e3) / ((frames / world) / 12.5),
                "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": int(toks.nbytes + spk.nbytes),
                        "d2h_bytes_per_step": FRAMES * 16 * 4},
                "gpu_launches": launches,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             # dram__bytes_read.sum + dram__bytes_write.sum of one 40-frame launch, from the
                             # ncu --set full capture summarised in profiles/r02s_frame_ncu.md
                             "traffic": NCU_DRAM_BYTES_PER_LAUNCH, "traffic_source": "profiles/r02s_frame_ncu.md",
                             "frames_per_launch": 40, "algorithmic_bytes_per_launch": bytes_per_frame * 40,
                             "kernel": "qtts_run_program", "peak_source": peak_src,
                             "algorithmic_bytes_per_frame": bytes_per_frame},
                "clocks": clocks.summary()}
        line["gpu_launches"] = launches + sum(b["launches"] for b in batched)
        tpeak, tpeak_src = measured_peak_tensor()
        line["batched"] = []
        for b in batched:
            S = b["streams_per_gpu"]
            fps = b["frames"] / (b["dev_ms"] / 1e3)
            step_ms = b["dev_ms"] / (b["frames"] / (S * world))
            # One launch of qtts_run_batched advances one lock-step group of <= 256 streams and streams the weights once
            # for it; S streams on a GPU are ceil(S / 256) groups launched back to back (qtts_step_frames), so the
            # kernel's algorithmic bytes count the weights once PER GROUP.  The job-level minimum (weights once per
            # frame-step whatever S is) is reported next to it: their ratio is the excess the grouping costs.
            groups = (S + LOCKSTEP_GROUP - 1) // LOCKSTEP_GROUP
            kv_bytes = S * KV_BYTES_PER_TOKEN * (b["n_kv_mean"] + 1)
            bytes_launched = groups * WEIGHT_BYTES_PER_FRAME + kv_bytes
            bytes_min = WEIGHT_BYTES_PER_FRAME + kv_bytes
            ach = bytes_min / (step_ms / 1e3) / 1e9
            flops = S * (FLOP_PER_FRAME_STREAM + ATTN_FLOP_PER_KEY * b["n_kv_mean"])
            tf = flops / (step_ms / 1e3) / 1e12
            ridge = tpeak * 1e12 / (peak * 1e9)        # streams per GPU above which the weight GEMMs are FLOP-bound
            wl = ("configs[2]: 64 concurrent utterances on 1 GPU" if S == 64 and world == 1 else
                  f"configs[3]: {B_TOTAL} streams sharded over {world} GPU(s)" if S * world == B_TOTAL else
                  f"full lock-step group: {S} streams x {world} GPU(s)")
            entry = {"workload": wl, "streams_per_gpu": S, "streams_total": S * world, "value": round(fps, 1), "unit": UNIT,
                     "ms_per_frame_step": round(step_ms, 3), "e2e": round(b["frames"] / (b["wall_ms"] / 1e3), 1),
                     "roofline": {"bound": "hbm", "achieved": round(ach, 1), "peak": peak, "unit": "GB/s", "frac": round(ach / peak, 4),
                                  "groups": groups, "bytes_min": int(bytes_min), "bytes_launched": int(bytes_launched),
                                  "excess": round(bytes_launched / bytes_min, 3)}}
            if S >= ridge:
                entry["roofline_tensor"] = {"bound": "tensor", "achieved": round(tf, 2), "peak": tpeak, "unit": "TFLOP/s",
                                            "frac": round(tf / tpeak, 4), "ridge_streams": round(ridge, 1)}
            else:
                entry["tensor_tflops"] = round(tf, 2)
            line["batched"].append(entry)
            # every leg also goes out as its own short line BEFORE the headline line, so a truncated tail keeps it
            print(json.dumps({"batched_leg": entry, "n_gpus": world, "frames_per_stream": B_FRAMES, "text_tokens": list(B_TEXT),
                              "temperature": TEMP, "top_k": TOP_K,
                              "kernel": "qtts_run_batched" if S <= 64 else "qtts_run_batched_ext (same body: 4 activation buffers, 3-slot weight ring)",
                              "e2e_includes": "stream_prepare per stream (H2D, embedding assembly) + batched prefill + step_frames with codes D2H",
                              "peak_source": peak_src, "tensor_peak_source": tpeak_src}), flush=True)
        if world == 1 and not args.no_cpu_baseline:
            try:
                cores = host_cores()
                r = time_reference(path, toks, spk, 1, 0, cores, lengths=(8, 72))
                r4 = time_reference(path, toks, spk, 1, 0, 4, lengths=(4, 36)) if cores != 4 else r
                line["cpu_baseline"] = {"value": r["loop_fps"], "unit": UNIT, "cores": cores, "kind": "reference",
                                        "sample": f"frame loop of the unmodified reference (ggml CPU backend, {cores} threads) by difference of "
                                                  f"generate({r['frames_short']}) and generate({r['frames_long']}), greedy, same GGUF / text / speaker: "
                                                  f"{r['ms_per_frame']:.1f} ms/frame; prefill {r['prefill_ms']:.0f} ms per call on top",
                                        "prefill_ms_per_call": r["prefill_ms"],
                                        "default_4_threads": {"value": r4["loop_fps"], "ms_per_frame": r4["ms_per_frame"], "cores": 4,
                                                              "note": "the reference ignores -j: ggml default (ggml/include/ggml.h:232)"}}
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
