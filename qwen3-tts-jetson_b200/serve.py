"""Batching front door for the codec-frame engine.

The reference's `--serve` loop (src/main.cpp:108-163) reads `text<TAB>output.wav` lines from stdin,
synthesises ONE request at a time and answers `OK\\t<seconds>\\t<ms>\\t<file>` or `ERR\\t<message>`
on stdout.  The hot path alone has no tokenizer and no vocoder, so the requests here carry token ids
and the product is the codec-frame stream the vocoder consumes:

    request :  <id> TAB <comma-separated text token ids> TAB <output file> [TAB <max frames> [TAB <seed>]]
    answer  :  OK TAB <id> TAB <frames> TAB <ms> TAB <output file>        (or ERR TAB <id> TAB <message>)
    `quit` / `exit` end the loop after the requests in flight have finished.

What it adds to the reference loop is CONTINUOUS BATCHING: requests are admitted into free engine
slots between two `step_frames` calls, all live slots advance in lock step on the batched kernel,
and a slot is reused as soon as its utterance ends.  The per-stream contract of the reference's
frame callback is kept (src/tts_transformer.cpp:2517-2523, 2563-2570): `on_chunk(id, codes)` is
called with every full block of `chunk_frames` (40) frames while the utterance is generated and
with the remainder at its end - that is the hand-off point to a vocoder.  The output file holds the
frames as little-endian int32 [n_frames][n_codebooks].
"""
from __future__ import annotations

import os
import sys
import time
from collections import deque

import numpy as np


class _Live:
    __slots__ = ("req_id", "out", "parts", "n", "sent", "t0")

    def __init__(self, req_id, out):
        self.req_id, self.out = req_id, out
        self.parts, self.n, self.sent, self.t0 = [], 0, 0, time.perf_counter()


class BatchingServer:
    """`engine` needs the TTSTransformer stream interface (max_streams, stream_begin, step_frames,
    get_error, config['n_codebooks'])."""

    def __init__(self, engine, max_len: int = 4096, frames_per_call: int = 40, chunk_frames: int = 40,
                 on_chunk=None, base_seed: int | None = None, **sampling):
        """base_seed: None (default) seeds the per-request sampler seeds from os.urandom, the way the reference seeds
        its mt19937 from std::random_device (src/tts_transformer.h:298); an integer makes a server run reproducible.
        Either way every request gets its OWN seed (the device uniform is philox(seed, frame, codebook): one shared
        seed would make all concurrent requests draw the same uniform sequence)."""
        self.engine, self.max_len = engine, max_len
        self._seed_base = int.from_bytes(os.urandom(8), "little") if base_seed is None else int(base_seed)
        self._n_submitted = 0
        self.frames_per_call, self.chunk_frames = frames_per_call, chunk_frames
        self.on_chunk, self.sampling = on_chunk, sampling
        self.queue: deque = deque()
        self.live: dict[int, _Live] = {}
        self.free = list(range(engine.max_streams))[::-1]
        self.nb = engine.config["n_codebooks"]

    # -- requests ---------------------------------------------------------------------------------
    def submit(self, req_id: str, tokens, out: str | None = None, speaker=None, max_len: int | None = None,
               seed: int | None = None):
        """seed: None = the server's next per-request seed; an explicit value reproduces a request exactly."""
        if seed is None:
            z = (self._seed_base + 0x9E3779B97F4A7C15 * (self._n_submitted + 1)) & 0xFFFFFFFFFFFFFFFF      # splitmix64
            z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
            z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
            seed = (z ^ (z >> 31)) & 0x7FFFFFFFFFFFFFFF
        self._n_submitted += 1
        self.queue.append((req_id, np.asarray(tokens, np.int32), out, speaker, max_len or self.max_len, int(seed)))

    def idle(self) -> bool:
        return not self.queue and not self.live

    # -- one scheduling round ---------------------------------------------------------------------
    def pump(self) -> list[tuple]:
        """Admit, advance every live stream by up to frames_per_call frames, retire.  Returns the
        finished requests as ("OK", id, codes, ms, out) / ("ERR", id, message)."""
        done = []
        while self.queue and self.free:
            req_id, tokens, out, speaker, max_len, seed = self.queue.popleft()
            slot = self.free.pop()
            # two-phase begin: the Talker prefill of everything admitted in this round is ONE batched pass, run by
            # the next step_frames (engines without stream_prepare, e.g. the CPU test double, use stream_begin)
            begin = getattr(self.engine, "stream_prepare", None) or self.engine.stream_begin
            if not begin(slot, tokens, speaker, max_len=max_len, seed=seed, **self.sampling):
                done.append(("ERR", req_id, self.engine.get_error()))
                self.free.append(slot)
                continue
            self.live[slot] = _Live(req_id, out)
        if not self.live:
            return done
        res = self.engine.step_frames(self.frames_per_call)
        if res is None:
            msg = self.engine.get_error()
            for slot, lv in list(self.live.items()):
                done.append(("ERR", lv.req_id, msg))
                self.free.append(slot)
            self.live.clear()
            return done
        codes, produced, active = res
        for slot, lv in list(self.live.items()):
            k = int(produced[slot])
            if k:
                lv.parts.append(np.asarray(codes[slot, :k], np.int32).copy())
                lv.n += k
            finished = not bool(active[slot])
            self._chunks(lv, finished)
            if finished:
                all_codes = np.concatenate(lv.parts) if lv.parts else np.zeros((0, self.nb), np.int32)
                done.append(("OK", lv.req_id, all_codes, (time.perf_counter() - lv.t0) * 1e3, lv.out))
                del self.live[slot]
                self.free.append(slot)
        return done

    def _chunks(self, lv: _Live, finished: bool):
        if self.on_chunk is None:
            return
        flat = None
        while lv.n - lv.sent >= self.chunk_frames or (finished and lv.n > lv.sent):
            if flat is None:
                flat = np.concatenate(lv.parts)
            k = min(self.chunk_frames, lv.n - lv.sent)
            self.on_chunk(lv.req_id, flat[lv.sent:lv.sent + k])
            lv.sent += k

    def drain(self) -> list[tuple]:
        out = []
        while not self.idle():
            out.extend(self.pump())
        return out

    # -- the line protocol ------------------------------------------------------------------------
    @staticmethod
    def parse(line: str):
        """-> ("quit",) | ("req", id, tokens, out, max_len, seed) | ("err", id, message)"""
        line = line.rstrip("\r\n")
        if not line:
            return None
        if line in ("quit", "exit"):
            return ("quit",)
        f = line.split("\t")
        if len(f) < 2:
            return ("err", f[0], "expected <id> TAB <token ids> [TAB <output file> [TAB <max frames>]]")
        try:
            tokens = [int(t) for t in f[1].replace(" ", "").split(",") if t]
            max_len = int(f[3]) if len(f) > 3 and f[3] else None
            seed = int(f[4]) if len(f) > 4 and f[4] else None
        except ValueError:
            return ("err", f[0], "token ids, max frames and seed must be integers")
        return ("req", f[0], tokens, f[2] if len(f) > 2 and f[2] else None, max_len, seed)

    def serve(self, lines, out=sys.stdout, err=sys.stderr):
        """`lines`: an iterable of request lines (sys.stdin, or a list in tests).  Requests that are
        available together are batched; answers are written as their utterances finish."""
        print(f"Server ready ({self.engine.max_streams} slots). Send: id<TAB>tokens<TAB>output  (or 'quit')", file=err, flush=True)

        def answer(results):
            for r in results:
                if r[0] == "OK":
                    _, req_id, codes, ms, path = r
                    if path:
                        try:
                            np.ascontiguousarray(codes, "<i4").tofile(path)
                        except OSError as ex:
                            print(f"ERR\t{req_id}\tfailed to save {path}: {ex}", file=out, flush=True)
                            continue
                    print(f"OK\t{req_id}\t{len(codes)}\t{ms:.0f}\t{path or ''}", file=out, flush=True)
                else:
                    print(f"ERR\t{r[1]}\t{r[2]}", file=out, flush=True)

        for line in lines:
            p = self.parse(line)
            if p is None:
                continue
            if p[0] == "quit":
                break
            if p[0] == "err":
                print(f"ERR\t{p[1]}\t{p[2]}", file=out, flush=True)
                continue
            _, req_id, tokens, path, max_len, seed = p
            self.submit(req_id, tokens, path, max_len=max_len, seed=seed)
            # keep reading while slots are free; once they are all taken, make room
            while self.queue and not self.free:
                answer(self.pump())
        answer(self.drain())
        print("Server shutting down.", file=err, flush=True)
        return 0


def main(argv=None):
    import argparse
    from . import transformer
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("model", help="qwen3-tts GGUF file (F16/F32 tensors)")
    ap.add_argument("--streams", type=int, default=64)
    ap.add_argument("--max-tokens", type=int, default=4096, help="maximum frames per utterance (reference: --max-tokens)")
    ap.add_argument("--temperature", type=float, default=0.9)
    ap.add_argument("--top-k", type=int, default=50)
    ap.add_argument("--repetition-penalty", type=float, default=1.05)
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--seed", type=int, default=None, help="base of the per-request sampler seeds (default: os.urandom)")
    a = ap.parse_args(argv)
    # KV slots per stream: prefill (<= 11) + frames + 8 (src/tts_transformer.cpp:2383); the engine's cap is 4608
    T = transformer.TTSTransformer(device=a.device, max_streams=a.streams, max_ctx=min(a.max_tokens + 32, 4608))
    if not T.load_model(a.model):
        print("ERR\t-\t" + T.get_error(), flush=True)
        return 1
    srv = BatchingServer(T, max_len=a.max_tokens, base_seed=a.seed, temperature=a.temperature, top_k=a.top_k,
                         repetition_penalty=a.repetition_penalty)
    return srv.serve(sys.stdin)


if __name__ == "__main__":
    sys.exit(main())
