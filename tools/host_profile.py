"""Host-side cost of the serving loop, without a GPU: the bench's provider-level turn (512 threads x 2k-token
history, 128 new tokens) driven through B200LLMProvider + Engine + Scheduler with a stepper that does no
arithmetic and whose results arrive `--gpu-ms` after launch (a timer stands in for the CUDA stream).

    python tools/host_profile.py --threads 512 --gpu-ms 0          # pure host time per step
    python tools/host_profile.py --threads 512 --gpu-ms 28         # how much of it hides behind the GPU
    python tools/host_profile.py --profile engine|loop             # cProfile of the engine thread / event loop

Says where the gap between the device-timed and the end-to-end tokens/s goes (bench.py `value` vs `e2e`).
"""
from __future__ import annotations

import argparse
import asyncio
import cProfile
import os
import pstats
import random
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "kafka-llm-service_b200"))

from kafka_b200.engine import Engine, Stepper          # noqa: E402
from kafka_b200.llm import Message                     # noqa: E402
from kafka_b200.provider import B200LLMProvider        # noqa: E402
from kafka_b200.tokenizer import ByteTokenizer         # noqa: E402


class TimerStepper(Stepper):
    """No arithmetic: every launch 'finishes' gpu_ms after the previous one (one stream, in order)."""

    def __init__(self, n_blocks, max_seqs, max_tokens, max_pos, eos_id, gpu_ms, prefill_ms_per_ktok):
        self.n_blocks, self.max_seqs, self.max_tokens, self.max_pos, self.eos_id = (n_blocks, max_seqs, max_tokens,
                                                                                     max_pos, eos_id)
        self.gpu_s, self.prefill_s = gpu_ms / 1e3, prefill_ms_per_ktok / 1e6
        self.stream_free_at = 0.0
        self.h2d_bytes = self.d2h_bytes = 0

    def _enqueue(self, dur):
        now = time.perf_counter()
        self.stream_free_at = max(self.stream_free_at, now) + dur
        return self.stream_free_at

    def prefill_launch(self, items, parity):
        done = self._enqueue(self.prefill_s * sum(len(i.tokens) for i in items))
        return done, [65 if i.is_last_chunk else None for i in items]

    def decode_launch(self, reqs, parity):
        return self._enqueue(self.gpu_s if self.gpu_s else 0.0), [66] * len(reqs)

    def fetch(self, handle):
        done, ids = handle
        dt = done - time.perf_counter()
        if dt > 0:
            time.sleep(dt)
        return ids


TIMERS = {}


def timed(owner, name):
    """Wrap owner.name with a CPU-time accumulator (cheap, unlike cProfile, whose per-call cost distorts
    the many small functions of this path)."""
    fn = getattr(owner, name)
    key = f"{getattr(owner, '__name__', type(owner).__name__)}.{name}"
    TIMERS[key] = [0.0, 0]

    def wrapper(*a, **k):
        t0 = time.thread_time()          # CPU time of this thread: waiting for the GIL is not counted
        try:
            return fn(*a, **k)
        finally:
            rec = TIMERS[key]
            rec[0] += time.thread_time() - t0
            rec[1] += 1
    setattr(owner, name, wrapper)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--threads", type=int, default=512)
    ap.add_argument("--history", type=int, default=2048)
    ap.add_argument("--user-tokens", type=int, default=64)
    ap.add_argument("--gen-tokens", type=int, default=128)
    ap.add_argument("--gpu-ms", type=float, default=0.0, help="decode step duration the timer stands in for")
    ap.add_argument("--prefill-ms-per-ktok", type=float, default=0.0)
    ap.add_argument("--profile", choices=["engine", "loop"], default=None)
    ap.add_argument("--no-pipeline", action="store_true")
    ap.add_argument("--timers", action="store_true", help="wall-clock per function of the path (setup excluded)")
    args = ap.parse_args()
    if args.timers:
        from kafka_b200 import scheduler as S
        for owner, names in ((S.Scheduler, ["add", "admit", "next_prefill_batch", "next_decode_batch", "on_prefill_done",
                                            "on_decode_done", "on_prefill_launched", "on_decode_launched", "_release"]),
                             (S.PrefixIndex, ["add", "remove", "lookup", "donor", "chain_of"]),
                             (B200LLMProvider, ["render"]), (Engine, ["submit", "_drain"])):
            for n in names:
                timed(owner, n)

    tok = ByteTokenizer(128256)
    n_blocks = args.threads * ((args.history + args.user_tokens + args.gen_tokens + 200) // 64 + 2)
    stepper = TimerStepper(n_blocks, args.threads, 32768, 8192, tok.eot_id, args.gpu_ms, args.prefill_ms_per_ktok)
    prof = cProfile.Profile() if args.profile else None
    if args.profile == "engine":
        run0 = Engine._run

        def run_profiled(self):
            prof.enable()
            try:
                run0(self)
            finally:
                prof.disable()
        Engine._run = run_profiled
    engine = Engine(stepper, max_prefill_tokens=stepper.max_tokens, pipeline=not args.no_pipeline)
    provider = B200LLMProvider(engine, tok, model="llama-3-8b", default_temperature=0.0)

    def text(rnd, n):
        return "".join(chr(rnd.randrange(32, 127)) for _ in range(n))

    def make_history(i):
        rnd = random.Random(100003 + i)
        msgs = [Message(role="system", content="You are a helpful assistant.")]
        while len(provider.render(msgs, None)) < args.history - 700:
            msgs.append(Message(role="user", content=text(rnd, 180)))
            msgs.append(Message(role="assistant", content=text(rnd, 180)))
        short = args.history - len(provider.render(msgs + [Message(role="user", content="")], None))
        msgs.append(Message(role="user", content=text(rnd, max(1, short))))
        return msgs

    async def run():
        hist = [make_history(i) for i in range(args.threads)]
        first = await asyncio.gather(*[provider.completion(h, temperature=0.0, max_tokens=1, thread_id=f"t{i}")
                                       for i, h in enumerate(hist)])
        turns = [h + [r.to_message(), Message(role="user", content=text(random.Random(i), args.user_tokens))]
                 for i, (h, r) in enumerate(zip(hist, first))]
        t_first = []
        for rec in TIMERS.values():
            rec[0], rec[1] = 0.0, 0

        async def one(i, msgs):
            t0 = time.perf_counter()
            n, ttft = 0, None
            async for ch in provider.stream_completion(msgs, temperature=0.0, max_tokens=args.gen_tokens,
                                                       thread_id=f"t{i}"):
                if ch.content:
                    if ttft is None:
                        ttft = time.perf_counter() - t0
                    n += len(ch.content)
            t_first.append(ttft)
            return n
        if args.profile == "loop":
            prof.enable()
        t0 = time.perf_counter()
        ns = await asyncio.gather(*[one(i, m) for i, m in enumerate(turns)])
        wall = time.perf_counter() - t0
        if args.profile == "loop":
            prof.disable()
        return wall, sum(ns), sorted(t_first)

    wall, n, ttfts = asyncio.run(run())
    steps = args.gen_tokens
    ideal = steps * args.gpu_ms / 1e3
    print(f"threads {args.threads}  tokens {n}  wall {wall * 1e3:.0f} ms  = {wall / steps * 1e3:.2f} ms per step "
          f"(timer GPU {args.gpu_ms} ms/step -> ideal {ideal * 1e3:.0f} ms)  ttft p50 {ttfts[len(ttfts) // 2] * 1e3:.0f} ms")
    print("engine timing:", {k: round(v, 3) if isinstance(v, float) else v for k, v in engine.timing.items()})
    engine.shutdown()
    for k, (t, c) in sorted(TIMERS.items(), key=lambda kv: -kv[1][0]):
        print(f"  {k:36s} {t * 1e3:8.1f} ms  {c:7d} calls")
    if prof is not None:
        pstats.Stats(prof).sort_stats("cumulative").print_stats(28)


if __name__ == "__main__":
    main()
