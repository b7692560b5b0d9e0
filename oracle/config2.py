"""ORACLE-side driver of BASELINE.json configs[1] ("TinyLlama-1.1B random-init fp32 CPU provider vs B200
provider, 64 threads x 8 turns").  TEST INFRASTRUCTURE ONLY: imported by tests/ and by bench.py's
`--config 2` leg (a CPU-vs-B200 measurement, i.e. a cpu_baseline leg); the product never imports it.

Per turn every thread posts a user message (48 printable-ASCII chars from
random.Random(1234 + i)); the assembled history (SQLite thread store, the reference's load -> convert ->
persist -> sanitise order, /root/reference/src/kafka/base.py:196-209) goes to the B200 provider — all
threads concurrently, continuous batching, KV resident between turns — and to the fp32 oracle
(oracle/cpu_provider.BatchedOracle).  Reports agreement per thread-turn and both sides' wall time.
"""
from __future__ import annotations

import asyncio
import random
import time
from typing import Dict


async def run_config2(gpu, oracle, store, n_threads: int = 64, n_turns: int = 8, max_tokens: int = 32,
                      user_len: int = 48) -> Dict:
    from kafka_b200.history import assemble_turn
    from kafka_b200.llm import Message
    mismatches, cached = [], []
    gpu_s = cpu_s = 0.0
    gen_gpu = gen_cpu = 0
    rnd = [random.Random(1234 + i) for i in range(n_threads)]
    keys = [f"bench-{i:04d}" for i in range(n_threads)]
    for turn in range(n_turns):
        hist = []
        for i, tid in enumerate(keys):
            text = "".join(chr(rnd[i].randrange(32, 127)) for _ in range(user_len))
            hist.append(await assemble_turn(store, tid, [{"role": "user", "content": text}]))
        t0 = time.perf_counter()
        got = await asyncio.gather(*[gpu.completion(h, temperature=0.0, max_tokens=max_tokens, thread_id=tid)
                                     for tid, h in zip(keys, hist)])
        gpu_s += time.perf_counter() - t0
        t0 = time.perf_counter()
        want = await asyncio.get_running_loop().run_in_executor(None, oracle.complete, keys, hist, max_tokens)
        cpu_s += time.perf_counter() - t0
        for tid, r, (text, reason) in zip(keys, got, want):
            gen_gpu += r.usage["completion_tokens"]
            if (r.content, r.finish_reason) != (text, reason):
                mismatches.append((tid, turn, r.content, text))
            cached.append((turn, r.usage["cached_tokens"], r.usage["prompt_tokens"]))
            await store.add_message(tid, Message(role="assistant", content=r.content))
        gen_cpu = oracle.tokens_generated
    bad = sorted({m[0] for m in mismatches})
    return {"threads": n_threads, "turns": n_turns, "max_tokens": max_tokens, "thread_turns": n_threads * n_turns,
            "mismatched_thread_turns": len(mismatches), "diverged_threads": len(bad),
            "agreement_fraction_of_threads": 1.0 - len(bad) / n_threads, "first_mismatches": mismatches[:3],
            "cached": cached, "b200_wall_s": gpu_s, "cpu_wall_s": cpu_s,
            "b200_tokens_per_s": gen_gpu / gpu_s, "cpu_tokens_per_s": gen_cpu / cpu_s,
            "generated_tokens": gen_gpu, "cpu_prefilled_tokens": oracle.tokens_prefilled}
