#!/usr/bin/env python
"""bench.py — decode tokens/s of the B200 completion path (BASELINE.json metric).

Workload (BASELINE.json configs[2], thread-sharded): Llama-3-8B bf16, per GPU `--threads` concurrent
synthetic threads (default 512 = 4096 / 8, the share one GPU owns at 8 GPUs and the most that is
resident in 180 GB), each with a 2 048-token history resident in paged KV, a 64-token user message
prefilled incrementally, then greedy decode.  A *step* is one decode step over all threads of the
rank (B tokens).  Weak scaling: per-GPU work is fixed, ranks share nothing (no collective on the
data path), value = N * B * K / max-over-ranks time.

Legs (all on the same resident state, CUDA-event timed on the launching stream):
  value      device-resident decode loop (token feedback on the device)
  roofline   the same loop with CUDA events between kernels -> per-kernel time; the dominant kernel
             (paged attention decode) reported as achieved HBM GB/s against MEASURED_PEAKS.json
  e2e        the same step driven through the C ABI with HOST buffers: pinned H2D of the step's
             input tokens and D2H of the sampled tokens inside the timed region
  cpu_baseline (rank 0, N=1 only)  the fp32 oracle on the host cores on a bounded sample
`--impl reference` times the reference arm: the CPU oracle port of the path (the reference's own
implementation is a remote service, unavailable offline — BASELINE.md §1).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT / "kafka-llm-service_b200"))
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

METRIC = "decode tokens/s (whole box)"
UNIT = "tokens/s"


# ------------------------------------------------------------------------------------------------
def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=32)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="llama-3-8b")
    ap.add_argument("--threads", type=int, default=512, help="concurrent threads per GPU")
    ap.add_argument("--history", type=int, default=2048)
    ap.add_argument("--user-tokens", type=int, default=64)
    ap.add_argument("--cpu-threads", type=int, default=0,
                    help="threads in the CPU sample (0 = 64 when the host has the RAM for it, else 8)")
    ap.add_argument("--cpu-steps", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-reference-server", action="store_true",
                    help="reference arm: skip the legs that run the reference's own host code (baseline/_ref)")
    ap.add_argument("--no-profile", action="store_true")
    ap.add_argument("--no-provider", action="store_true", help="skip the provider-level (plugin API) turn")
    ap.add_argument("--gen-tokens", type=int, default=128, help="new tokens per turn in the provider-level leg")
    ap.add_argument("--no-ttft-sweep", action="store_true")
    ap.add_argument("--agent-tool-rounds", type=int, default=0,
                    help="provider-level leg as BASELINE configs[3]: every thread's turn is an AGENT turn with this "
                         "many scripted tool rounds (64-token tool results) before the final text round")
    ap.add_argument("--no-pipeline", action="store_true",
                    help="serving loop reads every step back before enqueueing the next (A/B of the pipelined loop)")
    ap.add_argument("--tp-compare", action="store_true",
                    help="tensor-parallel run: also time the other all-reduce implementations on the same state")
    ap.add_argument("--tp-allreduce", default=None, choices=["peers", "peers-f16", "peers-fused", "peers-p2p", "nccl"],
                    help="all-reduce of the tensor-parallel run (default: own peer-memory kernel, NCCL if unavailable)")
    ap.add_argument("--config", type=int, default=3, choices=[2, 3],
                    help="3 (default): the headline workload. 2: BASELINE configs[1] whole — TinyLlama-1.1B shapes, 64 "
                         "threads x 8 turns x 32 tokens, B200 provider vs the fp32 CPU oracle, agreement + both tok/s")
    ap.add_argument("--tp", action="store_true",
                    help="tensor-parallel: ONE model sharded over all ranks (config 5: --model llama-3-70b --history 8192)")
    return ap.parse_args()


_JSON_FD = None   # the real stdout, when fd 1 has been pointed at stderr (see main)


def emit(line) -> None:
    """Print THE one JSON line on the real stdout."""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d["hbm_gbs"]), float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "measured"
    return 6650.0, 1590.0, "fallback"  # B200_PROFILING.md fallback


def ncu_dram_traffic(kernel_substr: str = "attn_decode_kernel<128, 4>"):
    """DRAM bytes per launch of the dominant kernel from the committed ncu capture
    (profiles/r02_ncu_attn_decode_dram.csv: `ncu --replay-mode application -k regex:attn_decode -c 2 --metrics
    dram__bytes_read.sum,dram__bytes_write.sum` on `bench.py --threads 512 --steps 2 --warmup 3 --no-e2e
    --no-provider --no-cpu-baseline --no-profile` — application replay, because kernel replay would have to save
    and restore the 140 GB pool). The captured launches are the first decode step's: 512 threads, kv_len =
    history + user tokens + 1. -> (bytes per launch, launches averaged) or None."""
    import csv
    path = ROOT / "profiles" / "r02_ncu_attn_decode_dram.csv"
    if not path.exists():
        return None
    per = {}
    with open(path, newline="") as f:
        rows = [r for r in csv.reader(f) if len(r) >= 15 and r[0].isdigit()]
    for r in rows:
        if kernel_substr in r[4] and r[12] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            per[r[0]] = per.get(r[0], 0.0) + float(r[14])
    if not per:
        return None
    return sum(per.values()) / len(per), len(per)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index: int):
        self.rows = []
        self.proc = None
        try:
            uuid = str(torch.cuda.get_device_properties(device_index).uuid)
            sel = uuid if uuid.startswith("GPU-") else "GPU-" + uuid
        except Exception:
            sel = str(device_index)
        self.cmd = ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200", "-i", sel]

    def start(self):
        try:
            self.proc = subprocess.Popen(self.cmd, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------
def pick_cpu_threads() -> int:
    """All host cores, unless the machine punishes that: on an over-subscribed VM an OpenMP team as wide as
    cpu_count() can spend milliseconds in the barrier of every small op (seen in the build container: 6.7 ms
    per RMSNorm-sized op at 8 threads, 50 us at 4). Halve until a small op costs what it should."""
    n = os.cpu_count() or 1
    x = torch.randn(64, 2048)
    while n > 1:
        torch.set_num_threads(n)
        for _ in range(5):
            x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + 1e-5)
        t0 = time.perf_counter()
        for _ in range(50):
            x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + 1e-5)
        if (time.perf_counter() - t0) / 50 < 5e-4:
            break
        n //= 2
    torch.set_num_threads(n)
    return n


def mem_available_gb() -> float:
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable:"):
                return int(line.split()[1]) / 1e6
    except OSError:
        pass
    return 0.0


def cpu_sample_size(cfg, ctx: int, want: int):
    """(threads, alias_weights) of the CPU sample. want <= 0 = automatic: 64 threads with every layer's fp32
    weights in memory of its own when the host has the RAM for it (Llama-3-8B: 32 GB of weights + 36 GB of fp32
    K/V), else 8 threads with one layer's tensors shared by all layers (872 MB, still far beyond the caches)."""
    w_gb = 4 * (cfg.matmul_params() + 2 * cfg.vocab * cfg.hidden) / 1e9
    kv_gb = lambda n: n * (ctx + 16) * cfg.kv_bytes_per_token * 2 / 1e9      # fp32 K/V
    avail = mem_available_gb()
    if want <= 0:
        want = 64 if avail > 1.3 * (w_gb + kv_gb(64)) + 8 else 8
    alias = avail < 1.3 * (w_gb + kv_gb(want)) + 8
    return want, alias


def cpu_baseline(cfg, n_threads: int, ctx: int, steps: int, warmup: int = 1, alias: bool = True):
    """The fp32 oracle (oracle/llama_fp32.py) decoding `n_threads` threads at context `ctx` on the host
    cores. Values are irrelevant for timing, shapes, FLOPs and DRAM traffic are the workload's. With
    `alias` (small hosts) ONE layer's fp32 weight tensors (872 MB for Llama-3-8B, far larger than the
    last-level cache) are shared by all layers and embed / lm_head share one table; without it every layer
    and both tables own their memory (copies of one draw)."""
    from oracle.llama_fp32 import OracleLlama
    cores = pick_cpu_threads()
    g = torch.Generator().manual_seed(0)
    H, F = cfg.hidden, cfg.ffn

    def rn(*shape, std=0.02):
        return torch.empty(*shape).normal_(0, std, generator=g)

    table = rn(cfg.vocab, H, std=1.0 / H ** 0.5)
    shared = {"attn_norm": torch.ones(H), "mlp_norm": torch.ones(H), "wq": rn(cfg.q_dim, H),
              "wk": rn(cfg.kv_dim, H), "wv": rn(cfg.kv_dim, H), "wo": rn(H, cfg.q_dim), "w_gate": rn(F, H),
              "w_up": rn(F, H), "w_down": rn(H, F)}
    w = {"embed": table, "lm_head": table if alias else table.clone(), "final_norm": torch.ones(H)}
    for i in range(cfg.n_layers):
        for k, v in shared.items():
            w[f"l{i}.{k}"] = v if (alias or i == 0) else v.clone()
    oracle = OracleLlama(cfg, w)
    caches = []
    for _ in range(n_threads):
        c = oracle.new_cache()
        for i in range(cfg.n_layers):
            # distinct memory per (thread, layer) so the cache streams from DRAM; values are irrelevant
            c._k[i] = torch.full((ctx + steps + warmup + 8, cfg.n_kv_heads, cfg.head_dim), 0.01)
            c._v[i] = torch.full((ctx + steps + warmup + 8, cfg.n_kv_heads, cfg.head_dim), 0.02)
            c._len[i] = ctx
        caches.append(c)
    toks = torch.randint(0, 256, (n_threads,), generator=g)
    for _ in range(warmup):
        toks = oracle.decode_batch(toks, caches).argmax(-1)
    t0 = time.perf_counter()
    for _ in range(steps):
        toks = oracle.decode_batch(toks, caches).argmax(-1)
    dt = time.perf_counter() - t0
    return {"value": n_threads * steps / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{n_threads} threads x {steps} decode steps at context {ctx}, fp32 torch eager oracle, "
                      f"{cfg.name} shapes, " + (f"one layer's weights shared by all {cfg.n_layers} layers"
                                                if alias else "every layer's weights in memory of its own"),
            "threads": n_threads, "weights_aliased": alias, "host_cpus": os.cpu_count(),
            "ms_per_step": dt / steps * 1e3}


# ------------------------------------------------------------------------------------------------
def provider_turn(core, cfg, n_threads: int, history: int, user_tokens: int, gen_tokens: int, seed: int,
                  pipeline: bool = True, tool_rounds: int = 0):
    """The call a user makes: B200LLMProvider.stream_completion for `n_threads` concurrent threads.

    Untimed set-up: every thread's ~`history`-token conversation is prefilled once through the provider
    (a first completion of 1 token), which leaves its KV resident. Timed: one turn of every thread —
    the conversation re-sent in full plus a new `user_tokens`-char user message, `gen_tokens` new tokens,
    streamed chunk by chunk. The engine finds the resident prefix, prefills only the new messages and
    decodes with continuous batching. Wall-clock around asyncio.gather; TTFT = first content chunk."""
    import asyncio
    import random
    import statistics as st
    from kafka_b200.engine import Engine
    from kafka_b200.gpu_stepper import GpuStepper
    from kafka_b200.llm import Message
    from kafka_b200.provider import B200LLMProvider
    from kafka_b200.tokenizer import ByteTokenizer

    tok = ByteTokenizer(cfg.vocab)
    stepper = GpuStepper(core, tok, generable_only=True)
    stepper.banned_token = tok.eot_id                   # EOS masked: every turn generates gen_tokens tokens
    stepper.close = lambda: None                        # the core outlives this leg
    engine = Engine(stepper, max_prefill_tokens=core.max_tokens, pipeline=pipeline)
    tools = tool_schemas = None
    if tool_rounds > 0:
        from kafka_b200.tools import LocalToolProvider, Tool

        async def lookup(key: str = ""):
            return "v" * 64                                   # a 64-token tool result (BASELINE configs[3])
        tools = LocalToolProvider([Tool("lookup", "look something up",
                                        {"type": "object", "properties": {"key": {"type": "string"}}}, lookup)])
        from kafka_b200.agent import StreamingAgent
        StreamingAgent(None, tools)                           # registers the agent's `idle` tool, as every turn will
        tool_schemas = asyncio.run(tools.get_tools())         # the schemas are part of every prompt of these threads
    provider = B200LLMProvider(engine, tok, model=cfg.name, default_temperature=0.0, tool_provider=tools)

    def text(rnd, n):
        return "".join(chr(rnd.randrange(32, 127)) for _ in range(n))

    def make_history(i):
        rnd = random.Random(seed * 100003 + i)
        msgs = [Message(role="system", content="You are a helpful assistant.")]
        while len(provider.render(msgs, tool_schemas)) < history - 700:      # ~200 tokens per rendered message
            msgs.append(Message(role="user", content=text(rnd, 180)))
            msgs.append(Message(role="assistant", content=text(rnd, 180)))
        # the last user message is sized so that the rendered prompt is exactly `history` tokens
        short = history - len(provider.render(msgs + [Message(role="user", content="")], tool_schemas))
        msgs.append(Message(role="user", content=text(rnd, max(1, short))))
        return msgs

    async def run():
        hist = [make_history(i) for i in range(n_threads)]
        n_hist = [len(provider.render(h, tool_schemas)) for h in hist]
        t0 = time.perf_counter()
        first = await asyncio.gather(*[provider.completion(h, temperature=0.0, max_tokens=1, thread_id=f"bench-{i:04d}")
                                       for i, h in enumerate(hist)])
        setup_s = time.perf_counter() - t0
        turns = []
        for i, (h, r) in enumerate(zip(hist, first)):
            rnd = random.Random(seed * 7 + i)
            turns.append(h + [r.to_message(), Message(role="user", content=text(rnd, user_tokens))])
        stats0 = engine.stats()
        h2d0, d2h0, l0 = stepper.h2d_bytes, stepper.d2h_bytes, __import__("kafka_b200.cabi", fromlist=["x"]).launch_count()

        async def one(i, msgs):
            t_call = time.perf_counter()
            ttft, n = None, 0
            if tool_rounds <= 0:
                async for ch in provider.stream_completion(msgs, temperature=0.0, max_tokens=gen_tokens,
                                                           thread_id=f"bench-{i:04d}"):
                    if ch.content:
                        if ttft is None:
                            ttft = time.perf_counter() - t_call
                        n += len(ch.content)
                return ttft, n, time.perf_counter() - t_call
            # agent turn: tool_rounds x (16 tokens of text + one scripted tool call + its result), then text
            from kafka_b200.agent import StreamingAgent
            from kafka_b200.provider import B200LLMProvider as _P, ScriptedToolCallProvider
            inner = _P(engine, tok, model=cfg.name, default_temperature=0.0, tool_provider=tools,
                       thread_id=f"bench-{i:04d}")
            agent = StreamingAgent(ScriptedToolCallProvider(inner, "lookup", {"key": "k"}, tool_rounds=tool_rounds,
                                                            think_tokens=16), tools, system_prompt=None)
            async for ev in agent.run(msgs, model=cfg.name, temperature=0.0, max_tokens=gen_tokens,
                                      thread_id=f"bench-{i:04d}"):
                ch = ev.get("choices")
                if ch and ch[0].get("delta", {}).get("content"):
                    if ttft is None:
                        ttft = time.perf_counter() - t_call
                    n += len(ch[0]["delta"]["content"])
            return ttft, n, time.perf_counter() - t_call

        t0 = time.perf_counter()
        res = await asyncio.gather(*[one(i, m) for i, m in enumerate(turns)])
        wall = time.perf_counter() - t0
        s1 = engine.stats()
        steps = max(1, s1["decode_steps"] - stats0["decode_steps"])
        new_tok = sum(n for _, n, _ in res)
        cached = s1["cached_tokens"] - stats0["cached_tokens"]
        prompt = s1["prompt_tokens"] - stats0["prompt_tokens"]
        ttfts = sorted(t for t, _, _ in res if t is not None)
        from kafka_b200 import cabi as _c
        return {
            "value": new_tok / wall, "unit": UNIT, "wall_s": wall, "new_tokens": new_tok,
            "h2d_bytes_per_step": (stepper.h2d_bytes - h2d0) / steps, "d2h_bytes_per_step": (stepper.d2h_bytes - d2h0) / steps,
            "decode_steps": steps, "gpu_launches": _c.launch_count() - l0,
            "ttft_p50_ms": 1e3 * st.median(ttfts), "ttft_p99_ms": 1e3 * ttfts[int(0.99 * (len(ttfts) - 1))],
            "history_tokens_mean": sum(n_hist) / len(n_hist), "cached_fraction_of_prompt": cached / max(1, prompt),
            "prefilled_new_tokens": prompt - cached, "setup_prefill_s": setup_s,
            "engine_prefill_s": s1["prefill_s"] - stats0["prefill_s"], "engine_decode_s": s1["decode_s"] - stats0["decode_s"],
            "provider_calls": s1["requests"] - stats0["requests"], "agent_tool_rounds": tool_rounds,
            "api": ("StreamingAgent.run over " if tool_rounds > 0 else "") +
                   "B200LLMProvider.stream_completion (LLMProvider plugin API), one coroutine per thread, "
                   "pinned H2D of step inputs (ctx lens, block tables on membership change) and D2H of sampled "
                   "ids every step",
        }

    try:
        return asyncio.run(run())
    finally:
        engine.shutdown()


# ------------------------------------------------------------------------------------------------
def http_turn(core, cfg, n_threads: int, history: int, user_tokens: int, gen_tokens: int):
    """HTTP-level TTFT: the repo's server (same routes as the reference's server.py) under
    uvicorn on 127.0.0.1, `n_threads` concurrent POST /v1/threads/{id}/chat/completions with stream=true on
    threads whose history (SQLite + KV) was built by an earlier request. TTFT = request sent -> first SSE
    `data:` line carrying delta.content."""
    import asyncio
    import random
    import socket
    import statistics as st
    import tempfile
    import threading as th
    import aiohttp
    import uvicorn
    from kafka_b200.db import ThreadStore
    from kafka_b200.engine import Engine
    from kafka_b200.gpu_stepper import GpuStepper
    from kafka_b200.provider import B200LLMProvider
    from kafka_b200.server import create_app
    from kafka_b200.tokenizer import ByteTokenizer

    tok = ByteTokenizer(cfg.vocab)
    stepper = GpuStepper(core, tok, generable_only=True)
    stepper.banned_token = tok.eot_id
    stepper.close = lambda: None
    engine = Engine(stepper, max_prefill_tokens=core.max_tokens)
    tmp = tempfile.mkdtemp(prefix="kb-http-")

    def factory(model, tool_provider=None, thread_id=None, **kw):
        return B200LLMProvider(engine, tok, model=cfg.name, tool_provider=tool_provider, thread_id=thread_id,
                               default_temperature=0.0)
    app = create_app(factory, ThreadStore(os.path.join(tmp, "threads.db")), models=[cfg.name])
    sock = socket.socket()
    sock.bind(("127.0.0.1", 0))
    port = sock.getsockname()[1]
    sock.close()
    server = uvicorn.Server(uvicorn.Config(app, host="127.0.0.1", port=port, log_level="error"))
    t = th.Thread(target=server.run, daemon=True)
    t.start()
    for _ in range(200):
        if server.started:
            break
        time.sleep(0.05)

    def text(rnd, n):
        return "".join(chr(rnd.randrange(97, 123)) for _ in range(n))

    async def post(session, tid, content, max_tokens):
        t0 = time.perf_counter()
        ttft, n = None, 0
        body = {"model": cfg.name, "stream": True, "temperature": 0, "max_tokens": max_tokens,
                "messages": [{"role": "user", "content": content}]}
        async with session.post(f"http://127.0.0.1:{port}/v1/threads/{tid}/chat/completions", json=body) as r:
            async for raw in r.content:
                line = raw.decode().strip()
                if not line.startswith("data: ") or line == "data: [DONE]":
                    continue
                ev = json.loads(line[6:])
                ch = ev.get("choices") or []
                if ch and (ch[0].get("delta") or {}).get("content"):
                    if ttft is None:
                        ttft = time.perf_counter() - t0
                    n += len(ch[0]["delta"]["content"])
        return ttft, n, time.perf_counter() - t0

    async def run():
        async with aiohttp.ClientSession(timeout=aiohttp.ClientTimeout(total=600)) as session:
            rnd = random.Random(7)
            await asyncio.gather(*[post(session, f"http-{i:04d}", text(rnd, history - 64), 1) for i in range(n_threads)])
            t0 = time.perf_counter()
            res = await asyncio.gather(*[post(session, f"http-{i:04d}", text(rnd, user_tokens), gen_tokens)
                                         for i in range(n_threads)])
            wall = time.perf_counter() - t0
        ttfts = sorted(x[0] for x in res if x[0] is not None)
        return {"threads": n_threads, "history_tokens": history, "new_tokens_per_thread": gen_tokens,
                "ttft_p50_ms": 1e3 * st.median(ttfts), "ttft_max_ms": 1e3 * ttfts[-1],
                "turn_p50_ms": 1e3 * st.median(x[2] for x in res), "tokens_per_s": sum(x[1] for x in res) / wall,
                "api": "POST /v1/threads/{id}/chat/completions stream=true over uvicorn on 127.0.0.1 (aiohttp client)"}

    try:
        return asyncio.run(run())
    finally:
        server.should_exit = True
        t.join(timeout=10)
        engine.shutdown()


# ------------------------------------------------------------------------------------------------
def _load_reference():
    """The reference's own host code: /root/reference where it is mounted (the build container), else its
    unmodified copy under baseline/_ref (made by __graft_entry__.build(); the GPU box only has that)."""
    sys.path.insert(0, str(ROOT / "tests"))
    import ref_harness
    ref_harness.ensure_copy()
    return ref_harness if ref_harness.load() else None


def reference_config1(n_threads: int = 64, n_turns: int = 8):
    """BASELINE configs[0]: thread replay on CPU over SQLite with an echo provider — the host plumbing around
    the path (history load, assembly, sanitising, persistence, agent loop), timed through the REFERENCE's
    LocalDBClient + KafkaAgent.run_with_thread + Agent.run (/root/reference/src/db/local.py:87-177,
    src/kafka/base.py:171-310, src/agents/base.py:160) and through this repo's ThreadStore + ThreadRunner +
    StreamingAgent, same messages, fresh database files, one turn after the other."""
    import asyncio
    import contextlib
    import random
    import tempfile
    ref = _load_reference()
    from kafka_b200.agent import StreamingAgent
    from kafka_b200.db import ThreadStore
    from kafka_b200.orchestrator import ThreadRunner
    from kafka_b200.provider import EchoLLMProvider
    from kafka_b200.tools import LocalToolProvider
    tmp = tempfile.mkdtemp(prefix="kb-config1-")

    def text(i, turn):
        rnd = random.Random(1234 + i * 100 + turn)
        return "".join(chr(rnd.randrange(32, 127)) for _ in range(48))

    async def ours():
        store = ThreadStore(os.path.join(tmp, "ours.db"))
        t0 = time.perf_counter()
        for turn in range(n_turns):
            for i in range(n_threads):
                tools = LocalToolProvider([])
                runner = ThreadRunner(StreamingAgent(EchoLLMProvider("echo", tools), tools), store, f"bench-{i:04d}")
                async for _ in runner.run([{"role": "user", "content": text(i, turn)}], model="echo"):
                    pass
        dt = time.perf_counter() - t0
        n = len(await store.get_thread_messages(f"bench-{n_threads - 1:04d}"))
        return dt, n

    async def theirs():
        from src.agents.base import Agent
        from src.db.local import LocalDBClient
        from src.kafka.base import KafkaAgent
        from src.llm.base import LLMProvider
        from src.llm.types import CompletionResponse, Message, StreamChunk
        from src.tools.agent import AgentToolProvider

        class Echo(LLMProvider):
            async def stream_completion(self, messages, *, temperature=None, max_tokens=None, stop=None, **kw):
                last = next((m.content for m in reversed(messages) if m.role == "user" and m.content), "")
                yield StreamChunk(role="assistant")
                if last:
                    yield StreamChunk(content=last)
                yield StreamChunk(finish_reason="stop")

            async def completion(self, messages, **kw):
                return CompletionResponse(content="", finish_reason="stop")

        class Kafka(KafkaAgent):
            """Smallest concrete KafkaAgent: what KafkaV1Provider.initialize sets up (src/kafka/v1.py:177-194)
            minus sandboxes, prompts and remote tools."""
            def __init__(self, thread_id, db):
                super().__init__(thread_id)
                self._db = db

            async def initialize(self):
                self._db_client = self._db
                self._tool_provider = AgentToolProvider()
                await self._tool_provider.connect()
                self._agent = Agent(llm_provider=Echo(), tool_provider=self._tool_provider)
                self._initialized = True

            async def cleanup(self):
                await self._tool_provider.disconnect()

            async def get_tools(self):
                return await self._tool_provider.get_tools()

            async def run(self, messages, model, temperature=0.7, max_tokens=None):
                async for ev in self._agent.run(messages, model=model, temperature=temperature, max_tokens=max_tokens):
                    yield ev

        db = LocalDBClient(os.path.join(tmp, "ref.db"))
        if hasattr(db, "initialize"):
            await db.initialize()
        t0 = time.perf_counter()
        for turn in range(n_turns):
            for i in range(n_threads):
                async with Kafka(f"bench-{i:04d}", db) as k:      # a new agent per request, like server.py:237-245
                    async for _ in k.run_with_thread([Message(role="user", content=text(i, turn))], model="echo"):
                        pass
        dt = time.perf_counter() - t0
        n = len(await db.get_thread_messages(f"bench-{n_threads - 1:04d}"))
        return dt, n

    out = {"threads": n_threads, "turns": n_turns, "provider": "echo",
           "what": "history load + assembly + persistence + agent loop per thread-turn, SQLite, no model"}
    dt, n = asyncio.run(ours())
    out["repo"] = {"turns_per_s": n_threads * n_turns / dt, "wall_s": dt, "messages_in_last_thread": n,
                   "api": "ThreadStore + ThreadRunner + StreamingAgent"}
    if ref is None:
        out["reference"] = {"unavailable": "neither /root/reference nor baseline/_ref holds the reference tree"}
        return out
    with contextlib.redirect_stdout(sys.stderr):             # the reference's agent prints progress lines
        dt, n = asyncio.run(theirs())
    out["reference"] = {"turns_per_s": n_threads * n_turns / dt, "wall_s": dt, "messages_in_last_thread": n,
                        "api": "LocalDBClient + KafkaAgent.run_with_thread + Agent.run (unmodified reference code, "
                               "aiosqlite shim over sqlite3)", "tree": str(ref.REF)}
    return out


def reference_oracle_in_agent(model: str = "tinyllama-1.1b", n_threads: int = 4, max_tokens: int = 16):
    """north_star's CPU baseline as worded: the fp32 CPU provider (oracle/cpu_provider.CPUOracleProvider)
    running INSIDE the reference's own KafkaAgent.run_with_thread / Agent.run over the reference's
    LocalDBClient, one turn per thread, config-2 shapes and message sizes. Bounded: a few threads."""
    import asyncio
    import contextlib
    import random
    import tempfile
    ref = _load_reference()
    if ref is None:
        return {"unavailable": "neither /root/reference nor baseline/_ref holds the reference tree"}
    from kafka_b200.llm import register_with_reference
    from kafka_b200.model_config import get_config
    from kafka_b200.weights import generate_weights
    from oracle.cpu_provider import CPUOracleProvider
    from src.agents.base import Agent
    from src.db.local import LocalDBClient
    from src.kafka.base import KafkaAgent
    from src.llm.types import Message
    from src.tools.agent import AgentToolProvider
    register_with_reference()
    cores = pick_cpu_threads()
    cfg = get_config(model)
    provider = CPUOracleProvider(cfg, generate_weights(cfg, seed=0, init="margin"), default_max_tokens=max_tokens)

    class Kafka(KafkaAgent):
        def __init__(self, thread_id, db):
            super().__init__(thread_id)
            self._db = db

        async def initialize(self):
            self._db_client = self._db
            self._tool_provider = AgentToolProvider()
            await self._tool_provider.connect()
            self._agent = Agent(llm_provider=provider, tool_provider=self._tool_provider)
            self._initialized = True

        async def cleanup(self):
            await self._tool_provider.disconnect()

        async def get_tools(self):
            return await self._tool_provider.get_tools()

        async def run(self, messages, model, temperature=0.7, max_tokens=None):
            async for ev in self._agent.run(messages, model=model, temperature=temperature, max_tokens=max_tokens):
                yield ev

    async def go():
        db = LocalDBClient(os.path.join(tempfile.mkdtemp(prefix="kb-ref-oracle-"), "ref.db"))
        if hasattr(db, "initialize"):
            await db.initialize()
        n_chars = 0
        t0 = time.perf_counter()
        for i in range(n_threads):
            rnd = random.Random(1234 + i)
            text = "".join(chr(rnd.randrange(32, 127)) for _ in range(48))
            async with Kafka(f"bench-{i:04d}", db) as k:
                async for ev in k.run_with_thread([Message(role="user", content=text)], model=provider.model,
                                                  temperature=0.0, max_tokens=max_tokens):
                    ch = ev.get("choices")
                    if ch and ch[0].get("delta", {}).get("content"):
                        n_chars += len(ch[0]["delta"]["content"])
        return time.perf_counter() - t0, n_chars
    before = provider.tokens_generated
    with contextlib.redirect_stdout(sys.stderr):
        dt, n_chars = asyncio.run(go())
    n_tok = provider.tokens_generated - before
    return {"model": cfg.name, "threads": n_threads, "max_tokens": max_tokens, "tokens": n_tok, "wall_s": dt,
            "tokens_per_s": n_tok / dt, "cores": cores, "streamed_chars": n_chars,
            "api": "CPUOracleProvider inside the reference's KafkaAgent.run_with_thread -> Agent.run -> "
                   "stream_completion (unmodified reference host code), one thread after the other; every call "
                   "re-runs its whole prompt (agent tool schemas included), as the reference's provider does"}


def run_reference(args, rank, world):
    if rank != 0:
        return
    from kafka_b200.model_config import get_config
    cfg = get_config(args.model)
    ctx = args.history + args.user_tokens
    steps = max(1, min(args.steps, 8))
    n_thr, alias = cpu_sample_size(cfg, ctx, args.cpu_threads)
    cb = cpu_baseline(cfg, n_thr, ctx, steps, warmup=max(1, min(args.warmup, 2)), alias=alias)
    server = {}
    if not args.no_reference_server:
        for name, fn in (("config1_thread_replay", reference_config1), ("cpu_provider_in_reference_agent",
                                                                         reference_oracle_in_agent)):
            try:
                server[name] = fn()
            except Exception as ex:          # the headline sample above must not be lost to a host-side surprise
                server[name] = {"failed": repr(ex)}
    line = {
        "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": args.warmup, "ms_per_step": cb["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{cfg.name} decode, {n_thr} threads x {args.history}-token history "
                               f"+ {args.user_tokens}-token turn (bounded CPU sample of the {args.threads}-thread "
                               "workload the B200 arm runs per GPU)", "threads": n_thr, "context": ctx,
                   "b200_arm_threads_per_gpu": args.threads, "weights_aliased": alias},
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "reference_server": server or None,
        "note": "the reference's own implementation of the path is a remote gateway call "
                "(src/llm/portkey.py:379), not runnable offline: `value` is the fp32 CPU oracle port of the path on "
                "the host cores; `reference_server` times the reference's own (unmodified) host code around the "
                "path — config 1 whole, and the CPU provider inside its agent / thread loop",
    }
    emit(line)


# ------------------------------------------------------------------------------------------------
def run_b200(args, rank, world, local_rank):
    from kafka_b200 import cabi
    from kafka_b200.cabi import KB_BLOCK_TOKENS
    from kafka_b200.engine_core import EngineCore
    from kafka_b200.model_config import get_config
    from kafka_b200.weights import generate_fused_on_device

    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device; there is no CPU fallback")
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    cfg = get_config(args.model)
    K, W = args.steps, max(args.warmup, 3)
    passes = 1 + (0 if args.no_profile else 1) + (0 if args.no_e2e else 1)
    tokens_per_thread = args.history + args.user_tokens + W * passes + K * passes + 8
    if not args.no_provider:  # the provider-level turn: history + re-templated turn (~90 tokens) + gen_tokens
        tokens_per_thread = max(tokens_per_thread,
                                args.history + 192 + args.gen_tokens + 320 * args.agent_tool_rounds)
    bpt = (tokens_per_thread + KB_BLOCK_TOKENS - 1) // KB_BLOCK_TOKENS
    B = args.threads

    weights = generate_fused_on_device(cfg, seed=rank, device=dev)
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    free, total = torch.cuda.mem_get_info(dev)
    block_bytes = EngineCore.kv_block_bytes(cfg)
    prefill_chunk = 4096      # measured (tools/prefill_profile.py --chunk): 433 ms per 32 768-token turn prefill against 452 at 8 192 and 444 at 2 048
    ws_guess = 3 * (1 << 30)
    fit = int((free - ws_guess) // (block_bytes * bpt))
    resident_note = None
    if fit < B:
        resident_note = f"only {fit} of {B} threads fit in {free / 2**30:.0f} GiB free HBM"
        B = fit
    eng = EngineCore(cfg, weights, dev, n_blocks=B * bpt, max_tokens=max(prefill_chunk, B), max_seqs=B)
    # synthetic resident history: random K/V in every block (values do not affect timing)
    for L in range(cfg.n_layers):
        eng.k_pool[L].normal_(0, 1)
        eng.v_pool[L].normal_(0, 1)
    g = torch.Generator(device="cpu").manual_seed(1234 + rank)
    block_ids = torch.randperm(B * bpt, generator=g).int().view(B, bpt).to(dev)  # scattered blocks
    ctx_lens = torch.full((B,), args.history, dtype=torch.int32, device=dev)
    sp = cabi.make_sampling(temperature=0.0, banned_token=cfg.vocab - 1)  # greedy, EOS masked

    # ---- the turn's incremental prefill (64 new tokens per thread), chunked --------------------
    U = args.user_tokens
    user = torch.randint(0, 256, (B, U), generator=g, dtype=torch.int32).to(dev)
    first = torch.empty(B, dtype=torch.int32, device=dev)
    per_chunk = max(1, prefill_chunk // U)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = cabi.launch_count()
    ev0.record()
    for s0 in range(0, B, per_chunk):
        s1 = min(B, s0 + per_chunk)
        n = s1 - s0
        cu = torch.arange(0, (n + 1) * U, U, dtype=torch.int32, device=dev)
        eng.prefill(user[s0:s1].reshape(-1), cu, ctx_lens[s0:s1], block_ids[s0:s1], U, sp, None, first[s0:s1])
    ev1.record()
    torch.cuda.synchronize()
    prefill_ms = ev0.elapsed_time(ev1)
    ctx_lens += U
    ctx0 = args.history + U
    prefill_flops = B * (U * 2 * cfg.matmul_params() + 2 * cfg.lm_head_params()
                         + 4 * cfg.head_dim * cfg.n_heads * cfg.n_layers * (args.history * U + U * U / 2))

    bufs = [first, torch.empty_like(first)]
    state = {"cur": 0, "ctx": ctx0}

    def decode_steps(n, host_io=False, pinned_in=None, pinned_out=None):
        for _ in range(n):
            a, b = bufs[state["cur"]], bufs[state["cur"] ^ 1]
            if host_io:
                a.copy_(pinned_in, non_blocking=True)           # H2D of this step's input tokens
            eng.decode(a, ctx_lens, block_ids, sp, b, max_ctx=state["ctx"], advance_ctx=True)
            if host_io:
                pinned_out.copy_(b, non_blocking=True)          # D2H of the sampled tokens
                torch.cuda.current_stream().synchronize()
                pinned_in.copy_(pinned_out)                     # host feeds the next step
            state["cur"] ^= 1
            state["ctx"] += 1

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            torch.distributed.barrier()

    def timed(n, **kw):
        """K steps between barrier+sync, CUDA events on the launching stream; max over ranks."""
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ctx_start = state["ctx"]
        e0.record()
        decode_steps(n, **kw)
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
            ms = float(t.item())
        return ms, ctx_start

    # ---- value: device-resident decode -----------------------------------------------------------
    decode_steps(W)
    sampler = ClockSampler(local_rank)
    sampler.start()
    l0 = cabi.launch_count()
    # KB_PROFILER_RANGE=1: cudaProfilerStart / Stop around exactly the timed steps, so that
    # `ncu --profile-from-start off` lists the launches of the timed region and nothing of the set-up
    ranged = os.environ.get("KB_PROFILER_RANGE") == "1"
    if ranged:
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
    ms_total, ctx_start = timed(K)
    if ranged:
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    gpu_launches = cabi.launch_count() - l0
    clocks = sampler.stop()
    ms_step = ms_total / K
    value = world * B * K / (ms_total / 1e3)

    kv_tok = cfg.kv_bytes_per_token
    hbm_peak, tf_peak, peak_kind = peaks()

    def step_bytes(ctx):  # W + sum (ctx_i + 1) kv_tok + B kv_tok
        return cfg.weight_bytes() + B * (ctx + 1) * kv_tok + B * kv_tok

    avg_step_bytes = sum(step_bytes(ctx_start + i) for i in range(K)) / K
    step_gbs = avg_step_bytes / (ms_step / 1e3) / 1e9

    # ---- roofline: per-kernel times from CUDA events between launches ----------------------------
    roofline = None
    kernels = None
    if not args.no_profile:
        decode_steps(W)
        barrier()
        ctx_p = state["ctx"]
        eng.profile_begin()
        decode_steps(K)
        prof = eng.profile_end()
        kernels = {k: {"ms_per_step": v[0] / K, "launches_per_step": v[1] / K} for k, v in prof.items() if v[1]}
        attn_ms, attn_n = prof["attention"]
        # algorithmic bytes of one attention launch = K and V of every cached token of every thread, one layer
        avg_kv_len = ctx_p + (K - 1) / 2 + 1
        attn_bytes = B * avg_kv_len * cfg.kv_dim * 2 * 2
        achieved = attn_bytes / (attn_ms / attn_n / 1e3) / 1e9
        gemm_ms = sum(prof[k][0] for k in ("gemm_qkv", "gemm_o", "gemm_gate_up", "gemm_down")) / K
        gemm_tf = B * 2 * cfg.matmul_params() / (gemm_ms / 1e3) / 1e12
        roofline = {
            "bound": "hbm", "kernel": "attn_decode_kernel<128,4>", "achieved": achieved, "peak": hbm_peak,
            "unit": "GB/s", "frac": achieved / hbm_peak, "traffic": None, "peak_kind": peak_kind,
            "bytes_per_launch": attn_bytes, "us_per_launch": attn_ms / attn_n * 1e3,
            "share_of_step": attn_ms / sum(v[0] for v in prof.values()),
            "step": {"algorithmic_bytes": avg_step_bytes, "achieved": step_gbs, "frac": step_gbs / hbm_peak},
            "gemm": {"bound": "tensor", "achieved": gemm_tf, "peak": tf_peak, "unit": "TFLOP/s",
                     "frac": gemm_tf / tf_peak, "ms_per_step": gemm_ms},
        }

        # The projections on their own: every layer's weights once per shape (32 distinct matrices = cold in L2,
        # as in the step, where 4.5 GB of K/V pass between two GEMMs), back to back with programmatic dependent
        # launch and NO event between launches. The per-kernel profile above brackets every launch with event
        # records, which cost device time and keep a kernel's set-up from overlapping its predecessor's tail
        # (~10 us per launch, profiles/r02_gemm_splitk.md) — it overstates the GEMMs' share of the real step.
        try:
            a_h = torch.randn(B, cfg.hidden, device=dev).half()
            a_q = torch.randn(B, cfg.q_dim, device=dev).half()
            a_f = torch.randn(B, cfg.ffn, device=dev).half()
            xres = torch.zeros(B, cfg.hidden, device=dev)
            o_qkv = torch.empty(B, cfg.qkv_dim, dtype=torch.float16, device=dev)
            o_h = torch.empty(B, cfg.ffn, dtype=torch.float16, device=dev)
            shapes = {"gemm_qkv": (a_h, "wqkv", cabi.EPI_F16, None, o_qkv),
                      "gemm_o": (a_q, "wo", cabi.EPI_RESID_F32, xres, xres),
                      "gemm_gate_up": (a_h, "w_gate_up", cabi.EPI_SWIGLU_F16, None, o_h),
                      "gemm_down": (a_f, "w_down", cabi.EPI_RESID_F32, xres, xres)}
            alone, tot_us = {}, 0.0
            for name, (a_, wname, epi, resid, out_) in shapes.items():
                ws_ = [weights[f"l{i}.{wname}"] for i in range(cfg.n_layers)]
                for w_ in ws_[:3]:
                    cabi.gemm(a_, w_, epi, resid=resid, out=out_)
                g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize()
                g0.record()
                for w_ in ws_:
                    cabi.gemm(a_, w_, epi, resid=resid, out=out_)
                g1.record()
                torch.cuda.synchronize()
                us = g0.elapsed_time(g1) / len(ws_) * 1e3
                n_, k_ = ws_[0].shape
                alone[name] = {"us_per_launch": us, "tflops": 2.0 * B * n_ * k_ / (us * 1e-6) / 1e12}
                tot_us += us
            tf_alone = B * 2 * cfg.matmul_params() / cfg.n_layers / (tot_us * 1e-6) / 1e12
            roofline["gemm"]["standalone"] = {
                "achieved": tf_alone, "frac": tf_alone / tf_peak, "ms_per_step": tot_us * cfg.n_layers / 1e3,
                "per_launch": alone,
                "what": "each projection over all layers' weights back to back, no events between launches (the "
                        "profile's per-kernel events add ~10 us to every launch)"}
            # Which of the two GEMM figures the un-instrumented step agrees with: the profile's per-kernel times
            # sum to MORE than the step the headline `value` was timed on (events between launches), the
            # standalone figure plus the profile's other kernels should land on it.
            prof_sum = sum(v[0] for v in prof.values()) / K
            roofline["gemm"]["consistency"] = {
                "step_ms_timed_without_events": ms_step, "sum_of_profiled_kernels_ms": prof_sum,
                "profiled_non_gemm_ms_plus_standalone_gemm_ms": prof_sum - gemm_ms + tot_us * cfg.n_layers / 1e3}
            est = roofline["gemm"]["consistency"]["profiled_non_gemm_ms_plus_standalone_gemm_ms"]
            g = roofline["gemm"]
            g["in_step_profile"] = {"achieved": g["achieved"], "frac": g["frac"], "ms_per_step": g["ms_per_step"]}
            if abs(est - ms_step) < abs(prof_sum - ms_step):
                # the standalone times explain the un-instrumented step better: they are the GEMM roofline figure
                g.update(achieved=tf_alone, frac=tf_alone / tf_peak, ms_per_step=tot_us * cfg.n_layers / 1e3,
                         source="standalone (see consistency)")
            else:
                g["source"] = "in-step profile"
            del a_h, a_q, a_f, xres, o_qkv, o_h
        except Exception as ex:      # a measurement aid must not cost the bench line
            roofline["gemm"]["standalone"] = {"failed": repr(ex)}
        # what a kernel that only READS reaches here: one pass over ~16 GB of the K pool (>> L2), best of 3
        n_l = max(1, min(cfg.n_layers, int(16e9 // (eng.k_pool[0].numel() * 2))))
        region = eng.k_pool[:n_l]
        sink = torch.zeros(4 * 160, dtype=torch.int32, device=dev)
        best = None
        for _ in range(4):
            r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            r0.record()
            nbytes = cabi.hbm_read_probe(region, sink)
            r1.record()
            torch.cuda.synchronize()
            t = r0.elapsed_time(r1)
            best = t if best is None else min(best, t)
        read_gbs = nbytes / (best / 1e3) / 1e9
        roofline["read_only_probe"] = {"achieved": read_gbs, "unit": "GB/s", "bytes": nbytes,
                                       "frac_of_it": achieved / read_gbs,
                                       "what": "kb_hbm_read_probe: 16-byte streaming loads over a slice of the K pool, "
                                               "best of 4 — the ceiling for a read-only kernel on this GPU, next to "
                                               "the copy-measured `peak`"}
        cap = ncu_dram_traffic()
        if cap is not None:
            # the capture's own launch: 512 threads at kv_len = history + user tokens + 1 (BASELINE configs[2])
            cap_alg = 512 * (2048 + 64 + 1) * cfg.kv_dim * 2 * 2
            roofline["traffic"] = cap[0]
            roofline["traffic_detail"] = {
                "source": "profiles/r02_ncu_attn_decode_dram.csv (ncu application replay, dram__bytes_read.sum + "
                          "dram__bytes_write.sum, mean of %d launches)" % cap[1],
                "launch": "512 threads, kv_len 2113", "algorithmic_bytes_of_that_launch": cap_alg,
                "traffic_over_algorithmic": cap[0] / cap_alg,
                "same_shape_as_this_run": bool(B == 512 and args.history == 2048 and U == 64 and cfg.name == "llama-3-8b")}
        if roofline["frac"] > 1.0:
            # the peak is a COPY measurement (half reads, half writes, with the bus turn-arounds that costs); a
            # read-only stream can exceed it. Not an error of the byte count: ncu's DRAM counters agree with the
            # algorithmic bytes to 2 % (traffic_detail).
            roofline["frac_note"] = ("above 1: MEASURED_PEAKS' hbm_gbs is a read+write copy; this kernel only reads "
                                     "(ncu dram bytes = %.3f x algorithmic)" %
                                     (roofline.get("traffic_detail", {}).get("traffic_over_algorithmic", float("nan"))))

    # ---- e2e: host buffers through the C ABI -----------------------------------------------------
    e2e = None
    if not args.no_e2e:
        pin_in = torch.empty(B, dtype=torch.int32, pin_memory=True)
        pin_out = torch.empty(B, dtype=torch.int32, pin_memory=True)
        pin_in.copy_(bufs[state["cur"]].cpu())
        decode_steps(W, host_io=True, pinned_in=pin_in, pinned_out=pin_out)
        ms_e2e, _ = timed(K, host_io=True, pinned_in=pin_in, pinned_out=pin_out)
        e2e = {"value": world * B * K / (ms_e2e / 1e3), "unit": UNIT, "h2d_bytes_per_step": B * 4,
               "d2h_bytes_per_step": B * 4, "ms_per_step": ms_e2e / K,
               "api": "kb_decode via ctypes, pinned host token buffers, per-step stream sync"}

    # ---- e2e through the plugin API (the call a user makes) ----------------------------------------
    e2e_cabi = e2e
    if not args.no_provider:
        barrier()
        pt = provider_turn(eng, cfg, B, args.history, U, args.gen_tokens, seed=rank, pipeline=not args.no_pipeline,
                           tool_rounds=args.agent_tool_rounds)
        pt["pipelined_loop"] = not args.no_pipeline
        wall = pt["wall_s"]
        if world > 1:
            t = torch.tensor([wall], device=dev)
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
            wall = float(t.item())
        pt["value"] = world * pt["new_tokens"] / wall
        pt["ms_per_step"] = 1e3 * wall / pt["decode_steps"]
        e2e = pt
        # per-turn TTFT against history length (north_star: independent of it): 16 concurrent threads
        if rank == 0 and world == 1 and not args.no_ttft_sweep:   # N = 1 only: no rank waits on rank 0's extras
            sweep = {}
            for h in (256, 1024, 2048, 4096, 7936):      # 7936 + turn = the 8192-position limit
                if h + 192 + 32 > cfg.max_pos:
                    continue
                # three turns on fresh threads per length, the median of their p50s: one turn of 16 threads is a
                # single prefill call, and a one-off (first launch of a GEMM shape, a GC pass) is ~10 ms of it
                rs = [provider_turn(eng, cfg, 16, h, U, 32, seed=1000 + h + 7919 * k) for k in range(3)]
                p50s = sorted(r["ttft_p50_ms"] for r in rs)
                r = rs[0]
                sweep[str(h)] = {"ttft_p50_ms": p50s[1], "ttft_p50_ms_each_turn": [x["ttft_p50_ms"] for x in rs],
                                 "cached_fraction": r["cached_fraction_of_prompt"],
                                 "prefilled_new_tokens": r["prefilled_new_tokens"]}
            e2e["ttft_vs_history_16_threads"] = sweep
            try:    # the same turn through the HTTP server (first SSE line with content)
                e2e["http_16_threads"] = http_turn(eng, cfg, 16, args.history, U, 32)
            except Exception as ex:   # a missing HTTP dependency must not cost the bench line
                e2e["http_16_threads"] = {"unavailable": repr(ex)}

    cb = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        n_thr, alias = cpu_sample_size(cfg, ctx0, args.cpu_threads)
        cb = cpu_baseline(cfg, n_thr, ctx0, args.cpu_steps, alias=alias)

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16",
            "dtype_detail": "16-bit operands, fp32 accumulation and residual stream: the weights are bf16 VALUES (the "
                            "five projection matrices bound re-encoded as fp16, exact for |w| >= 2^-14); every other "
                            "16-bit value — GEMM activations, q / k / v, the KV cache, the attention probabilities — "
                            "is fp16 (10 mantissa bits instead of 7, same bytes and MMA rate; conversions saturate at "
                            "65504). Reason and error budget: DESIGN.md §5, profiles/r02_stage_errors.md",
            "data": "synthetic",
            "config": {
                "workload": f"{cfg.name} bf16 decode, {B} resident threads/GPU x {args.history}-token history + "
                            f"{U}-token turn (BASELINE configs[2], thread-sharded: 4096 threads at 8 GPUs)",
                "model": cfg.name, "threads_per_gpu": B, "history": args.history, "user_tokens": U,
                "context_at_start": ctx_start, "kv_block_tokens": KB_BLOCK_TOKENS,
                "l2": "inputs larger than L2 (KV read per step >> 126 MB); no explicit flush",
                "parallelism": f"replica per GPU, thread-sharded x{world}, no collective",
                "resident_note": resident_note,
            },
            "clocks": clocks, "gpu_launches": gpu_launches, "e2e": e2e, "e2e_cabi": e2e_cabi, "roofline": roofline,
            "kernels": kernels,
            "cpu_baseline": cb,
            "prefill": {"ms": prefill_ms, "tokens": B * U, "tflops": prefill_flops / (prefill_ms / 1e3) / 1e12,
                        "frac_of_tensor_peak": prefill_flops / (prefill_ms / 1e3) / 1e12 / tf_peak,
                        "launches": l0 - launches0},
        }
        emit(line)


def run_config2(args):
    """BASELINE configs[1] at its stated size (oracle/config2.py): the CPU-vs-B200 measurement SURVEY.md §8(d)
    asks for, and the greedy-agreement criterion on the way."""
    import asyncio
    import tempfile
    from kafka_b200.db import ThreadStore
    from kafka_b200.factory import build_engine
    from kafka_b200.model_config import get_config
    from kafka_b200.provider import B200LLMProvider
    from kafka_b200.tokenizer import ByteTokenizer
    from kafka_b200.weights import generate_weights
    from oracle.config2 import run_config2 as drive
    from oracle.cpu_provider import BatchedOracle
    cores = pick_cpu_threads()
    cfg = get_config("tinyllama-1.1b")
    w = generate_weights(cfg, seed=0, init="margin")
    eng = build_engine(cfg, 0, n_blocks=64 * 16, max_seqs=64, max_tokens=8192, weights=w)
    gpu = B200LLMProvider(eng, ByteTokenizer(cfg.vocab), model=cfg.name, default_temperature=0.0)
    store = ThreadStore(os.path.join(tempfile.mkdtemp(prefix="kb-config2-"), "threads.db"))
    from kafka_b200 import cabi
    l0 = cabi.launch_count()
    try:
        r = asyncio.run(drive(gpu, BatchedOracle(cfg, w), store, n_threads=64, n_turns=8, max_tokens=32))
    finally:
        eng.shutdown()
    r.pop("cached")
    emit({"metric": METRIC, "value": r["b200_tokens_per_s"], "unit": UNIT, "n_gpus": 1, "steps": r["turns"],
          "warmup": 0, "ms_per_step": 1e3 * r["b200_wall_s"] / r["turns"], "higher_is_better": True,
          "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
          "config": {"workload": "BASELINE configs[1]: tinyllama-1.1b shapes (margin init), 64 threads x 8 turns x 32 "
                                 "greedy tokens through B200LLMProvider.completion, histories in SQLite; a step is "
                                 "one turn of all 64 threads; wall clock (first turn includes first-launch costs)"},
          "gpu_launches": cabi.launch_count() - l0,
          "cpu_baseline": {"value": r["cpu_tokens_per_s"], "unit": UNIT, "cores": cores, "kind": "port",
                           "sample": "the same 512 thread-turns, whole: fp32 oracle with packed projections and "
                                     "its K/V kept between turns (oracle/cpu_provider.BatchedOracle)",
                           "wall_s": r["cpu_wall_s"]},
          "parity": {k: r[k] for k in ("thread_turns", "mismatched_thread_turns", "diverged_threads",
                                       "agreement_fraction_of_threads", "first_mismatches")}})


def run_tp(args, rank, world, local_rank):
    """BASELINE config 5 shape: one model sharded over all `world` GPUs (tensor parallel; the sums after the
    O and down projections by our own kernels over NVSwitch multicast memory, see kafka_b200/tp.py), all
    ranks stepping the same batch in lockstep."""
    from kafka_b200 import cabi
    from kafka_b200.cabi import KB_BLOCK_TOKENS
    from kafka_b200.engine_core import EngineCore
    from kafka_b200.model_config import get_config
    from kafka_b200.tp import build_tp_core, local_config

    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    cfg = get_config(args.model)
    loc = local_config(cfg, world)
    K, W = args.steps, max(args.warmup, 3)
    U = args.user_tokens
    n_meas = 4 if args.tp_compare else 1
    bpt = (args.history + U + n_meas * (W + 2 * K) + 8 + KB_BLOCK_TOKENS - 1) // KB_BLOCK_TOKENS
    B = args.threads
    # weights first (they decide how much HBM is left for KV); a 1-block core, then the real one
    free0, _ = torch.cuda.mem_get_info(dev)
    w_bytes = 2 * (loc.matmul_params() + 2 * cfg.vocab * cfg.hidden)
    fit = int((free0 - w_bytes - 6 * (1 << 30)) // (EngineCore.kv_block_bytes(loc) * bpt))
    fit_t = torch.tensor([fit], device=dev)
    torch.distributed.all_reduce(fit_t, op=torch.distributed.ReduceOp.MIN)
    fit = int(fit_t.item())
    resident_note = None
    if fit < B:
        resident_note = f"only {fit} of {B} threads resident per the smallest rank's free HBM"
        B = fit
    core, grp = build_tp_core(cfg, rank, world, dev, n_blocks=B * bpt, max_tokens=max(4096, B), max_seqs=B, seed=0,
                              allreduce=args.tp_allreduce)
    for L in range(loc.n_layers):
        core.k_pool[L].normal_(0, 1)
        core.v_pool[L].normal_(0, 1)
    g = torch.Generator(device="cpu").manual_seed(1234)            # identical inputs on every rank
    block_ids = torch.randperm(B * bpt, generator=g).int().view(B, bpt).to(dev)
    ctx_lens = torch.full((B,), args.history, dtype=torch.int32, device=dev)
    sp = cabi.make_sampling(temperature=0.0, banned_token=cfg.vocab - 1)
    user = torch.randint(0, 256, (B, U), generator=g, dtype=torch.int32).to(dev)
    first = torch.empty(B, dtype=torch.int32, device=dev)
    per_chunk = max(1, 4096 // U)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.distributed.barrier()
    e0.record()
    for s0 in range(0, B, per_chunk):
        s1 = min(B, s0 + per_chunk)
        cu = torch.arange(0, (s1 - s0 + 1) * U, U, dtype=torch.int32, device=dev)
        core.prefill(user[s0:s1].reshape(-1), cu, ctx_lens[s0:s1], block_ids[s0:s1], U, sp, None, first[s0:s1])
    e1.record()
    torch.cuda.synchronize()
    prefill_ms = e0.elapsed_time(e1)
    ctx_lens += U
    bufs, cur, ctx = [first, torch.empty_like(first)], 0, args.history + U

    def steps(n):
        nonlocal cur, ctx
        for _ in range(n):
            core.decode(bufs[cur], ctx_lens, block_ids, sp, bufs[cur ^ 1], max_ctx=ctx, advance_ctx=True)
            cur ^= 1
            ctx += 1

    def measure():
        """W warm-up steps, K timed steps (max over ranks), K more under the per-kernel event profile."""
        nonlocal ctx
        steps(W)
        sampler = ClockSampler(local_rank)
        sampler.start()
        torch.cuda.synchronize()
        torch.distributed.barrier()
        l0 = cabi.launch_count()
        c0 = ctx
        e0.record()
        steps(K)
        e1.record()
        torch.cuda.synchronize()
        torch.distributed.barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        n_launch = cabi.launch_count() - l0
        clk = sampler.stop()
        core.profile_begin()
        cp = ctx
        steps(K)
        return {"ms_total": float(t.item()), "launches": n_launch, "clocks": clk, "prof": core.profile_end(),
                "ctx_start": c0, "ctx_p": cp, "mode": grp.mode}

    others = {}
    if args.tp_compare and hasattr(grp, "peers"):
        # the same resident state under the other all-reduce implementations (switchable at run time)
        from kafka_b200.tp import TPGroup
        blocks = int(os.environ.get("KAFKA_B200_TP_BLOCKS", "0"))
        main_wire = grp.wire
        if core.lm_head_slice is None:      # the NCCL callback needs a replicated lm_head (KAFKA_B200_TP_LM_HEAD)
            nccl = TPGroup()
            nccl.attach(core, extra_base=grp.buf)
            grp.mode = "nccl"
            r = measure()
            others["nccl"] = {"ms_per_step": r["ms_total"] / K, "allreduce_ms_per_step": r["prof"]["allreduce"][0] / K,
                              "rmsnorm_ms_per_step": r["prof"]["rmsnorm"][0] / K}
        for wire in ("f32", "f16", "fused"):
            if wire == main_wire:
                continue
            grp.attach(core, blocks, wire=wire)
            r = measure()
            others[grp.mode] = {"ms_per_step": r["ms_total"] / K,
                                "allreduce_ms_per_step": r["prof"]["allreduce"][0] / K,
                                "rmsnorm_ms_per_step": r["prof"]["rmsnorm"][0] / K}
        grp.attach(core, blocks, wire=main_wire)
    m = measure()
    ms_total, launches, clocks, prof, ctx_start, ctx_p = (m["ms_total"], m["launches"], m["clocks"], m["prof"],
                                                          m["ctx_start"], m["ctx_p"])
    if rank != 0:
        return
    hbm_peak, tf_peak, peak_kind = peaks()
    attn_ms, attn_n = prof["attention"]
    attn_bytes = B * (ctx_p + (K - 1) / 2 + 1) * loc.kv_dim * 2 * 2
    achieved = attn_bytes / (attn_ms / attn_n / 1e3) / 1e9
    ar_ms, ar_n = prof["allreduce"]
    ar_bytes = (4 if grp.mode in ("nccl", "peers-multicast", "peers-p2p") else 2) * B * cfg.hidden
    line = {
        "metric": METRIC, "value": B * K / (ms_total / 1e3), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "bf16", "dtype_detail": "bf16 weight values (projection matrices bound as fp16 re-encodings); activations, "
                                         "q / k / v, KV cache and attention probabilities fp16; fp32 accumulate",
        "data": "synthetic",
        "config": {"workload": f"{cfg.name} tensor-parallel x{world} decode, {B} threads x {args.history}-token history "
                               f"+ {U}-token turn (BASELINE configs[4])", "model": cfg.name, "threads": B,
                   "history": args.history, "context_at_start": ctx_start, "resident_note": resident_note,
                   "parallelism": f"tp{world}: heads/ffn sharded, 2 all-reduces ({grp.mode}) of [T,{cfg.hidden}] per layer, "
                                  + ("lm_head by vocabulary slice (local winner + exchange of (value, id) pairs)"
                                     if core.lm_head_slice else "lm_head replicated"),
                   "l2": "inputs larger than L2; no explicit flush"},
        "clocks": clocks, "gpu_launches": launches, "e2e": None,
        "roofline": {"bound": "hbm", "kernel": "attn_decode_kernel<128,8>", "achieved": achieved, "peak": hbm_peak,
                     "unit": "GB/s", "frac": achieved / hbm_peak, "traffic": None, "peak_kind": peak_kind,
                     "bytes_per_launch": attn_bytes, "us_per_launch": attn_ms / attn_n * 1e3,
                     "share_of_step": attn_ms / sum(v[0] for v in prof.values())},
        "allreduce": {"calls_per_step": ar_n / K, "ms_per_step": ar_ms / K, "bytes_per_call": ar_bytes,
                      "busbw_GBps": (2 * (world - 1) / world) * ar_bytes / (ar_ms / ar_n / 1e3) / 1e9,
                      "backend": {"nccl": "NCCL via torch.distributed on the kernels' stream",
                                  "peers-multicast": "own kernel: multimem.ld_reduce / multimem.st over the NVSwitch "
                                                     "multicast mapping of the residual buffers (kb_collective.cu)",
                                  "peers-multicast-f16": "own kernel: fp16 partial sums, multimem.ld_reduce (fp32 "
                                                         "accumulate in the switch) / multimem.st over the NVSwitch "
                                                         "multicast mapping (kb_collective.cu); the sum is added "
                                                         "to the fp32 residual by the following RMSNorm",
                                  "peers-p2p-f16": "own kernel: fp16 partial sums, peer loads / stores over NVLink",
                                  "peers-multicast-fused": "own kernel per projection: multimem.ld_reduce of the fp16 "
                                                           "partial sums of this rank's rows + residual add + RMSNorm "
                                                           "+ multimem.st of the normed fp16 rows to all ranks "
                                                           "(kb_tp_reduce_norm_gather, kb_collective.cu)",
                                  "peers-p2p-fused": "own kernel per projection: reduce + residual add + RMSNorm + "
                                                     "gather over NVLink peer loads / stores",
                                  "peers-p2p": "own kernel: peer loads / stores over NVLink (kb_collective.cu)"}[grp.mode],
                      "other_modes_same_state": others or None},
        "kernels": {k: {"ms_per_step": v[0] / K, "launches_per_step": v[1] / K} for k, v in prof.items() if v[1]},
        "prefill": {"ms": prefill_ms, "tokens": B * U}, "cpu_baseline": None,
    }
    emit(line)


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL is used for the barrier / max-over-ranks only. Its log level is the caller's (NCCL_DEBUG is
        # left as exported); whatever native code prints (the NCCL banner, INFO lines) goes to stderr: fd 1
        # is pointed at stderr for the rest of the run and the ONE JSON line is written to the saved descriptor
        global _JSON_FD
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)
        torch.cuda.set_device(local_rank)
        torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        if args.config == 2:
            if rank == 0:
                run_config2(args)
        elif args.tp and world > 1:
            run_tp(args, rank, world, local_rank)
        else:
            run_b200(args, rank, world, local_rank)
    finally:
        if world > 1:
            torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
