#!/usr/bin/env python
"""Generate the committed fixtures under tests/golden/.  Run in the build container, where
/root/reference is mounted:   python tests/golden/make_golden.py

  ref_threads_rows.json      rows of the reference's bundled data/threads.db (messages + threads), so
                             the history-assembly test can rebuild the database without /root/reference
  history_assembly.json      per-thread digests produced BY THE REFERENCE'S OWN CODE
                             (LocalDBClient.get_thread_messages + sanitize_messages_for_openai) over
                             that database: the one result near the path that the reference pins
  agent_events_echo.json     event stream of the reference's Agent.run driven by a scripted provider
                             (tool round + text round): what a drop-in provider's chunks must produce
  template_snapshots.json    this repo's tokenisation of every thread of the fixture (regression pin)
  oracle_tiny.json           fp32 oracle logits / greedy tokens on the two tiny test models
"""
from __future__ import annotations

import asyncio
import hashlib
import json
import sqlite3
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
for p in (str(ROOT), str(ROOT / "kafka-llm-service_b200"), str(ROOT / "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import ref_harness  # noqa: E402


def digest(obj) -> str:
    return hashlib.sha256(json.dumps(obj, sort_keys=True).encode()).hexdigest()


def dump(name, obj):
    (HERE / name).write_text(json.dumps(obj, indent=1, sort_keys=True) + "\n")
    print(f"wrote {name}")


def export_db_rows():
    con = sqlite3.connect(str(ref_harness.REF / "data" / "threads.db"))
    threads = [list(r) for r in con.execute("SELECT id, created_at, metadata, sandbox_id FROM threads ORDER BY rowid")]
    msgs = [list(r) for r in con.execute(
        "SELECT id, thread_id, message, metadata, created_at FROM messages ORDER BY rowid")]
    dump("ref_threads_rows.json", {"threads": threads, "messages": msgs})
    return threads, msgs


async def reference_history():
    from src.db.local import LocalDBClient
    from src.kafka.utils import sanitize_messages_for_openai
    import contextlib
    import io
    db = LocalDBClient(str(ref_harness.REF / "data" / "threads.db"))
    con = sqlite3.connect(str(ref_harness.REF / "data" / "threads.db"))
    ids = sorted(r[0] for r in con.execute("SELECT DISTINCT thread_id FROM messages"))
    per_thread = {}
    n_loaded = n_kept = 0
    all_dicts = []
    for tid in ids:
        msgs = await db.get_thread_messages(tid)
        with contextlib.redirect_stdout(io.StringIO()):
            kept = sanitize_messages_for_openai(msgs)
        dicts = [m.to_dict() for m in kept]
        per_thread[tid] = {"loaded": len(msgs), "kept": len(kept), "sha256": digest(dicts)}
        n_loaded += len(msgs)
        n_kept += len(kept)
        all_dicts.append(dicts)
    dump("history_assembly.json", {"threads": per_thread, "loaded": n_loaded, "kept": n_kept,
                                   "sha256_all": digest(all_dicts)})
    return ids


async def reference_agent_events():
    """Drive the reference's Agent.run with a scripted provider: round 1 emits a tool call (streamed in
    fragments), round 2 emits text. Records the event stream with volatile ids normalised."""
    from src.agents.base import Agent
    from src.llm.base import LLMProvider
    from src.llm.types import Message, StreamChunk, CompletionResponse
    from src.tools.base import ToolProvider
    from src.tools.types import Tool

    class Scripted(LLMProvider):
        def __init__(self):
            super().__init__()
            self.calls = []

        async def stream_completion(self, messages, *, temperature=None, max_tokens=None, stop=None, **kw):
            self.calls.append([m.to_dict() for m in messages])
            if len(self.calls) == 1:
                yield StreamChunk(role="assistant", model="m", id="x")
                yield StreamChunk(tool_calls=[{"index": 0, "id": "call_1", "type": "function",
                                               "function": {"name": "add", "arguments": ""}}])
                yield StreamChunk(tool_calls=[{"index": 0, "function": {"arguments": "{\"a\": 2,"}}])
                yield StreamChunk(tool_calls=[{"index": 0, "function": {"arguments": " \"b\": 3}"}}])
                yield StreamChunk(finish_reason="tool_calls")
            else:
                yield StreamChunk(role="assistant")
                yield StreamChunk(content="The answer")
                yield StreamChunk(content=" is 5")
                yield StreamChunk(finish_reason="stop")

        async def completion(self, messages, **kw):
            return CompletionResponse(content="")

    def add(a: int, b: int) -> str:
        return str(a + b)

    from src.tools.agent import AgentToolProvider
    tools = AgentToolProvider(tools=[Tool(name="add", description="add two ints",
                                          parameters={"type": "object",
                                                      "properties": {"a": {"type": "integer"},
                                                                     "b": {"type": "integer"}},
                                                      "required": ["a", "b"]}, handler=add)])
    await tools.connect()
    prov = Scripted()
    agent = Agent(llm_provider=prov, tool_provider=tools, system_prompt="SYS")
    events = []
    async for ev in agent.run([Message(role="user", content="2+3?")], model="m", temperature=0.0, max_tokens=16):
        ev = json.loads(json.dumps(ev))
        if "id" in ev and str(ev["id"]).startswith("chatcmpl-"):
            ev["id"] = "chatcmpl-*"
        if "created" in ev:
            ev["created"] = 0
        events.append(ev)
    dump("agent_events_echo.json", {"events": events, "provider_calls": prov.calls})


def template_snapshots(ids):
    from kafka_b200.db import message_from_json
    from kafka_b200.history import sanitize_messages
    from kafka_b200.template import ChatTemplate
    from kafka_b200.tokenizer import ByteTokenizer
    rows = json.loads((HERE / "ref_threads_rows.json").read_text())["messages"]
    tmpl = ChatTemplate(ByteTokenizer(512))
    out = {}
    for tid in ids:
        msgs = sanitize_messages([message_from_json(r[2]) for r in rows if r[1] == tid])
        toks = tmpl.render(msgs)
        out[tid] = {"n_tokens": len(toks), "sha256": hashlib.sha256(json.dumps(toks).encode()).hexdigest()}
    dump("template_snapshots.json", out)


def oracle_tiny():
    import torch
    from kafka_b200.model_config import get_config
    from kafka_b200.weights import generate_weights, generable_ids
    from oracle.llama_fp32 import OracleLlama
    out = {}
    for name in ("tiny-test", "tiny-test-hd128"):
        cfg = get_config(name)
        for init in ("iid", "margin"):
            o = OracleLlama(cfg, generate_weights(cfg, seed=0, init=init))
            g = torch.Generator().manual_seed(7)
            prompt = torch.randint(32, 127, (70,), generator=g)
            toks, lg = o.greedy(prompt, 8, allowed=generable_ids(cfg))
            out[f"{name}/{init}"] = {"prompt": prompt.tolist(), "tokens": toks,
                                     "logits_head": [[round(float(x), 5) for x in row[:8]] for row in lg],
                                     "logits_sum": [round(float(row.double().sum()), 4) for row in lg]}
    dump("oracle_tiny.json", out)


def main():
    if not ref_harness.load():
        raise SystemExit("/root/reference is not mounted: fixtures can only be regenerated in the build container")
    export_db_rows()
    ids = asyncio.run(reference_history())
    asyncio.run(reference_agent_events())
    template_snapshots(ids)
    oracle_tiny()


if __name__ == "__main__":
    main()
