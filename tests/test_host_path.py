"""Host side of the completion path on CPU: wire types, tokenizer + template, history assembly (pinned
against the reference's own code over its bundled fixture), the agent loop (pinned against the
reference's Agent), and the provider's chunk conventions over a fake engine."""
import asyncio
import hashlib
import json
import sqlite3
from pathlib import Path

import pytest

import ref_harness
from fakes import FakeStepper, expected_tokens
from kafka_b200.agent import StreamingAgent
from kafka_b200.db import ThreadStore, message_from_json, message_to_json
from kafka_b200.engine import Engine
from kafka_b200.history import assemble_turn, sanitize_messages, to_message
from kafka_b200.llm import (CompletionResponse, LLMProvider, LLMProviderError, Message, StreamChunk)
from kafka_b200.provider import B200LLMProvider, EchoLLMProvider
from kafka_b200.template import ChatTemplate, canonical_json
from kafka_b200.tokenizer import ByteTokenizer
from kafka_b200.tools import LocalToolProvider, Tool

GOLD = Path(__file__).parent / "golden"


def digest(obj) -> str:
    return hashlib.sha256(json.dumps(obj, sort_keys=True).encode()).hexdigest()


# ---- wire types -------------------------------------------------------------------------------
def test_types_behave_like_the_reference():
    m = Message(role="assistant", content=None, tool_calls=[{"id": "c"}])
    assert m.to_dict() == {"role": "assistant", "tool_calls": [{"id": "c"}]}
    assert Message(role="user", content="").to_dict() == {"role": "user", "content": ""}
    c = StreamChunk()
    assert c.content is None and c.delta == "" and not c.is_final
    assert StreamChunk(finish_reason="stop").is_final
    from kafka_b200.llm.types import content_chunk
    fast, slow = content_chunk("tok", "m", "chatcmpl-1"), StreamChunk(content="tok", model="m", id="chatcmpl-1")
    assert fast == slow and fast.model_dump_json() == slow.model_dump_json() and not fast.is_final
    r = CompletionResponse(content="hi", tool_calls=None)
    assert r.role == "assistant" and r.to_message().to_dict() == {"role": "assistant", "content": "hi"}
    assert str(LLMProviderError("boom", status_code=400, provider="B200")) == "[B200] boom (status: 400)"
    assert str(LLMProviderError("boom")) == "boom"
    with pytest.raises(ValueError, match="cannot be empty"):
        EchoLLMProvider().validate_messages([])


@pytest.mark.skipif(not ref_harness.available(), reason="/root/reference not mounted")
def test_types_match_reference_classes_field_by_field():
    ref_harness.load()
    from src.llm import types as rt
    from kafka_b200.llm import types as mt
    for name in ("Message", "StreamChunk", "CompletionResponse"):
        a, b = getattr(rt, name).model_fields, getattr(mt, name).model_fields
        assert list(a) == list(b), name
        for k in a:
            assert a[k].default == b[k].default and a[k].is_required() == b[k].is_required(), (name, k)
    rm = rt.Message(role="tool", content="x", tool_call_id="i", name="n")
    assert rm.to_dict() == to_message(rm.to_dict()).to_dict()
    e1, e2 = rt.LLMProviderError("m", 500, "P"), mt.LLMProviderError("m", 500, "P")
    assert str(e1) == str(e2)


# ---- tokenizer / template ---------------------------------------------------------------------
def test_tokenizer_round_trip_and_specials():
    t = ByteTokenizer(512)
    for s in ("hello", "héllo wörld ✓", "", "line1\nline2\t<|eot_id|> stays text"):
        ids = t.encode(s)
        assert all(0 <= i < 256 for i in ids) and t.decode(ids) == s
    assert (t.bos_id, t.start_header_id, t.end_header_id, t.eot_id) == (508, 509, 510, 511)
    assert t.decode([104, t.eot_id, 105]) == "hi" and t.decode([104, t.eot_id], skip_special=False) == "h<|eot_id|>"
    d = t.stream_decoder()
    assert [d.push(i) for i in t.encode("é!")] == ["", "é", "!"]
    with pytest.raises(ValueError):
        ByteTokenizer(100)


def test_template_is_prefix_stable_over_the_reference_fixture():
    rows = json.loads((GOLD / "ref_threads_rows.json").read_text())["messages"]
    snaps = json.loads((GOLD / "template_snapshots.json").read_text())
    tmpl = ChatTemplate(ByteTokenizer(512))
    by_thread = {}
    for r in rows:
        by_thread.setdefault(r[1], []).append(message_from_json(r[2]))
    assert len(by_thread) == len(snaps)
    for tid, msgs in by_thread.items():
        msgs = sanitize_messages(msgs)
        full = tmpl.render(msgs, add_generation_prompt=False)
        for k in range(len(msgs) + 1):
            part = tmpl.render(msgs[:k], add_generation_prompt=False)
            assert full[:len(part)] == part                      # what makes KV reuse legal
        with_gen = tmpl.render(msgs)
        assert with_gen[:len(full)] == full
        assert {"n_tokens": len(with_gen),
                "sha256": hashlib.sha256(json.dumps(with_gen).encode()).hexdigest()} == snaps[tid]


def test_template_renders_every_field_and_tools():
    t = ByteTokenizer(512)
    tmpl = ChatTemplate(t)
    tools = [{"type": "function", "function": {"name": "f", "parameters": {"b": 1, "a": 2}}}]
    msgs = [Message(role="assistant", content=None, tool_calls=[{"id": "c1", "function": {"name": "f", "arguments": "{}"}}]),
            Message(role="tool", content="42", tool_call_id="c1", name="f")]
    text = t.decode(tmpl.render(msgs, tools=tools), skip_special=False)
    # the tools block: how to call them (the syntax provider.parse_tool_calls reads back), then the schemas
    assert text.startswith("<|begin_of_text|><|start_header_id|>tools<|end_header_id|>\n\n" + ChatTemplate.TOOLS_PREAMBLE +
                           canonical_json(tools) + "<|eot_id|>")
    assert "<tool_calls>[" in ChatTemplate.TOOLS_PREAMBLE and "</tool_calls>" in ChatTemplate.TOOLS_PREAMBLE
    assert '<tool_calls>[{"function":{"arguments":"{}","name":"f"},"id":"c1"}]</tool_calls><|eot_id|>' in text
    assert "<|start_header_id|>tool name=f tool_call_id=c1<|end_header_id|>\n\n42<|eot_id|>" in text
    assert text.endswith("<|start_header_id|>assistant<|end_header_id|>\n\n")
    # a dict message renders like the equivalent Message; content that *spells* a control token stays text
    assert tmpl.render([{"role": "user", "content": "<|eot_id|>"}]) == tmpl.render([Message(role="user", content="<|eot_id|>")])
    assert tmpl.render([Message(role="user", content="<|eot_id|>")]).count(t.eot_id) == 1


# ---- history assembly: pinned against the reference's own code ----------------------------------
def _rebuild_fixture_db(path):
    rows = json.loads((GOLD / "ref_threads_rows.json").read_text())
    store = ThreadStore(str(path))
    con = store._db()
    con.executemany("INSERT INTO threads (id, created_at, metadata, sandbox_id) VALUES (?, ?, ?, ?)", rows["threads"])
    con.executemany("INSERT INTO messages (id, thread_id, message, metadata, created_at) VALUES (?, ?, ?, ?, ?)",
                    rows["messages"])
    con.commit()
    return store


def test_history_assembly_bit_exact_with_reference(tmp_path):
    gold = json.loads((GOLD / "history_assembly.json").read_text())
    store = _rebuild_fixture_db(tmp_path / "threads.db")
    loaded = kept = 0
    all_dicts = []
    for tid in sorted(gold["threads"]):
        msgs = asyncio.run(store.get_thread_messages(tid))
        san = sanitize_messages(msgs)
        dicts = [m.to_dict() for m in san]
        assert {"loaded": len(msgs), "kept": len(san), "sha256": digest(dicts)} == gold["threads"][tid], tid
        loaded += len(msgs)
        kept += len(san)
        all_dicts.append(dicts)
    assert (loaded, kept) == (gold["loaded"], gold["kept"]) == (201, 194)
    assert digest(all_dicts) == gold["sha256_all"]


@pytest.mark.skipif(not ref_harness.available(), reason="/root/reference not mounted")
def test_store_and_reference_client_read_each_others_files(tmp_path):
    ref_harness.load()
    from src.db.local import LocalDBClient
    from src.llm.types import Message as RefMessage

    async def go():
        path = str(tmp_path / "t.db")
        mine, theirs = ThreadStore(path), LocalDBClient(path)
        await mine.create_thread("T", user_id="u1")
        await mine.add_message("T", Message(role="user", content="hi"))
        await theirs.add_message("T", RefMessage(role="assistant", content=None,
                                                 tool_calls=[{"id": "c", "function": {"name": "f", "arguments": "{}"}}]))
        await mine.add_message("T", Message(role="tool", content="ok", tool_call_id="c", name="f"))
        a = [m.to_dict() for m in await mine.get_thread_messages("T")]
        b = [m.to_dict() for m in await theirs.get_thread_messages("T")]
        assert a == b and [m["role"] for m in a] == ["user", "assistant", "tool"]
        assert await theirs.thread_exists("T")
        # identical stored documents
        assert message_to_json(Message(role="user", content=None)) == '{"role": "user", "content": null}'
    asyncio.run(go())


def test_history_order_survives_same_second_timestamps(tmp_path):
    """The reference orders by a 1-second timestamp only; rowid tie-break = insertion order."""
    store = ThreadStore(str(tmp_path / "x.db"))

    async def go():
        await store.create_thread("T")
        for i in range(20):
            await store.add_message("T", Message(role="user", content=str(i)))
        con = store._db()
        con.execute("UPDATE messages SET created_at = '2025-01-01 00:00:00'")
        con.commit()
        got = [m.content for m in await store.get_thread_messages("T")]
        assert got == [str(i) for i in range(20)]
    asyncio.run(go())


def test_assemble_turn_persists_and_sanitises(tmp_path):
    store = ThreadStore(str(tmp_path / "y.db"))

    async def go():
        await store.create_thread("T")
        await store.add_message("T", Message(role="tool", content="orphan", tool_call_id="nope"))
        out = await assemble_turn(store, "T", [{"role": "user", "content": "hello"}])
        assert [m.role for m in out] == ["user"]
        assert [m.role for m in await store.get_thread_messages("T")] == ["tool", "user"]
        out2 = await assemble_turn(store, "NEW", [{"role": "user", "content": "x"}], save_new=False)
        assert await store.thread_exists("NEW") and [m.content for m in out2] == ["x"]
        assert await store.get_thread_messages("NEW") == []
    asyncio.run(go())


def test_kv_residency_table(tmp_path):
    store = ThreadStore(str(tmp_path / "z.db"))
    rows = [{"thread_id": "a", "entry_id": 1, "n_tokens": 130, "prefix_hash": 0xABCDEF, "block_ids": [5, 9, 2],
             "last_used": 12.5},
            {"thread_id": None, "entry_id": 2, "n_tokens": 10, "prefix_hash": 0, "block_ids": [1], "last_used": 1.0}]

    async def go():
        assert await store.record_residency(rows, gpu=3) == 1
        r = await store.get_residency("a")
        assert r == {"thread_id": "a", "gpu": 3, "n_tokens": 130, "prefix_hash": "0000000000abcdef",
                     "block_ids": [5, 9, 2], "last_used": 12.5}
        assert await store.get_residency("missing") is None
        await store.record_residency([], gpu=3)
        assert await store.get_residency("a") is None
    asyncio.run(go())


# ---- agent loop: same event stream as the reference's Agent -------------------------------------
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


def _normalise(ev):
    ev = json.loads(json.dumps(ev))
    if str(ev.get("id", "")).startswith("chatcmpl-"):
        ev["id"] = "chatcmpl-*"
    if "created" in ev:
        ev["created"] = 0
    return ev


def test_streaming_agent_reproduces_reference_event_stream():
    gold = json.loads((GOLD / "agent_events_echo.json").read_text())
    tools = LocalToolProvider([Tool("add", "add two ints",
                                    {"type": "object", "properties": {"a": {"type": "integer"}, "b": {"type": "integer"}},
                                     "required": ["a", "b"]}, handler=lambda a, b: str(a + b))])
    prov = Scripted()
    agent = StreamingAgent(prov, tools, system_prompt="SYS")

    async def go():
        return [_normalise(ev) async for ev in agent.run([Message(role="user", content="2+3?")], model="m",
                                                         temperature=0.0, max_tokens=16)]
    events = asyncio.run(go())
    assert events == gold["events"]
    assert prov.calls == gold["provider_calls"]


def test_streaming_agent_yields_before_the_provider_finishes():
    """The reference buffers the provider stream (agents/base.py:231-233); the twin must not."""
    order = []

    class Slow(LLMProvider):
        async def stream_completion(self, messages, **kw):
            for i in range(3):
                order.append(f"produced{i}")
                yield StreamChunk(content=str(i))
            yield StreamChunk(finish_reason="stop")

        async def completion(self, messages, **kw):
            return CompletionResponse()

    async def go():
        async for ev in StreamingAgent(Slow()).run([Message(role="user", content="x")]):
            if ev.get("choices") and ev["choices"][0]["delta"].get("content"):
                order.append("consumed" + ev["choices"][0]["delta"]["content"])
    asyncio.run(go())
    assert order[:4] == ["produced0", "consumed0", "produced1", "consumed1"]


def test_streaming_agent_idle_and_compaction():
    class Idle(LLMProvider):
        async def stream_completion(self, messages, **kw):
            yield StreamChunk(tool_calls=[{"index": 0, "id": "i1", "function": {"name": "idle",
                                                                                "arguments": "{\"summary\": \"done\"}"}}])
            yield StreamChunk(finish_reason="tool_calls")

        async def completion(self, messages, **kw):
            return CompletionResponse()

    async def run(agent):
        return [ev async for ev in agent.run([Message(role="user", content="x")])]
    evs = asyncio.run(run(StreamingAgent(Idle())))
    assert evs[-1] == {"type": "agent_done", "reason": "idle", "summary": "done", "iteration": 0}
    assert evs[-2]["delta"] == json.dumps({"status": "idle", "summary": "done"}) and evs[-2]["is_complete"]

    class Overflow(LLMProvider):
        def __init__(self):
            super().__init__()
            self.n = 0

        async def stream_completion(self, messages, **kw):
            self.n += 1
            if self.n == 1:
                raise LLMProviderError("This model's maximum context length is 8 tokens", provider="B200")
            yield StreamChunk(content=f"{len(messages)} msgs")
            yield StreamChunk(finish_reason="stop")

        async def completion(self, messages, **kw):
            return CompletionResponse()

    class KeepLast:
        async def compact(self, messages, system_prompt, model):
            return messages[-1:]
    evs = asyncio.run(run(StreamingAgent(Overflow(), context_compaction_provider=KeepLast(), system_prompt="S")))
    assert evs[-1]["final_content"] == "1 msgs"
    with pytest.raises(LLMProviderError):
        asyncio.run(run(StreamingAgent(Overflow())))              # no compaction provider: error propagates


# ---- provider over a fake engine ----------------------------------------------------------------
@pytest.fixture
def provider():
    tok = ByteTokenizer(512)
    eng = Engine(FakeStepper(eos_id=tok.eot_id, max_pos=1024, max_tokens=512))
    yield B200LLMProvider(eng, tok, model="tiny-test", default_temperature=0.0, default_max_tokens=16)
    eng.shutdown()


def test_provider_chunk_conventions(provider):
    msgs = [Message(role="system", content="be brief"), Message(role="user", content="hello")]

    async def go():
        return [c async for c in provider.stream_completion(msgs, max_tokens=6, model="b200/tiny-test")]
    chunks = asyncio.run(go())
    assert chunks[0].role == "assistant" and chunks[0].content is None
    assert chunks[-1].finish_reason == "length" and chunks[-1].content is None
    assert all(c.id == chunks[0].id and c.id.startswith("chatcmpl-") and c.model == "b200/tiny-test" for c in chunks)
    body = chunks[1:-1]
    assert len(body) == 6 and all(c.content and c.role is None and c.finish_reason is None for c in body)
    prompt = provider.render(msgs, None)
    want = expected_tokens(prompt, 6, provider.tokenizer.eot_id)
    assert "".join(c.content for c in body) == provider.tokenizer.decode(want)


def test_provider_completion_usage_and_thread_reuse(provider):
    async def go():
        m1 = [Message(role="user", content="first question")]
        r1 = await provider.completion(m1, max_tokens=8, thread_id="T1")
        assert r1.finish_reason == "length" and r1.usage["completion_tokens"] == 8 and r1.usage["cached_tokens"] == 0
        assert r1.usage["total_tokens"] == r1.usage["prompt_tokens"] + 8
        m2 = m1 + [r1.to_message(), Message(role="user", content="second")]
        r2 = await provider.completion(m2, max_tokens=8, thread_id="T1")
        # everything the first turn computed (prompt + 7 fed-back tokens) is served from resident KV
        assert r2.usage["cached_tokens"] == r1.usage["prompt_tokens"] + 7
        return r1, r2
    asyncio.run(go())


def test_provider_stop_sequences_and_errors(provider):
    msgs = [Message(role="user", content="stop test")]

    async def full():
        return (await provider.completion(msgs, max_tokens=12)).content
    text = asyncio.run(full())
    stop = text[5:7]

    async def stopped():
        return await provider.completion(msgs, max_tokens=12, stop=[stop])
    r = asyncio.run(stopped())
    assert r.finish_reason == "stop" and r.content == text[:text.find(stop)]

    async def empty():
        async for _ in provider.stream_completion([]):
            pass
    with pytest.raises(ValueError):
        asyncio.run(empty())

    async def too_long():
        async for _ in provider.stream_completion([Message(role="user", content="x" * 2000)]):
            pass
    with pytest.raises(LLMProviderError) as ei:
        asyncio.run(too_long())
    assert "maximum context length" in str(ei.value) and ei.value.provider == "B200"


def test_provider_abandoned_stream_releases_slot(provider):
    async def go():
        agen = provider.stream_completion([Message(role="user", content="bye")], max_tokens=400)
        n = 0
        async for _ in agen:
            n += 1
            if n == 3:
                break
        await agen.aclose()
        for _ in range(200):
            if not provider.engine.sched.running:
                break
            await asyncio.sleep(0.01)
        assert not provider.engine.sched.running
    asyncio.run(go())


@pytest.mark.skipif(not ref_harness.available(), reason="/root/reference not mounted")
def test_provider_runs_inside_the_reference_agent(provider):
    """Drop-in check: the reference's own Agent.run drives this provider unchanged."""
    ref_harness.load()
    from src.agents.base import Agent
    from src.llm.types import Message as RefMessage
    from src.tools.agent import AgentToolProvider

    async def go():
        tools = AgentToolProvider()
        await tools.connect()
        agent = Agent(llm_provider=provider, tool_provider=tools, system_prompt="SYS")
        return [ev async for ev in agent.run([RefMessage(role="user", content="hi")], model="tiny-test",
                                             temperature=0.0, max_tokens=5)]
    import contextlib
    import io
    with contextlib.redirect_stdout(io.StringIO()):
        evs = asyncio.run(go())
    assert evs[-1]["type"] == "agent_done" and evs[-1]["reason"] == "text_response"
    assert len(evs[-1]["final_content"]) == 5
    assert evs[0]["choices"][0]["delta"] == {"role": "assistant"}
    assert evs[-2]["choices"][0]["finish_reason"] == "length"


def test_echo_provider():
    async def go():
        p = EchoLLMProvider()
        r = await p.completion([Message(role="user", content="ping"), Message(role="assistant", content="x")])
        assert r.content == "ping" and r.finish_reason == "stop"
    asyncio.run(go())


def test_config4_scripted_tool_rounds_prefill_incrementally():
    """BASELINE config 4 shape: two tool rounds then text, through the streaming agent. Each round's
    prompt extends the previous one, so after round 1 almost everything is served from resident KV."""
    from kafka_b200.provider import ScriptedToolCallProvider
    tok = ByteTokenizer(512)
    eng = Engine(FakeStepper(eos_id=tok.eot_id, n_blocks=128, max_pos=4096, max_tokens=2048))
    provider = B200LLMProvider(eng, tok, model="tiny-test", default_temperature=0.0)
    try:
        _config4_body(provider)
    finally:
        eng.shutdown()


def _config4_body(provider):
    from kafka_b200.provider import ScriptedToolCallProvider

    async def lookup(key: str):
        return "v" * 64                                           # a 64-token tool result

    tools = LocalToolProvider([Tool("lookup", "look something up",
                                    {"type": "object", "properties": {"key": {"type": "string"}}}, lookup)])
    scripted = ScriptedToolCallProvider(provider, "lookup", {"key": "k"}, tool_rounds=2, think_tokens=8)
    agent = StreamingAgent(scripted, tools, system_prompt="SYS")
    st = provider.engine.stepper

    async def go():
        return [ev async for ev in agent.run([Message(role="user", content="use the tool twice")], model="tiny-test",
                                             temperature=0.0, max_tokens=24, thread_id="agent-1")]
    events = asyncio.run(go())
    finishes = [e["choices"][0]["finish_reason"] for e in events if e.get("choices") and e["choices"][0]["finish_reason"]]
    assert finishes == ["tool_calls", "tool_calls", "length"]
    results = [e for e in events if e.get("type") == "tool_result" and e["is_complete"]]
    assert len(results) == 2 and events[-1]["type"] == "agent_done" and len(events[-1]["final_content"]) == 24
    # 3 provider calls; the 2nd and 3rd found the earlier rounds resident
    assert provider.engine.sched.stats["requests"] == 3
    cached, prompt = provider.engine.sched.stats["cached_tokens"], provider.engine.sched.stats["prompt_tokens"]
    assert cached / prompt > 0.45 and st.prefilled_tokens == prompt - cached


def test_render_cache_is_invisible():
    """template.RenderCache (tokens of the messages a thread's previous call already rendered are reused): whatever
    happens to a thread's conversation between calls — messages appended, the tail dropped, a message in the middle
    edited, tool calls, a different tool list, another vocabulary under the same thread id — the cached render is
    the same token list as a render from scratch."""
    import random
    from kafka_b200.template import RenderCache
    rnd = random.Random(3)
    toks = [ByteTokenizer(512), ByteTokenizer(32000)]
    tmpls = [ChatTemplate(t) for t in toks]
    tool_sets = [None, [{"type": "function", "function": {"name": "f", "parameters": {"a": 1}}}],
                 [{"type": "function", "function": {"name": "g", "parameters": {}}}]]

    def msg():
        r = rnd.random()
        text = "".join(chr(rnd.randrange(32, 127)) for _ in range(rnd.randrange(0, 40)))
        if r < 0.15:
            return Message(role="assistant", content=text or None,
                           tool_calls=[{"id": f"c{rnd.randrange(9)}", "function": {"name": "f", "arguments": "{}"}}])
        if r < 0.3:
            return {"role": "tool", "content": text, "tool_call_id": f"c{rnd.randrange(9)}", "name": "f"}
        return Message(role=rnd.choice(["user", "assistant", "system"]), content=text)
    RenderCache.clear()
    convs = {f"t{i}": [msg() for _ in range(rnd.randrange(1, 6))] for i in range(4)}
    hits = 0
    for step in range(600):
        tid = rnd.choice(sorted(convs))
        c = convs[tid]
        r = rnd.random()
        if r < 0.5:
            c = c + [msg() for _ in range(rnd.randrange(1, 4))]
        elif r < 0.65 and len(c) > 1:
            c = c[:rnd.randrange(1, len(c))]
        elif r < 0.8 and c:
            i = rnd.randrange(len(c))
            c = c[:i] + [msg()] + c[i + 1:]
        if len(c) > 30:
            c = c[-3:]
        convs[tid] = c
        which = rnd.randrange(2)
        tools = rnd.choice(tool_sets)
        gen = rnd.random() < 0.7
        before = RenderCache._total
        got = tmpls[which].render(c, tools=tools, add_generation_prompt=gen, cache_key=tid)
        assert got == tmpls[which].render(c, tools=tools, add_generation_prompt=gen), (step, tid)
        got.append(-1)                                       # callers may mutate the list they get
        hits += RenderCache._total >= before
    assert RenderCache._total == sum(len(e[3]) for e in RenderCache._entries.values()) and hits > 100
    # bounded: least recently used threads go first
    old_max = RenderCache.max_tokens
    try:
        RenderCache.max_tokens = 200
        for i in range(20):
            tmpls[0].render([Message(role="user", content="x" * 50)], cache_key=f"lru{i}")
        assert RenderCache._total <= 200 + 80 and ("byte", 512), "lru19" in RenderCache._entries
        assert (("byte", 512), "lru0") not in RenderCache._entries
    finally:
        RenderCache.max_tokens = old_max
        RenderCache.clear()
