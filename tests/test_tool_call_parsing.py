"""Model-emitted tool calls: text in the syntax the chat template renders tool
calls in is turned into StreamChunk.tool_calls deltas with finish_reason "tool_calls", the shape the
reference's agent consumes (/root/reference/src/agents/base.py:286-334). A scripted stepper plays the
model, since random-init weights never produce the syntax."""
import asyncio
import json

from fakes import FakeStepper
from kafka_b200.agent import StreamingAgent
from kafka_b200.engine import Engine
from kafka_b200.llm import Message
from kafka_b200.provider import B200LLMProvider, parse_tool_calls
from kafka_b200.tokenizer import ByteTokenizer
from kafka_b200.tools import LocalToolProvider, Tool


class ScriptStepper(FakeStepper):
    """Emits the bytes of `scripts[k]` for the k-th request, then EOS."""

    def __init__(self, scripts, **kw):
        super().__init__(**kw)
        self.scripts = [list(s.encode()) for s in scripts]
        self.by_req = {}

    def _next(self, req):
        if req.req_id not in self.by_req:
            self.by_req[req.req_id] = self.scripts[len(self.by_req) % len(self.scripts)]
        s = self.by_req[req.req_id]
        i = len(req.generated) + req.inflight   # ids already read back + ids still in flight
        return s[i] if i < len(s) else self.eos_id

    def prefill_launch(self, items, parity):
        return [self._next(it.req) if it.is_last_chunk else None for it in items]

    def decode_launch(self, reqs, parity):
        return [self._next(r) for r in reqs]

    def fetch(self, handle):
        return handle


def _provider(scripts, tools):
    tok = ByteTokenizer(512)
    eng = Engine(ScriptStepper(scripts, eos_id=tok.eot_id, n_blocks=256, max_pos=4096, max_tokens=2048))
    return eng, B200LLMProvider(eng, tok, model="tiny-test", tool_provider=tools, default_temperature=0.0,
                                default_max_tokens=400)


def test_parse_tool_calls_shapes():
    full = parse_tool_calls('[{"id":"c1","type":"function","function":{"name":"add","arguments":"{\\"a\\":1}"}}]')
    assert full == [{"id": "c1", "type": "function", "function": {"name": "add", "arguments": '{"a":1}'}}]
    short = parse_tool_calls('{"name": "add", "arguments": {"a": 1, "b": 2}}')
    assert short[0]["function"] == {"name": "add", "arguments": '{"a": 1, "b": 2}'} and short[0]["id"].startswith("call_")
    assert parse_tool_calls("not json") is None and parse_tool_calls("[]") is None and parse_tool_calls('[{"x":1}]') is None


def test_stream_turns_tool_call_text_into_deltas():
    tools = LocalToolProvider([Tool("add", "add", {"type": "object", "properties": {}}, lambda a, b: str(a + b))])
    call = '[{"id":"call_7","type":"function","function":{"name":"add","arguments":"{\\"a\\": 2, \\"b\\": 3}"}}]'
    eng, prov = _provider(["Let me add.\n<tool_calls>" + call + "</tool_calls>ignored tail"], tools)
    try:
        async def go():
            return [c async for c in prov.stream_completion([Message(role="user", content="2+3?")])]
        chunks = asyncio.run(go())
        text = "".join(c.content or "" for c in chunks)
        assert text == "Let me add.\n" and "<tool_calls>" not in text
        assert chunks[-1].finish_reason == "tool_calls"
        deltas = [tc for c in chunks for tc in (c.tool_calls or [])]
        assert deltas[0] == {"index": 0, "id": "call_7", "type": "function", "function": {"name": "add", "arguments": ""}}
        assert "".join(d["function"].get("arguments", "") for d in deltas) == '{"a": 2, "b": 3}'
        r = asyncio.run(prov.completion([Message(role="user", content="again")]))
        assert r.finish_reason == "tool_calls" and r.tool_calls[0]["function"]["name"] == "add"
    finally:
        eng.shutdown()


def test_agent_executes_a_model_emitted_tool_call_and_continues():
    tools = LocalToolProvider([Tool("add", "add", {"type": "object", "properties": {}}, lambda a, b: str(a + b))])
    eng, prov = _provider(['<tool_calls>{"name": "add", "arguments": {"a": 2, "b": 3}}</tool_calls>', "It is 5."], tools)
    try:
        agent = StreamingAgent(prov, tools)

        async def go():
            return [e async for e in agent.run([Message(role="user", content="2+3?")], model="tiny-test", temperature=0.0)]
        evs = asyncio.run(go())
        results = [e for e in evs if e.get("type") == "tool_result" and not e["is_complete"]]
        assert [r["delta"] for r in results] == ["5"]
        assert evs[-1] == {"type": "agent_done", "reason": "text_response", "final_content": "It is 5.", "iteration": 1}
    finally:
        eng.shutdown()


def test_malformed_block_and_no_tools_fall_back_to_text():
    tools = LocalToolProvider([Tool("add", "add", {"type": "object", "properties": {}}, lambda a, b: str(a + b))])
    eng, prov = _provider(["x <tool_calls>oops</tool_calls> y"], tools)
    try:
        r = asyncio.run(prov.completion([Message(role="user", content="q")]))
        assert r.content == "x <tool_calls>oops</tool_calls> y" and r.finish_reason == "stop" and r.tool_calls is None
    finally:
        eng.shutdown()
    eng, prov = _provider(['<tool_calls>{"name":"add","arguments":{}}</tool_calls>'], None)   # no tools on offer
    try:
        r = asyncio.run(prov.completion([Message(role="user", content="q")]))
        assert r.content.startswith("<tool_calls>") and r.finish_reason == "stop"
    finally:
        eng.shutdown()
