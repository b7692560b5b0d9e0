"""HTTP front of the path on CPU: same routes and wire shapes as the reference server, token-by-token
SSE, thread persistence, and BASELINE config 1 (thread replay on SQLite with an echo provider)."""
import asyncio
import json
import random

import httpx
import pytest

from fakes import FakeStepper, expected_tokens
from kafka_b200.db import ThreadStore
from kafka_b200.engine import Engine
from kafka_b200.provider import B200LLMProvider, EchoLLMProvider
from kafka_b200.server import create_app
from kafka_b200.tokenizer import ByteTokenizer
from kafka_b200.tools import Tool


def _client(app):
    return httpx.AsyncClient(transport=httpx.ASGITransport(app=app), base_url="http://test")


def _sse_events(text):
    out = []
    for line in text.splitlines():
        if line.startswith("data: "):
            payload = line[len("data: "):]
            out.append(payload if payload == "[DONE]" else json.loads(payload))
    return out


@pytest.fixture
def echo_app(tmp_path):
    store = ThreadStore(str(tmp_path / "threads.db"))
    return create_app(lambda model, tool_provider=None, thread_id=None: EchoLLMProvider(model, tool_provider), store=store)


def test_routes_match_the_reference(echo_app):
    paths = {(m, r.path) for r in echo_app.routes for m in getattr(r, "methods", [])}
    for want in [("POST", "/v1/threads/{thread_id}/chat/completions"), ("POST", "/v1/chat/completions"),
                 ("POST", "/v1/agent/run"), ("POST", "/v1/threads/{thread_id}/agent/run"),
                 ("POST", "/v1/threads/{thread_id}/messages"), ("GET", "/v1/threads/{thread_id}/messages"),
                 ("POST", "/v1/threads"), ("DELETE", "/v1/threads/{thread_id}/messages"), ("GET", "/v1/models"),
                 ("GET", "/health")]:                              # /root/reference/server.py:384-620
        assert want in paths, want


def test_thread_chat_completion_json_and_persistence(echo_app):
    async def go():
        async with _client(echo_app) as c:
            assert (await c.get("/health")).json() == {"status": "healthy", "kafka_initialized": True}
            r = await c.post("/v1/threads/T1/chat/completions",
                             json={"model": "echo", "messages": [{"role": "user", "content": "hello"}]})
            body = r.json()
            assert r.status_code == 200 and body["object"] == "chat.completion"
            assert body["choices"][0]["message"] == {"role": "assistant", "content": "hello"}
            r = await c.post("/v1/threads/T1/chat/completions",
                             json={"model": "echo", "messages": [{"role": "user", "content": "again"}]})
            msgs = (await c.get("/v1/threads/T1/messages")).json()["messages"]
            assert [(m["role"], m["content"]) for m in msgs] == [("user", "hello"), ("assistant", "hello"),
                                                                 ("user", "again"), ("assistant", "again")]
            assert (await c.delete("/v1/threads/T1/messages")).json()["deleted"] == 4
            t = (await c.post("/v1/threads", json={"system_message": "SYS", "user_id": "u"})).json()
            msgs = (await c.get(f"/v1/threads/{t['thread_id']}/messages")).json()["messages"]
            assert msgs == [{"role": "system", "content": "SYS"}]
            assert (await c.post("/v1/chat/completions", json={"model": "echo", "messages": []})).status_code == 500
    asyncio.run(go())


@pytest.fixture
def fake_engine_app(tmp_path):
    tok = ByteTokenizer(512)
    eng = Engine(FakeStepper(eos_id=tok.eot_id, max_pos=2048, max_tokens=512))
    store = ThreadStore(str(tmp_path / "threads.db"))

    def factory(model, tool_provider=None, thread_id=None):
        return B200LLMProvider(eng, tok, model=model, tool_provider=tool_provider, thread_id=thread_id,
                               default_max_tokens=16)

    async def slow_count(n: int):
        for i in range(n):
            yield f"{i} "
    tools = [Tool("count", "count slowly", {"type": "object", "properties": {"n": {"type": "integer"}}}, slow_count)]
    app = create_app(factory, store=store, tools=tools, system_prompt="You are terse.")
    app.state.engine, app.state.tok = eng, tok
    yield app
    eng.shutdown()


def test_sse_streams_token_by_token_and_honours_temperature_zero(fake_engine_app):
    app = fake_engine_app

    async def go():
        async with _client(app) as c:
            r = await c.post("/v1/threads/S1/chat/completions",
                             json={"model": "tiny-test", "stream": True, "temperature": 0, "max_tokens": 9,
                                   "messages": [{"role": "user", "content": "stream please"}]})
            assert r.headers["content-type"].startswith("text/event-stream")
            evs = _sse_events(r.text)
            assert evs[-1] == "[DONE]"
            chunks = [e for e in evs[:-1] if e.get("object") == "chat.completion.chunk"]
            # wire shape of the reference's StreamChunkResponse, nulls included (server.py:338-356)
            assert set(chunks[0]) == {"id", "object", "created", "model", "choices"}
            assert chunks[0]["choices"][0] == {"index": 0, "delta": {"role": "assistant", "content": None,
                                                                     "tool_calls": None}, "finish_reason": None}
            assert chunks[-1]["choices"][0]["finish_reason"] == "stop"
            pieces = [ch["choices"][0]["delta"]["content"] for ch in chunks[1:-1]]
            assert len(pieces) == 9 and all(len(p) == 1 for p in pieces)     # one token per SSE chunk
            msgs = (await c.get("/v1/threads/S1/messages")).json()["messages"]
            assert msgs[-1] == {"role": "assistant", "content": "".join(pieces)}
            # what the model saw: system prompt + user message, rendered with the tool schemas
            from kafka_b200.llm import Message
            prov = B200LLMProvider(app.state.engine, app.state.tok, model="tiny-test")
            tools_schema = [{"type": "function", "function": {"name": "count", "description": "count slowly",
                                                              "parameters": {"type": "object", "properties": {
                                                                  "n": {"type": "integer"}}}}}]
            from kafka_b200.agent import IDLE_DESCRIPTION
            tools_schema.append({"type": "function", "function": {
                "name": "idle", "description": IDLE_DESCRIPTION,
                "parameters": {"type": "object", "properties": {"summary": {
                    "type": "string", "description": "Optional brief summary of what was accomplished"}},
                    "required": []}}})
            prompt = prov.render([Message(role="system", content="You are terse."),
                                  Message(role="user", content="stream please")], tools_schema)
            want = app.state.tok.decode(expected_tokens(prompt, 9, app.state.tok.eot_id))
            assert "".join(pieces) == want
    asyncio.run(go())


def test_agent_run_stream_with_thread(fake_engine_app):
    async def go():
        async with _client(fake_engine_app) as c:
            r = await c.post("/v1/threads/A1/agent/run",
                             json={"model": "tiny-test", "temperature": 0.0, "max_tokens": 5,
                                   "messages": [{"role": "user", "content": "go"}]})
            evs = _sse_events(r.text)
            assert evs[-1] == "[DONE]" and evs[-2]["type"] == "agent_done" and evs[-2]["reason"] == "text_response"
            assert evs[0]["choices"][0]["delta"] == {"role": "assistant"}
            text = "".join(e["choices"][0]["delta"].get("content", "") for e in evs if isinstance(e, dict) and e.get("choices"))
            assert text == evs[-2]["final_content"] and len(text) == 5
            msgs = (await c.get("/v1/threads/A1/messages")).json()["messages"]
            assert [m["role"] for m in msgs] == ["user", "assistant"] and msgs[1]["content"] == text
            kv = (await c.get("/v1/threads/A1/kv")).json()
            # the turn left the thread's KV resident and the server recorded it in kv_residency
            assert kv["resident"] and kv["residency"]["gpu"] == 0 and kv["residency"]["n_tokens"] > 300
            assert len(kv["residency"]["block_ids"]) == (kv["residency"]["n_tokens"] + 63) // 64
            assert (await c.get("/v1/threads/nope/kv")).json() == {"thread_id": "nope", "resident": False,
                                                                   "residency": None}
    asyncio.run(go())


def test_context_overflow_is_reported_as_sse_error(fake_engine_app):
    async def go():
        async with _client(fake_engine_app) as c:
            r = await c.post("/v1/threads/O1/chat/completions",
                             json={"model": "tiny-test", "stream": True,
                                   "messages": [{"role": "user", "content": "x" * 3000}]})
            evs = _sse_events(r.text)
            assert evs[-1] == "[DONE]" and "maximum context length" in evs[-2]["error"]["message"]
            assert evs[-2]["error"]["type"] == "server_error"
    asyncio.run(go())


def test_config1_thread_replay_echo(tmp_path):
    """BASELINE config 1: 64 synthetic threads x 8 turns through the server on a fresh SQLite file."""
    store = ThreadStore(str(tmp_path / "replay.db"))
    app = create_app(lambda model, tool_provider=None, thread_id=None: EchoLLMProvider(model, tool_provider), store=store)

    async def go():
        async with _client(app) as c:
            for turn in range(8):
                for i in range(64):
                    rnd = random.Random(1234 + i * 100 + turn)
                    text = "".join(chr(rnd.randrange(32, 127)) for _ in range(48))
                    r = await c.post(f"/v1/threads/bench-{i:04d}/chat/completions",
                                     json={"model": "echo", "messages": [{"role": "user", "content": text}]})
                    assert r.json()["choices"][0]["message"]["content"] == text
            msgs = await store.get_thread_messages("bench-0063")
            assert len(msgs) == 16 and [m.role for m in msgs] == ["user", "assistant"] * 8
            assert len(await store.list_threads()) == 64
    asyncio.run(go())


def test_non_streaming_response_carries_real_usage(fake_engine_app):
    async def go():
        async with _client(fake_engine_app) as c:
            r = await c.post("/v1/threads/U1/chat/completions",
                             json={"model": "tiny-test", "temperature": 0, "max_tokens": 7,
                                   "messages": [{"role": "user", "content": "count my tokens"}]})
            u = r.json()["usage"]
            assert u["completion_tokens"] == 7 and u["prompt_tokens"] > 20
            assert u["total_tokens"] == u["prompt_tokens"] + u["completion_tokens"]
    asyncio.run(go())


def test_thread_that_outgrew_the_context_is_compacted_and_answered(fake_engine_app):
    """A stored history longer than the model's context: the provider raises the context-length error, the
    server's agent compacts by truncation (kafka_b200/compaction.py) and the retry answers."""
    app = fake_engine_app

    async def go():
        async with _client(app) as c:
            msgs = []
            for i in range(30):
                msgs += [{"role": "user", "content": f"q{i} " + "x" * 40}, {"role": "assistant", "content": f"a{i} " + "y" * 40}]
            for m in msgs:
                assert (await c.post("/v1/threads/C1/messages", json=m)).status_code == 200
            r = await c.post("/v1/threads/C1/chat/completions",
                             json={"model": "tiny-test", "temperature": 0, "max_tokens": 6,
                                   "messages": [{"role": "user", "content": "still there?"}]})
            body = r.json()
            assert r.status_code == 200 and len(body["choices"][0]["message"]["content"]) == 6
            assert body["usage"]["prompt_tokens"] <= int(0.75 * 2048)
    asyncio.run(go())


def test_module_entry_point_serves_over_http(tmp_path):
    """`python -m kafka_b200.server` (the counterpart of the reference's `python server.py`, server.py:627-631):
    a real process, a real socket; the echo model keeps it GPU-free. Health, the model list (built-in shapes + echo)
    and one streamed thread completion whose SSE lines end in [DONE]."""
    import json
    import os
    import socket
    import subprocess
    import sys
    import time
    import urllib.request
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    pkg = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "kafka-llm-service_b200")
    env = dict(os.environ, PYTHONPATH=pkg + os.pathsep + os.environ.get("PYTHONPATH", ""))
    proc = subprocess.Popen([sys.executable, "-m", "kafka_b200.server", "--host", "127.0.0.1", "--port", str(port),
                             "--db", str(tmp_path / "threads.db")], env=env, stdout=subprocess.PIPE,
                            stderr=subprocess.STDOUT)
    base = f"http://127.0.0.1:{port}"
    try:
        for _ in range(200):
            try:
                with urllib.request.urlopen(base + "/health", timeout=1) as r:
                    assert json.load(r)["status"] == "healthy"
                break
            except OSError:
                assert proc.poll() is None, proc.stdout.read().decode()[-2000:]
                time.sleep(0.1)
        else:
            raise AssertionError("server did not come up")
        with urllib.request.urlopen(base + "/v1/models", timeout=5) as r:
            ids = [m["id"] for m in json.load(r)["data"]]
        assert "echo" in ids and "llama-3-8b" in ids
        req = urllib.request.Request(base + "/v1/threads/t-main/chat/completions", method="POST",
                                     headers={"Content-Type": "application/json"},
                                     data=json.dumps({"model": "echo", "stream": True,
                                                      "messages": [{"role": "user", "content": "ping 123"}]}).encode())
        with urllib.request.urlopen(req, timeout=20) as r:
            lines = [ln.decode().strip() for ln in r if ln.strip()]
        assert lines[-1] == "data: [DONE]"
        text = "".join(json.loads(ln[6:])["choices"][0]["delta"].get("content") or "" for ln in lines[:-1])
        assert "ping 123" in text
    finally:
        proc.terminate()
        try:
            proc.wait(timeout=10)
        except subprocess.TimeoutExpired:
            proc.kill()


def test_example_script_runs_on_the_echo_model():
    """examples/agent.py (the shape of the reference's examples/agent.py: two local tools, a stateless run and two
    turns on a thread) end to end with the echo model — no GPU."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "examples", "agent.py"), "--model", "echo"], cwd=root,
                         capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "What is the weather in Paris?" in out.stdout and "Which number did I mention?" in out.stdout
    assert "stored messages: 4" in out.stdout
