"""HTTP front of the completion path — same routes and wire shapes as the reference server
(/root/reference/server.py:384-620), backed by the local provider and the SQLite thread store.

Routes (identical paths / request models, /root/reference/src/kafka/types.py):
  POST /v1/threads/{id}/chat/completions   thread history + new messages -> completion (SSE or JSON)
  POST /v1/chat/completions                stateless
  POST /v1/agent/run, POST /v1/threads/{id}/agent/run     raw agent event stream (SSE)
  POST/GET/DELETE /v1/threads/{id}/messages, POST /v1/threads, GET /v1/models, GET /health

SSE framing is the reference's: `data: {json}\n\n` per event and a closing `data: [DONE]\n\n`
(:195-197); chat chunks are `chat.completion.chunk` objects with null fields kept (pydantic dump of
StreamChunkResponse, :338-356); failures become `{"error": {"message", "type"}}` + [DONE] (:375-377).

Differences, all deliberate (§8f):
  * token-by-token streaming: content deltas are forwarded as the provider yields them. The reference
    waits for the whole agent loop and then re-slices the final text in 20-character pieces
    (:289-356), so its HTTP time-to-first-token equals the total generation time.
  * `temperature: 0` is honoured (the reference's `request.temperature or 0.7` turns it into 0.7, :292).
  * tool rounds of /chat/completions are persisted (the reference listens for events nothing emits,
    :309-324), so stored history == model-visible history and the KV prefix stays valid.
  * threads live in the local SQLite store; the server starts offline (the reference's lifespan needs
    Supabase credentials, :97).
  * non-streaming responses carry the real token usage (the reference returns zeros, :452).
"""
from __future__ import annotations

import json
import time
import uuid
from typing import Any, AsyncGenerator, Callable, Dict, List, Optional

from fastapi import FastAPI, HTTPException
from fastapi.responses import StreamingResponse
from pydantic import BaseModel, Field

from .agent import StreamingAgent
from .db import ThreadStore
from .history import sanitize_messages, to_message
from .orchestrator import ThreadRunner
from .tools import LocalToolProvider, Tool


# ---- request / response models (field-for-field the reference's kafka/types.py) -------------------
class ChatMessage(BaseModel):
    role: str
    content: Optional[str] = None
    name: Optional[str] = None
    tool_calls: Optional[List[Dict[str, Any]]] = None
    tool_call_id: Optional[str] = None


class ChatCompletionRequest(BaseModel):
    model: str
    messages: List[ChatMessage]
    temperature: Optional[float] = Field(None, ge=0, le=2)
    max_tokens: Optional[int] = Field(None, gt=0)
    stream: Optional[bool] = False
    stop: Optional[List[str]] = None
    top_p: Optional[float] = Field(None, ge=0, le=1)
    frequency_penalty: Optional[float] = Field(None, ge=-2, le=2)
    presence_penalty: Optional[float] = Field(None, ge=-2, le=2)
    user: Optional[str] = None


class AgentRunRequest(BaseModel):
    messages: List[ChatMessage]
    model: str = "gpt-4o"
    temperature: float = 0.7
    max_tokens: Optional[int] = None


class CreateThreadRequest(BaseModel):
    system_message: Optional[str] = None
    user_id: Optional[str] = None
    kafka_profile_id: Optional[str] = None
    metadata: Optional[Dict[str, Any]] = None


class DeltaContent(BaseModel):
    role: Optional[str] = None
    content: Optional[str] = None
    tool_calls: Optional[List[Dict[str, Any]]] = None


class StreamChoice(BaseModel):
    index: int = 0
    delta: DeltaContent
    finish_reason: Optional[str] = None


class StreamChunkResponse(BaseModel):
    id: str
    object: str = "chat.completion.chunk"
    created: int
    model: str
    choices: List[StreamChoice]


def _sse(obj: Any) -> str:
    return f"data: {obj if isinstance(obj, str) else json.dumps(obj)}\n\n"


def create_app(provider_factory: Callable[..., Any], store: Optional[ThreadStore] = None,
               tools: Optional[List[Tool]] = None, system_prompt: Optional[str] = None,
               models: Optional[List[str]] = None) -> FastAPI:
    """provider_factory(model, tool_provider, thread_id) -> LLMProvider  (kafka_b200.factory.make_provider)."""
    app = FastAPI(title="kafka-b200")
    store = store or ThreadStore()
    models = models or ["llama-3-8b", "tinyllama-1.1b", "echo"]

    def make_agent(model: str, thread_id: Optional[str], top_p: Optional[float] = None) -> StreamingAgent:
        tool_provider = LocalToolProvider(list(tools or []))
        # the request's `top_p` becomes the per-request provider's default (the reference accepts the field,
        # src/kafka/types.py:35, and drops it; here the sampling kernel applies it). Only passed when given, so that
        # provider factories without the argument keep working.
        extra = {"default_top_p": top_p} if top_p is not None and 0.0 < top_p < 1.0 else {}
        provider = provider_factory(model, tool_provider=tool_provider, thread_id=thread_id, **extra)
        # context overflow -> offline compaction (truncation at safe points until the prompt fits the model's
        # context, kafka_b200/compaction.py) -> one retry; the reference builds a summarising provider here,
        # which needs a remote LLM (src/kafka/v1.py:185-194)
        compaction = provider.compaction_provider() if hasattr(provider, "compaction_provider") else None
        return StreamingAgent(provider, tool_provider, system_prompt=system_prompt,
                              context_compaction_provider=compaction)

    def make_runner(model: str, thread_id: str, top_p: Optional[float] = None) -> ThreadRunner:
        agent = make_agent(model, thread_id, top_p)
        prov = agent.llm_provider
        engine = getattr(prov, "engine", None) or getattr(getattr(prov, "inner", None), "engine", None)
        return ThreadRunner(agent, store, thread_id, engine=engine)   # engine: feeds the kv_residency table

    def temperature_of(req: ChatCompletionRequest) -> float:
        return 0.7 if req.temperature is None else req.temperature

    async def completion_stream(events: AsyncGenerator[Dict[str, Any], None], model: str) -> AsyncGenerator[str, None]:
        cid, created = f"chatcmpl-{uuid.uuid4().hex[:24]}", int(time.time())

        def chunk(delta: DeltaContent, finish: Optional[str] = None) -> str:
            return _sse(StreamChunkResponse(id=cid, created=created, model=model,
                                            choices=[StreamChoice(delta=delta, finish_reason=finish)]).model_dump_json())
        try:
            sent_role = False
            async for ev in events:
                if ev.get("type") == "tool_result":
                    yield _sse({"type": "tool_result", "tool_call_id": ev.get("tool_call_id"),
                                "tool_name": ev.get("tool_name"), "delta": ev.get("delta", ""),
                                "is_complete": ev.get("is_complete", False)})
                elif ev.get("choices"):
                    text = ev["choices"][0].get("delta", {}).get("content")
                    if text:
                        if not sent_role:
                            yield chunk(DeltaContent(role="assistant"))
                            sent_role = True
                        yield chunk(DeltaContent(content=text))
            if not sent_role:
                yield chunk(DeltaContent(role="assistant"))
            yield chunk(DeltaContent(), "stop")
            yield _sse("[DONE]")
        except Exception as e:
            yield _sse({"error": {"message": str(e), "type": "server_error"}})
            yield _sse("[DONE]")

    async def agent_stream(events: AsyncGenerator[Dict[str, Any], None]) -> AsyncGenerator[str, None]:
        try:
            async for ev in events:
                yield _sse(ev)
            yield _sse("[DONE]")
        except Exception as e:
            yield _sse({"error": {"message": str(e), "type": "agent_error"}})
            yield _sse("[DONE]")

    async def collect(events: AsyncGenerator[Dict[str, Any], None], model: str, provider: Any = None) -> Dict[str, Any]:
        final, finish, usage = "", "stop", None
        async for ev in events:
            if ev.get("type") == "agent_done":
                final = ev.get("final_content", "") or ""
            elif ev.get("choices") and ev["choices"][0].get("finish_reason") in ("stop", "length"):
                finish = ev["choices"][0]["finish_reason"]
        total = getattr(provider, "usage_total", None)
        if total:                                    # the request's real token counts, tool rounds included
            usage = {k: int(total.get(k, 0)) for k in ("prompt_tokens", "completion_tokens", "total_tokens")}
        return {"id": f"chatcmpl-{uuid.uuid4().hex[:24]}", "object": "chat.completion", "created": int(time.time()),
                "model": model, "choices": [{"index": 0, "message": {"role": "assistant", "content": final},
                                             "finish_reason": finish}],
                "usage": usage or {"prompt_tokens": 0, "completion_tokens": 0, "total_tokens": 0}}

    # ---- completions -----------------------------------------------------------------------------
    @app.post("/v1/threads/{thread_id}/chat/completions")
    async def chat_completions(thread_id: str, request: ChatCompletionRequest):
        runner = make_runner(request.model, thread_id, request.top_p)
        events = runner.run(request.messages, model=request.model, temperature=temperature_of(request),
                            max_tokens=request.max_tokens)
        if request.stream:
            return StreamingResponse(completion_stream(events, request.model), media_type="text/event-stream",
                                     headers={"Cache-Control": "no-cache", "Connection": "keep-alive",
                                              "X-Accel-Buffering": "no"})
        try:
            return await collect(events, request.model, runner.agent.llm_provider)
        except Exception as e:
            raise HTTPException(status_code=500, detail=str(e))

    @app.post("/v1/chat/completions")
    async def chat_completions_standard(request: ChatCompletionRequest):
        agent = make_agent(request.model, None, request.top_p)
        msgs = sanitize_messages([to_message(m) for m in request.messages])
        events = agent.run(msgs, model=request.model, temperature=temperature_of(request),
                           max_tokens=request.max_tokens)
        if request.stream:
            return StreamingResponse(completion_stream(events, request.model), media_type="text/event-stream")
        try:
            return await collect(events, request.model, agent.llm_provider)
        except Exception as e:
            raise HTTPException(status_code=500, detail=str(e))

    @app.post("/v1/agent/run")
    async def run_agent(request: AgentRunRequest):
        agent = make_agent(request.model, None)
        msgs = [to_message(m) for m in request.messages]
        return StreamingResponse(agent_stream(agent.run(msgs, model=request.model, temperature=request.temperature,
                                                        max_tokens=request.max_tokens)),
                                 media_type="text/event-stream")

    @app.post("/v1/threads/{thread_id}/agent/run")
    async def run_agent_with_thread(thread_id: str, request: AgentRunRequest):
        runner = make_runner(request.model, thread_id)
        return StreamingResponse(agent_stream(runner.run(request.messages, model=request.model,
                                                         temperature=request.temperature,
                                                         max_tokens=request.max_tokens)),
                                 media_type="text/event-stream")

    # ---- threads ---------------------------------------------------------------------------------
    @app.post("/v1/threads/{thread_id}/messages")
    async def add_message_to_thread(thread_id: str, message: ChatMessage):
        if not await store.thread_exists(thread_id):
            await store.create_thread(thread_id=thread_id)
        res = await store.add_message(thread_id, to_message(message))
        return {"status": "ok", "message_id": res["id"], "thread_id": thread_id}

    @app.get("/v1/threads/{thread_id}/messages")
    async def get_thread_messages(thread_id: str):
        msgs = await store.get_thread_messages(thread_id)
        return {"thread_id": thread_id, "messages": [m.to_dict() for m in msgs]}

    @app.post("/v1/threads")
    async def create_thread(request: Optional[CreateThreadRequest] = None):
        r = request or CreateThreadRequest()
        t = await store.create_thread(metadata=r.metadata, system_message=r.system_message, user_id=r.user_id,
                                      kafka_profile_id=r.kafka_profile_id)
        return {"thread_id": t["id"], "created_at": t["created_at"]}

    @app.delete("/v1/threads/{thread_id}/messages")
    async def clear_thread(thread_id: str):
        n = await store.delete_thread_messages(thread_id)
        return {"status": "ok", "thread_id": thread_id, "deleted": n}

    @app.get("/v1/threads/{thread_id}/kv")
    async def thread_residency(thread_id: str):
        """Where this thread's KV is resident (new; the reference has no such notion)."""
        row = await store.get_residency(thread_id)
        return {"thread_id": thread_id, "resident": row is not None, "residency": row}

    @app.get("/v1/models")
    async def list_models():
        return {"object": "list", "data": [{"id": m, "object": "model", "owned_by": "kafka-b200"} for m in models]}

    @app.get("/health")
    async def health_check():
        return {"status": "healthy", "kafka_initialized": True}

    app.state.store = store
    return app


def main(argv: Optional[List[str]] = None) -> None:
    """`python -m kafka_b200.server`: the app on uvicorn, the way the reference starts its own
    (/root/reference/server.py:627-631: 0.0.0.0, $PORT, default 8081), with the process-wide engines of
    kafka_b200.factory behind it. Models: the built-in shapes (seeded random weights) plus every checkpoint
    named in $KAFKA_B200_CHECKPOINTS (real weights and tokenizer); threads in $LOCAL_DB_PATH (data/threads.db)."""
    import argparse
    import os
    import uvicorn
    from . import checkpoint, factory
    from .model_config import REGISTRY
    ap = argparse.ArgumentParser(prog="python -m kafka_b200.server")
    ap.add_argument("--host", default="0.0.0.0")
    ap.add_argument("--port", type=int, default=int(os.environ.get("PORT", 8081)))
    ap.add_argument("--db", default=None, help="SQLite thread store (default: $LOCAL_DB_PATH or data/threads.db)")
    ap.add_argument("--system-prompt", default=None)
    args = ap.parse_args(argv)
    checkpoint.register_from_env()
    models = sorted(checkpoint.CHECKPOINTS) + [m for m in sorted(REGISTRY) if m not in checkpoint.CHECKPOINTS] + ["echo"]
    app = create_app(factory.make_provider, store=ThreadStore(args.db), system_prompt=args.system_prompt, models=models)
    try:
        uvicorn.run(app, host=args.host, port=args.port)
    finally:
        factory.shutdown_engines()


if __name__ == "__main__":
    main()
