"""StreamingAgent — the agent loop around the provider, without the buffering.

Same contract and event stream as the reference's `Agent.run`
(/root/reference/src/agents/base.py:160-440): prepend the system prompt when the conversation does
not start with one (:199-202), fetch tools (:205), then up to `max_iterations` rounds of
"stream a completion -> merge tool-call deltas by index -> execute tools -> append results", emitting
OpenAI `chat.completion.chunk` dicts, `tool_result` events and a final `agent_done`. The `idle` tool is
added automatically and short-circuits the loop (:384-411).

The one intended difference: the reference collects the whole provider stream
into a list before yielding anything (:231-233), so its time-to-first-token equals the full generation
time. Here every StreamChunk is forwarded the moment the provider yields it; a context-length failure
can only occur before the first chunk (the prompt is checked at submission), so the compaction retry
of :234-271 keeps its semantics.

tests/test_agent.py replays the scripted conversation of tests/golden/agent_events_echo.json (recorded
from the reference's own Agent) and requires an identical event sequence.
"""
from __future__ import annotations

import json
import logging
import time
import uuid
from typing import Any, AsyncGenerator, Dict, List, Optional

from .llm.types import Message
from .tools import LocalToolProvider, Tool

IDLE_DESCRIPTION = ("Call this after using tools to signal you are done with your task. Only needed after tool "
                    "usage, not for simple text responses.")

CONTEXT_ERROR_MARKERS = ("context_length_exceeded", "maximum context length", "token limit", "too many tokens",
                         "prompt is too long", "input is too long")


def looks_like_context_length_error(e: Exception) -> bool:
    """Subset of /root/reference/src/llm/context_compaction/base.py:10-65 sufficient for local providers."""
    s = str(e).lower()
    return any(m in s for m in CONTEXT_ERROR_MARKERS)


class StreamingAgent:
    def __init__(self, llm_provider: Any, tool_provider: Optional[Any] = None, max_iterations: int = 50,
                 system_prompt: Optional[str] = None, context_compaction_provider: Optional[Any] = None,
                 logger: Optional[logging.Logger] = None):
        self.llm_provider = llm_provider
        self.tool_provider = tool_provider if tool_provider is not None else LocalToolProvider()
        self.max_iterations = max_iterations
        self.system_prompt = system_prompt
        self.context_compaction_provider = context_compaction_provider
        self.logger = logger or logging.getLogger(__name__)
        self.tool_provider.add_tool(Tool(
            name="idle", description=IDLE_DESCRIPTION,
            parameters={"type": "object",
                        "properties": {"summary": {"type": "string",
                                                   "description": "Optional brief summary of what was accomplished"}},
                        "required": []},
            handler=lambda summary="": {"status": "idle", "summary": summary}))
        if getattr(llm_provider, "tool_provider", None) is None:
            try:
                llm_provider.tool_provider = self.tool_provider
            except Exception:
                pass

    async def run(self, messages: List[Message], model: str = "gpt-4o", temperature: float = 0.7,
                  max_tokens: Optional[int] = None, **kwargs: Any) -> AsyncGenerator[Dict[str, Any], None]:
        working = list(messages)
        if self.system_prompt and (not working or working[0].role != "system"):
            working.insert(0, Message(role="system", content=self.system_prompt))
        await self.tool_provider.get_tools()
        compaction_attempted = False

        iteration = 0
        while iteration < self.max_iterations:
            completion_id = f"chatcmpl-{uuid.uuid4().hex[:24]}"
            created = int(time.time())
            full_content = ""
            calls: Dict[int, Dict[str, Any]] = {}
            finish_reason = None
            yielded_any = False
            try:
                stream = self.llm_provider.stream_completion(working, model=model, temperature=temperature,
                                                             max_tokens=max_tokens, **kwargs)
                async for chunk in stream:
                    delta: Dict[str, Any] = {}
                    if chunk.role:
                        delta["role"] = chunk.role
                    if chunk.content:
                        delta["content"] = chunk.content
                        full_content += chunk.content
                    if chunk.tool_calls:
                        delta["tool_calls"] = [self._merge_tool_call(calls, tc) for tc in chunk.tool_calls]
                    if chunk.finish_reason:
                        finish_reason = chunk.finish_reason
                    yielded_any = True
                    yield {"id": completion_id, "object": "chat.completion.chunk", "created": created, "model": model,
                           "choices": [{"index": 0, "delta": delta,
                                        "finish_reason": finish_reason if chunk.finish_reason else None}]}
            except Exception as e:
                if (not yielded_any and looks_like_context_length_error(e) and not compaction_attempted
                        and self.context_compaction_provider is not None):
                    self.logger.info("context length error: %s — compacting", e)
                    try:
                        compacted = await self.context_compaction_provider.compact(
                            messages=[m.to_dict() for m in working], system_prompt=self.system_prompt or "",
                            model=model)
                    except Exception as ce:
                        self.logger.error("context compaction failed: %s", ce)
                        raise e
                    working = [Message(role=d.get("role", "user"), content=d.get("content"),
                                       tool_calls=d.get("tool_calls"), tool_call_id=d.get("tool_call_id"),
                                       name=d.get("name")) for d in compacted]
                    compaction_attempted = True
                    iteration += 1  # the reference's `continue` inside `for iteration in range(...)` also advances
                    continue        # retry with the compacted conversation
                raise

            tool_calls = [calls[i] for i in sorted(calls)]
            if not tool_calls:
                yield {"type": "agent_done", "reason": "text_response", "final_content": full_content,
                       "iteration": iteration}
                return
            working.append(Message(role="assistant", content=full_content or None, tool_calls=tool_calls))

            for tc in tool_calls:
                name, args_str, tcid = tc["function"]["name"], tc["function"]["arguments"], tc["id"]
                try:
                    args = json.loads(args_str) if args_str else {}
                except json.JSONDecodeError:
                    args = {}
                if name == "idle":
                    summary = args.get("summary", "")
                    payload = json.dumps({"status": "idle", "summary": summary})
                    working.append(Message(role="tool", content=payload, tool_call_id=tcid, name=name))
                    yield {"type": "tool_result", "tool_call_id": tcid, "tool_name": name, "delta": payload,
                           "is_complete": True}
                    yield {"type": "agent_done", "reason": "idle", "summary": summary, "iteration": iteration}
                    return
                result = ""
                async for ch in self.tool_provider.run_tool_stream(name, args, tcid):
                    result += ch.delta
                    yield {"type": "tool_result", "tool_call_id": tcid, "tool_name": name, "delta": ch.delta,
                           "is_complete": ch.is_complete}
                working.append(Message(role="tool", content=result, tool_call_id=tcid, name=name))
            iteration += 1

        yield {"type": "agent_done", "reason": "max_iterations", "iteration": self.max_iterations}

    @staticmethod
    def _merge_tool_call(calls: Dict[int, Dict[str, Any]], tc: Dict[str, Any]) -> Dict[str, Any]:
        """Accumulate one streamed tool-call delta (by `index`) and return the delta to forward."""
        idx = tc.get("index", 0)
        fn = tc.get("function") or {}
        if idx not in calls:
            calls[idx] = {"id": tc.get("id", ""), "type": "function",
                          "function": {"name": fn.get("name", ""), "arguments": ""}}
        acc = calls[idx]
        if fn.get("arguments"):
            acc["function"]["arguments"] += fn["arguments"]
        if tc.get("id"):
            acc["id"] = tc["id"]
        if fn.get("name"):
            acc["function"]["name"] = fn["name"]
        out: Dict[str, Any] = {"index": idx}
        if tc.get("id"):
            out["id"] = tc["id"]
            out["type"] = "function"
        if tc.get("function"):
            out["function"] = {k: fn[k] for k in ("name", "arguments") if fn.get(k)}
        return out
