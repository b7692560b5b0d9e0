"""ThreadRunner — runs the agent over a persisted thread and records what the model actually saw.

Mirror of `KafkaAgent.run_with_thread` (/root/reference/src/kafka/base.py:171-310): ensure the thread,
load + sanitise history, persist the new user/system messages, forward every agent event, and rebuild
the assistant / tool messages from the event stream (tool_result deltas accumulated per call id until
is_complete, :232-249; content and tool-call deltas merged until finish_reason "tool_calls" or "stop",
:252-296), saving them at the end, plus the final text when no saved message already carries it
(:302-310).

One addition: a completion that ends with finish_reason "length" (max_tokens) also closes the current
assistant message. The reference only closes on "stop"/"tool_calls" and relies on the final_content
fallback; the result in the database is the same, the code path is explicit.

Because the stored history then equals the model-visible history, the next turn's rendered prompt
extends the resident token prefix and only the new messages are prefilled.
"""
from __future__ import annotations

from typing import Any, AsyncGenerator, Dict, Iterable, List, Optional

from .history import sanitize_messages, to_message
from .llm.types import Message


class ThreadRunner:
    def __init__(self, agent: Any, store: Any, thread_id: str, engine: Any = None):
        self.agent, self.store, self.thread_id, self.engine = agent, store, thread_id, engine

    async def _record_residency(self) -> None:
        """Refresh the thread -> KV-block directory (kv_residency table) from the engine(s)."""
        eng = self.engine
        if eng is None or not hasattr(self.store, "record_residency"):
            return
        tid = self.thread_id
        by_gpu = eng.residency_by_gpu(tid) if hasattr(eng, "residency_by_gpu") else {0: eng.residency(tid)}
        found = False
        for gpu, rows in by_gpu.items():
            for row in rows:
                await self.store.upsert_residency(row, gpu=gpu)
                found = True
        if not found:
            await self.store.drop_residency(tid)

    async def run(self, new_messages: Iterable[Any], model: str, temperature: float = 0.7,
                  max_tokens: Optional[int] = None, save_to_thread: bool = True,
                  **kwargs: Any) -> AsyncGenerator[Dict[str, Any], None]:
        store, tid = self.store, self.thread_id
        if not await store.thread_exists(tid):
            await store.create_thread(thread_id=tid)
        history = await store.get_thread_messages(tid)
        new = [to_message(m) for m in new_messages]
        all_messages = sanitize_messages(list(history) + new)
        if save_to_thread:
            for m in new:
                if m.role in ("user", "system"):
                    await store.add_message(tid, m)

        to_save: List[Message] = []
        final_content = ""
        content = ""
        calls: Dict[int, Dict[str, Any]] = {}
        results: Dict[str, str] = {}

        async for event in self.agent.run(all_messages, model=model, temperature=temperature, max_tokens=max_tokens,
                                          thread_id=tid, **kwargs):
            yield event
            if not save_to_thread:
                continue
            kind = event.get("type")
            if kind == "tool_result":
                cid = event.get("tool_call_id", "")
                results[cid] = results.get(cid, "") + event.get("delta", "")
                if event.get("is_complete", False):
                    to_save.append(Message(role="tool", content=results[cid], tool_call_id=cid,
                                           name=event.get("tool_name")))
            elif event.get("choices"):
                choice = event["choices"][0]
                delta = choice.get("delta", {})
                if delta.get("content"):
                    content += delta["content"]
                for tc in delta.get("tool_calls") or []:
                    slot = calls.setdefault(tc.get("index", 0),
                                            {"id": "", "type": "function", "function": {"name": "", "arguments": ""}})
                    fn = tc.get("function") or {}
                    if tc.get("id"):
                        slot["id"] = tc["id"]
                    if fn.get("name"):
                        slot["function"]["name"] = fn["name"]
                    if fn.get("arguments"):
                        slot["function"]["arguments"] += fn["arguments"]
                finish = choice.get("finish_reason")
                if finish == "tool_calls":
                    to_save.append(Message(role="assistant", content=content or None,
                                           tool_calls=[calls[i] for i in sorted(calls)]))
                    content, calls = "", {}
                elif finish in ("stop", "length") and content:
                    to_save.append(Message(role="assistant", content=content))
                    content = ""
            elif kind == "agent_done":
                final_content = event.get("final_content", "")

        if save_to_thread:
            await store.add_messages(tid, to_save)
            if final_content and not any(m.role == "assistant" and m.content == final_content for m in to_save):
                await store.add_message(tid, Message(role="assistant", content=final_content))
        await self._record_residency()
