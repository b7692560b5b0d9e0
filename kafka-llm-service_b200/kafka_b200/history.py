"""History assembly for a thread turn — the host steps that precede every provider call.

Mirrors, with the same observable results:
  * sanitising: /root/reference/src/kafka/utils.py:25-61 (`sanitize_messages_for_openai`) — a tool
    message survives only if its tool_call_id was announced by the closest preceding assistant message
    with tool_calls and has not been answered yet; anything else in between resets the pending set;
  * the turn sequence of /root/reference/server.py:391-404 and
    /root/reference/src/kafka/base.py:196-209: ensure thread, load history, convert the new messages,
    persist the new user/system messages, sanitise history + new.

BASELINE requires this to be bit-exact with the reference; tests/test_history_golden.py pins it against
a digest produced by running the reference's own functions over its bundled data/threads.db.
"""
from __future__ import annotations

from typing import Any, Iterable, List, Sequence

from .llm.types import Message


def to_message(m: Any) -> Message:
    """OpenAI-shaped dict / pydantic ChatMessage / Message -> Message (kafka/utils.py:14-22)."""
    if isinstance(m, Message):
        return m
    get = (lambda k: m.get(k)) if isinstance(m, dict) else (lambda k: getattr(m, k, None))
    return Message(role=get("role"), content=get("content"), name=get("name"), tool_calls=get("tool_calls"),
                   tool_call_id=get("tool_call_id"))


def sanitize_messages(messages: Sequence[Message]) -> List[Message]:
    """Drop orphan tool messages so that every tool message answers a pending call."""
    kept: List[Message] = []
    pending: set = set()
    for m in messages:
        if m.role == "tool":
            if m.tool_call_id and m.tool_call_id in pending:
                pending.discard(m.tool_call_id)
                kept.append(m)
            continue  # orphan: silently dropped (the reference prints a warning)
        if m.role == "assistant" and m.tool_calls:
            pending = {tc.get("id") for tc in m.tool_calls if tc.get("id")}
        else:
            pending = set()
        kept.append(m)
    return kept


async def assemble_turn(store, thread_id: str, new_messages: Iterable[Any], save_new: bool = True) -> List[Message]:
    """Load -> convert -> persist new user/system messages -> sanitise. Returns what the provider sees."""
    if not await store.thread_exists(thread_id):
        await store.create_thread(thread_id=thread_id)
    history = await store.get_thread_messages(thread_id)
    new = [to_message(m) for m in new_messages]
    if save_new:
        for m in new:
            if m.role in ("user", "system"):
                await store.add_message(thread_id, m)
    return sanitize_messages(list(history) + new)
