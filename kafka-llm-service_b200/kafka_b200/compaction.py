"""Offline context compaction: what the agent does when the provider reports a context-length overflow.

Twin of the reference's `TruncationCompactionProvider`
(/root/reference/src/llm/context_compaction/v1.py:242-311) and of the two helpers it relies on,
`find_safe_split_point` and `validate_message_structure`
(/root/reference/src/llm/context_compaction/base.py:68-112, :115-167): keep the leading system message(s) and
the most recent `keep_count` messages, never cut between an assistant message that calls tools and the results
of those calls, drop tool results whose call was cut away and assistant messages left empty. Needs no model, so
it works offline — the reference's default (`ContextCompactionProviderV1`) summarises with a remote LLM and
cannot.  tests/test_compaction.py runs both implementations on the same random conversations.

One addition for local models, whose context (2 k - 8 k tokens) is far smaller than a remote model's: with
`count_tokens` and `token_budget` the cut moves further towards the present — always to a safe point — until the
compacted conversation fits the budget, so that the retry the agent makes after compaction can succeed instead
of overflowing again with its last 50 messages.
"""
from __future__ import annotations

import logging
from typing import Any, Callable, Dict, List, Optional, Sequence

Msg = Dict[str, Any]


def safe_cut(messages: Sequence[Msg], target: int) -> int:
    """Largest index <= target at which the list can be cut without separating a tool-calling assistant
    message from its tool results (same rule and result as the reference's find_safe_split_point)."""
    n = len(messages)
    if target <= 0:
        return 0
    if target >= n:
        return n
    cut = target
    while cut > 0:
        before, after = messages[cut - 1], messages[cut]
        opens_tools = before.get("role") == "assistant" and bool(before.get("tool_calls"))
        if opens_tools or after.get("role") == "tool":
            cut -= 1           # the cut would fall inside a call/result group: move in front of it
            continue
        break
    return cut


def drop_orphans(messages: Sequence[Msg], logger: Optional[logging.Logger] = None) -> List[Msg]:
    """Remove tool results whose call id no assistant message in the list announces, and assistant messages
    with neither text nor tool calls (the reference's validate_message_structure)."""
    announced = {tc["id"] for m in messages if m.get("role") == "assistant"
                 for tc in (m.get("tool_calls") or []) if tc.get("id")}
    kept: List[Msg] = []
    for m in messages:
        role = m.get("role")
        if role == "tool" and m.get("tool_call_id") not in announced:
            if logger:
                logger.warning("dropping a tool result without its call: %s", m.get("tool_call_id"))
            continue
        if role == "assistant" and not m.get("content") and not m.get("tool_calls"):
            if logger:
                logger.warning("dropping an empty assistant message")
            continue
        kept.append(m)
    return kept


class TruncationCompactionProvider:
    """compact(messages, system_prompt, model) -> shorter message list; same call as the reference's
    ContextCompactionProvider.compact (base.py:187-210), so either agent loop can use it."""

    def __init__(self, keep_count: int = 50, logger: Optional[logging.Logger] = None,
                 count_tokens: Optional[Callable[[List[Msg]], int]] = None, token_budget: Optional[int] = None):
        self.keep_count = keep_count
        self.logger = logger or logging.getLogger(__name__)
        self.count_tokens = count_tokens
        self.token_budget = token_budget

    async def compact(self, messages: List[Msg], system_prompt: str = "", model: str = "", **kwargs: Any) -> List[Msg]:
        head: List[Msg] = []
        body: List[Msg] = []
        for m in messages:
            (head if m.get("role") == "system" and not body else body).append(m)
        fits = self.count_tokens is None or self.token_budget is None or \
            self.count_tokens(list(messages)) <= self.token_budget
        if len(body) <= self.keep_count and fits:
            return messages                          # nothing to cut (the reference returns its input unchanged)
        cut = safe_cut(body, max(0, len(body) - self.keep_count))
        out = drop_orphans(head + body[cut:], self.logger)
        if self.count_tokens is not None and self.token_budget is not None:
            # local models: keep cutting, one safe point at a time, until the prompt fits
            while self.count_tokens(out) > self.token_budget and cut < len(body) - 1:
                nxt = cut + 1
                while nxt < len(body) - 1 and safe_cut(body, nxt) <= cut:
                    nxt += 1
                cut = max(cut + 1, safe_cut(body, nxt))
                out = drop_orphans(head + body[cut:], self.logger)
        self.logger.info("context compaction by truncation: %d -> %d messages", len(messages), len(out))
        return out
