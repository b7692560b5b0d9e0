"""Chat template: messages (+ tool schemas) -> token ids, prefix-stably.

The reference never templates anything: it forwards `[m.to_dict() for m in messages]` plus the tool
list to the gateway (/root/reference/src/llm/portkey.py:263, :317-319). A local engine has to choose
a rendering; this one is Llama-3-shaped and has the one property KV reuse depends on:

    render(messages[:k]) is a token PREFIX of render(messages[:k+1])      (add_generation_prompt=False)

because each message renders independently of what follows, and the tool block sits before all
messages. Every field the reference's Message carries (role, content, name, tool_calls, tool_call_id
— /root/reference/src/llm/types.py:47-51) is rendered, tool_calls as canonical JSON
(sort_keys, compact separators) so equal structures give equal tokens.
"""
from __future__ import annotations

import json
from typing import Any, Dict, List, Optional, Sequence

from .tokenizer import ByteTokenizer


def canonical_json(obj: Any) -> str:
    return json.dumps(obj, sort_keys=True, separators=(",", ":"), ensure_ascii=False)


def _field(msg: Any, key: str):
    return msg.get(key) if isinstance(msg, dict) else getattr(msg, key, None)


class ChatTemplate:
    def __init__(self, tokenizer: ByteTokenizer):
        self.tok = tokenizer
        self._headers: Dict[str, List[int]] = {}     # label -> ids (a handful of distinct labels per process)

    # ---- pieces ------------------------------------------------------------------------------
    def _header(self, label: str) -> List[int]:
        ids = self._headers.get(label)
        if ids is None:
            t = self.tok
            ids = [t.start_header_id] + t.encode(label) + [t.end_header_id] + t.encode("\n\n")
            if len(self._headers) < 4096:                # labels with tool_call_ids are unbounded: stop caching
                self._headers[label] = ids
        return ids

    # How a model is told to call the tools: the block provider.py parses out of the generated text and turns into
    # StreamChunk.tool_calls (`<tool_calls>[...]</tool_calls>`, the form assistant tool calls are rendered in below,
    # so a call that comes back as history is the same tokens). A real instruction-tuned checkpoint needs to be told;
    # the random-init models of the tests ignore it.
    TOOLS_PREAMBLE = ('You may call the tools described below. To call tools, end your reply with '
                      '<tool_calls>[{"name": "<tool name>", "arguments": {<arguments as JSON>}}]</tool_calls> '
                      '(one list entry per call) and write nothing after it; the results come back as tool messages.\n')

    def render_tools(self, tools: Optional[Sequence[Dict[str, Any]]]) -> List[int]:
        if not tools:
            return []
        return (self._header("tools") + self.tok.encode(self.TOOLS_PREAMBLE + canonical_json(list(tools))) +
                [self.tok.eot_id])

    def message_body(self, msg: Any) -> str:
        content = _field(msg, "content")
        body = content if isinstance(content, str) else ("" if content is None else canonical_json(content))
        tool_calls = _field(msg, "tool_calls")
        if tool_calls:
            # on its own line unless the text already ends one (what a streamed reply that opened the
            # block looks like when it comes back as history: the same tokens, so the KV prefix holds)
            if body and not body.endswith("\n"):
                body += "\n"
            body += "<tool_calls>" + canonical_json(tool_calls) + "</tool_calls>"
        return body

    def render_message(self, msg: Any) -> List[int]:
        label = str(_field(msg, "role"))
        name = _field(msg, "name")
        if name:
            label += f" name={name}"
        tcid = _field(msg, "tool_call_id")
        if tcid:
            label += f" tool_call_id={tcid}"
        ids = list(self._header(label))
        ids += self.message_body(msg).encode("utf-8") if type(self.tok) is ByteTokenizer \
            else self.tok.encode(self.message_body(msg))
        ids.append(self.tok.eot_id)
        return ids

    def generation_prompt(self) -> List[int]:
        return self._header("assistant")

    # ---- whole conversation ------------------------------------------------------------------
    def render(self, messages: Sequence[Any], tools: Optional[Sequence[Dict[str, Any]]] = None,
               add_generation_prompt: bool = True, cache_key: Optional[str] = None) -> List[int]:
        """`cache_key` (a thread id): a follow-up turn re-sends the whole conversation, and because every message
        renders independently of what follows, the tokens of the messages it has in common with the thread's
        previous call are taken from RenderCache instead of being tokenised again (the result is the same list)."""
        if cache_key is not None and RenderCache.enabled:
            return RenderCache.render(self, cache_key, messages, tools, add_generation_prompt)
        ids: List[int] = [self.tok.bos_id]
        ids += self.render_tools(tools)
        for m in messages:
            ids += self.render_message(m)
        if add_generation_prompt:
            ids += self.generation_prompt()
        return ids


def _message_key(msg: Any):
    """Everything render_message looks at, cheap to compare (strings compare at memcmp speed)."""
    tc = _field(msg, "tool_calls")
    content = _field(msg, "content")
    if content is not None and not isinstance(content, str):
        content = canonical_json(content)
    return (str(_field(msg, "role")), content, _field(msg, "name"), _field(msg, "tool_call_id"),
            canonical_json(tc) if tc else None)


class RenderCache:
    """Per (tokenizer, thread): the message keys of the last rendered conversation, the token offset after each
    message, and the tokens. Process-wide (the server builds a provider — and a ChatTemplate — per request),
    least-recently-used, bounded by a total token count. KAFKA_B200_RENDER_CACHE=0 switches it off."""
    import os as _os
    enabled = _os.environ.get("KAFKA_B200_RENDER_CACHE", "1") != "0"
    max_tokens = 16_000_000      # ~36 B per cached id in CPython: at most ~0.6 GB
    _lock = __import__("threading").Lock()
    _entries: "Dict[Any, Any]" = {}          # (tokenizer.cache_id, cache_key) -> [tools_key, keys, ends, ids]; dict order = LRU
    _total = 0

    @classmethod
    def clear(cls) -> None:
        with cls._lock:
            cls._entries.clear()
            cls._total = 0

    @classmethod
    def render(cls, tmpl: "ChatTemplate", cache_key: str, messages, tools, add_generation_prompt: bool) -> List[int]:
        tok_id = getattr(tmpl.tok, "cache_id", None)      # which vocabulary the cached ids belong to
        if tok_id is None:
            return tmpl.render(messages, tools, add_generation_prompt)
        tools_key = canonical_json(list(tools)) if tools else ""
        keys = [_message_key(m) for m in messages]
        k = (tok_id, cache_key)
        with cls._lock:
            ent = cls._entries.pop(k, None)
            if ent is not None:
                cls._total -= len(ent[3])
        n_same = 0
        if ent is not None and ent[0] == tools_key:
            old_keys = ent[1]
            lim = min(len(old_keys), len(keys))
            while n_same < lim and old_keys[n_same] == keys[n_same]:
                n_same += 1
        if n_same:
            ends = ent[2][:n_same]
            ids = ent[3][:ends[-1]]
        else:
            ids = [tmpl.tok.bos_id] + tmpl.render_tools(tools)
            ends = []
        for m in messages[n_same:]:
            ids += tmpl.render_message(m)
            ends.append(len(ids))
        with cls._lock:
            cls._entries[k] = [tools_key, keys, ends, ids]
            cls._total += len(ids)
            while cls._total > cls.max_tokens and len(cls._entries) > 1:
                _, old = next(iter(cls._entries.items()))
                cls._total -= len(old[3])
                del cls._entries[next(iter(cls._entries))]
        out = list(ids)                      # the cache keeps its own list: callers may mutate what they get
        if add_generation_prompt:
            out += tmpl.generation_prompt()
        return out
