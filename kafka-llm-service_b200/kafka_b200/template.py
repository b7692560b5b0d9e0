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
               add_generation_prompt: bool = True) -> List[int]:
        ids: List[int] = [self.tok.bos_id]
        ids += self.render_tools(tools)
        for m in messages:
            ids += self.render_message(m)
        if add_generation_prompt:
            ids += self.generation_prompt()
        return ids
