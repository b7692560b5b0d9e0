"""B200LLMProvider — the drop-in for the reference's PortkeyLLMProvider on local Llama models.

Same plugin surface as /root/reference/src/llm/portkey.py (stream_completion :225-483, completion
:485-669, get_model_info :671, close :691) and the same chunk conventions the reference's agent
reads (/root/reference/src/agents/base.py:274-348): a first chunk with role="assistant", content
chunks (content is None — never "" — when absent), a final chunk whose finish_reason is "stop" or
"length", `id` and `model` on every chunk. Failures surface as LLMProviderError(provider="B200")
(:477-483); an over-long prompt raises an error that is_context_length_error
(/root/reference/src/llm/context_compaction/base.py:10-65) recognises, so the agent's compaction
retry still fires.

What differs by design: the completion is computed on this box (engine.py -> C ABI -> sm_100a
kernels), one StreamChunk per generated token, and the conversation's KV stays resident between
calls — the prompt is re-templated and matched against resident token prefixes, so only the new
messages are prefilled.
"""
from __future__ import annotations

import asyncio
import collections
import uuid
import weakref
from typing import Any, AsyncGenerator, Dict, List, Optional

from .llm.base import LLMProvider
from .llm.types import CompletionResponse, LLMProviderError, Message, StreamChunk, content_chunk
from .scheduler import ContextLengthError, SamplingParams
from .template import ChatTemplate
from .tokenizer import ByteTokenizer

PROVIDER_NAME = "B200"
TOOL_OPEN, TOOL_CLOSE = "<tool_calls>", "</tool_calls>"      # the block template.py renders tool calls in


def parse_tool_calls(payload: str) -> Optional[List[Dict[str, Any]]]:
    """JSON between the tool-call tags -> normalised OpenAI tool_calls, or None if it is not a list of
    calls. Accepts both the full shape ({"id","type","function":{"name","arguments"}}) and the short
    {"name","arguments"}; `arguments` may be an object (re-serialised) or a string."""
    import json
    try:
        data = json.loads(payload)
    except ValueError:
        return None
    if isinstance(data, dict):
        data = [data]
    if not isinstance(data, list) or not data:
        return None
    out = []
    for item in data:
        if not isinstance(item, dict):
            return None
        fn = item.get("function") if isinstance(item.get("function"), dict) else item
        name, args = fn.get("name"), fn.get("arguments", {})
        if not isinstance(name, str) or not name:
            return None
        if not isinstance(args, str):
            args = json.dumps(args)
        out.append({"id": item.get("id") or f"call_{uuid.uuid4().hex[:12]}", "type": "function",
                    "function": {"name": name, "arguments": args}})
    return out


def tool_call_chunks(calls: List[Dict[str, Any]], model: Optional[str], cid: Optional[str]):
    """Tool calls as the delta stream the reference's agent merges by `index`
    (/root/reference/src/agents/base.py:286-331): id + name first, then the argument string."""
    for i, c in enumerate(calls):
        yield StreamChunk(tool_calls=[{"index": i, "id": c["id"], "type": "function",
                                       "function": {"name": c["function"]["name"], "arguments": ""}}],
                          model=model, id=cid)
        yield StreamChunk(tool_calls=[{"index": i, "function": {"arguments": c["function"]["arguments"]}}],
                          model=model, id=cid)


class _LoopMailbox:
    """Engine thread -> event loop hand-off, batched: a decode step produces one event for each of
    hundreds of streams; instead of one `call_soon_threadsafe` (a lock and a wake-up write) per event,
    events are appended to a deque and the loop is woken once per batch."""
    _by_loop: "weakref.WeakKeyDictionary" = weakref.WeakKeyDictionary()

    def __init__(self, loop: asyncio.AbstractEventLoop):
        self.loop = loop
        self.items: "collections.deque" = collections.deque()
        self.scheduled = False

    @classmethod
    def for_running_loop(cls) -> "_LoopMailbox":
        loop = asyncio.get_running_loop()
        mb = cls._by_loop.get(loop)
        if mb is None:
            mb = cls._by_loop[loop] = cls(loop)
        return mb

    def post(self, queue: "asyncio.Queue", ev) -> None:        # any thread
        self.items.append((queue, ev))
        if not self.scheduled:                                  # a redundant wake-up is harmless
            self.scheduled = True
            self.loop.call_soon_threadsafe(self._drain)

    def _drain(self) -> None:                                   # event-loop thread
        self.scheduled = False
        items = self.items
        while items:
            queue, ev = items.popleft()
            queue.put_nowait(ev)


class B200LLMProvider(LLMProvider):
    def __init__(self, engine, tokenizer: ByteTokenizer, model: str, tool_provider: Optional[Any] = None,
                 default_temperature: float = 0.7, default_max_tokens: Optional[int] = None,
                 thread_id: Optional[str] = None, seed: int = 0, default_top_p: Optional[float] = None):
        super().__init__(tool_provider=tool_provider)
        self.engine = engine            # Engine or ShardedEngine (process-wide, shared by providers)
        self.tokenizer = tokenizer
        self.template = ChatTemplate(tokenizer)
        self.model = model
        self.default_temperature = default_temperature
        self.default_max_tokens = default_max_tokens
        self.thread_id = thread_id
        self.seed = seed
        self.default_top_p = default_top_p
        # token counts summed over this object's completions (the server builds one provider per request, so
        # this is the request's usage, agent tool rounds included)
        self.usage_total: Dict[str, int] = {"prompt_tokens": 0, "completion_tokens": 0, "total_tokens": 0,
                                            "cached_tokens": 0}

    def _add_usage(self, usage: Dict[str, int]) -> None:
        for k in self.usage_total:
            self.usage_total[k] += int(usage.get(k, 0))

    # ---- helpers -------------------------------------------------------------------------------
    def render(self, messages: List[Message], tools: Optional[List[Dict[str, Any]]],
               thread_id: Optional[str] = None) -> List[int]:
        return self.template.render(messages, tools=tools or None, add_generation_prompt=True, cache_key=thread_id)

    @staticmethod
    def _first_marker(text: str, markers: List[str], start: int) -> int:
        if not markers:
            return -1
        hits = [i for i in (text.find(s, start) for s in markers) if i >= 0]
        return min(hits) if hits else -1

    @staticmethod
    def _partial_marker(text: str, markers: List[str], hold: int) -> int:
        """Length of the longest suffix of `text` that is a proper prefix of some marker."""
        if not markers:
            return 0
        for k in range(min(hold, len(text)), 0, -1):
            tail = text[-k:]
            for m in markers:
                if m.startswith(tail):
                    return k
        return 0

    def count_prompt_tokens(self, messages: List[Any]) -> int:
        """Tokens the prompt of `messages` (Message objects or OpenAI-style dicts) would take, tool schemas of
        the most recent call included — what a compaction provider needs to know whether its result fits."""
        msgs = [m if isinstance(m, Message) else
                Message(role=m.get("role", "user"), content=m.get("content"), name=m.get("name"),
                        tool_calls=m.get("tool_calls"), tool_call_id=m.get("tool_call_id")) for m in messages]
        return len(self.render(msgs, getattr(self, "_last_tools", None)))

    def compaction_provider(self, keep_count: int = 50, fraction: float = 0.75, reserve: int = 0):
        """The offline compaction for this provider's model (kafka_b200/compaction.py): the reference's
        truncation rule, continued until the prompt takes at most `fraction` of the context minus `reserve`
        tokens for the answer. Pass it to the agent as `context_compaction_provider`."""
        from .compaction import TruncationCompactionProvider
        budget = max(64, int(self.engine.max_pos * fraction) - reserve)
        return TruncationCompactionProvider(keep_count=keep_count, count_tokens=self.count_prompt_tokens,
                                            token_budget=budget)

    # ---- the plugin API ------------------------------------------------------------------------
    async def stream_completion(self, messages: List[Message], *, temperature: Optional[float] = None,
                                max_tokens: Optional[int] = None, stop: Optional[List[str]] = None,
                                **kwargs: Any) -> AsyncGenerator[StreamChunk, None]:
        validated = self.validate_messages(messages)          # ValueError on empty, like the reference
        model = kwargs.pop("model", None) or self.model
        thread_id = kwargs.pop("thread_id", None) or self.thread_id
        seed = int(kwargs.pop("seed", self.seed) or 0)
        usage_out = kwargs.pop("_usage_out", None)            # completion() collects usage through this
        # `top_p`: one of the extra parameters the reference's provider documents and forwards
        # (/root/reference/src/llm/portkey.py:245-249, :375); applied by the sampling kernel
        top_p = kwargs.pop("top_p", None)
        top_p = self.default_top_p if top_p is None else top_p
        top_p = float(top_p) if top_p is not None and 0.0 < float(top_p) < 1.0 else 0.0
        completion_id = f"chatcmpl-{uuid.uuid4().hex[:24]}"
        req = None
        try:
            tools = await self.get_tools()
            self._last_tools = tools                          # count_prompt_tokens() sizes prompts like this call
            prompt = self.render(validated, tools, thread_id)
            temp = self.default_temperature if temperature is None else float(temperature)
            limit = max_tokens if max_tokens is not None else (self.default_max_tokens or 512)
            events: "asyncio.Queue" = asyncio.Queue()
            mailbox = _LoopMailbox.for_running_loop()

            def on_event(ev, _post=mailbox.post, _q=events):
                _post(_q, ev)

            req = self.engine.submit(prompt, SamplingParams(temperature=temp, seed=seed, max_tokens=int(limit), top_p=top_p),
                                     on_event, thread_hint=thread_id)
        except ContextLengthError as e:
            raise LLMProviderError(str(e), status_code=400, provider=PROVIDER_NAME, original_error=e) from e
        except (ValueError, LLMProviderError):
            raise
        except Exception as e:
            raise LLMProviderError(str(e), provider=PROVIDER_NAME, original_error=e) from e

        try:
            yield StreamChunk(role="assistant", model=model, id=completion_id)
            decoder = self.tokenizer.stream_decoder()
            text = ""
            emitted = 0
            # strings that end plain-text streaming: the caller's stop sequences and, when tools are on
            # offer, the opening tag of a tool-call block (the syntax template.py renders tool calls in)
            markers = [s for s in (stop or []) if s] + ([TOOL_OPEN] if tools else [])
            hold = max((len(s) for s in markers), default=1) - 1   # longest partial marker a tail can be
            collecting = -1                            # >= 0: index where a tool-call block opened
            n_tokens = 0

            def cut_short():       # the stream ends before the engine's own finish event (stop string, tool call)
                req.cancelled = True
                self._add_usage({"prompt_tokens": len(prompt), "completion_tokens": n_tokens,
                                 "total_tokens": len(prompt) + n_tokens, "cached_tokens": req.n_cached})
            while True:
                ev = await events.get()
                kind = ev[0]
                if kind == "token":
                    n_tokens += 1
                    text += decoder.push(ev[1])
                    if collecting >= 0:
                        end = text.find(TOOL_CLOSE, collecting)
                        if end < 0:
                            continue
                        calls = parse_tool_calls(text[collecting + len(TOOL_OPEN):end])
                        if calls is None:              # not valid JSON after all: hand it over as text
                            yield StreamChunk(content=text[emitted:], model=model, id=completion_id)
                            emitted, collecting = len(text), -1
                            continue
                        cut_short()
                        for ch in tool_call_chunks(calls, model, completion_id):
                            yield ch
                        yield StreamChunk(finish_reason="tool_calls", model=model, id=completion_id)
                        return
                    cut = self._first_marker(text, markers, max(0, emitted - hold))
                    if cut >= 0:
                        if tools and text.startswith(TOOL_OPEN, cut):
                            body = text[emitted:cut].rstrip("\n")
                            if body:
                                yield StreamChunk(content=body, model=model, id=completion_id)
                            emitted = collecting = cut
                            continue
                        if cut > emitted:
                            yield StreamChunk(content=text[emitted:cut], model=model, id=completion_id)
                        cut_short()
                        yield StreamChunk(finish_reason="stop", model=model, id=completion_id)
                        return
                    # keep back only a tail that is the beginning of a marker (usually nothing)
                    upto = len(text) - self._partial_marker(text, markers, hold)
                    if upto > emitted:
                        yield content_chunk(text[emitted:upto], model, completion_id)   # per-token hot spot
                        emitted = upto
                elif kind == "finish":
                    text += decoder.flush()
                    if len(text) > emitted:
                        yield StreamChunk(content=text[emitted:], model=model, id=completion_id)
                    self._add_usage(ev[2])
                    if usage_out is not None:
                        usage_out.update(ev[2])
                    yield StreamChunk(finish_reason=ev[1], model=model, id=completion_id)
                    return
                else:  # ("error", exc)
                    exc = ev[1]
                    raise LLMProviderError(str(exc), provider=PROVIDER_NAME, original_error=exc)
        finally:
            # abandoned mid-stream (client disconnect / aclose): free the slot, keep the KV computed so far
            if req is not None:
                req.cancelled = True

    async def completion(self, messages: List[Message], *, temperature: Optional[float] = None,
                         max_tokens: Optional[int] = None, stop: Optional[List[str]] = None,
                         **kwargs: Any) -> CompletionResponse:
        usage: Dict[str, int] = {}
        parts: List[str] = []
        calls: Dict[int, Dict[str, Any]] = {}
        finish = None
        cid = model = None
        async for ch in self.stream_completion(messages, temperature=temperature, max_tokens=max_tokens, stop=stop,
                                               _usage_out=usage, **kwargs):
            if ch.content:
                parts.append(ch.content)
            for tc in ch.tool_calls or []:
                slot = calls.setdefault(tc.get("index", 0), {"id": "", "type": "function",
                                                             "function": {"name": "", "arguments": ""}})
                fn = tc.get("function") or {}
                slot["id"] = tc.get("id") or slot["id"]
                slot["function"]["name"] = fn.get("name") or slot["function"]["name"]
                slot["function"]["arguments"] += fn.get("arguments") or ""
            finish = ch.finish_reason or finish
            cid, model = ch.id, ch.model
        return CompletionResponse(content="".join(parts), finish_reason=finish, model=model, id=cid,
                                  usage=usage or None, tool_calls=[calls[i] for i in sorted(calls)] or None)

    def get_model_info(self) -> Dict[str, Any]:
        return {"model": self.model, "provider": PROVIDER_NAME, "max_context_tokens": self.engine.max_pos,
                "supports_vision": False, "supports_tools": True, "default_temperature": self.default_temperature}

    async def close(self) -> None:
        """The engine is process-wide and outlives provider objects (one is built per request on
        /agent/run, /root/reference/server.py:237-245), so closing a provider is a no-op like the
        reference's (/root/reference/src/llm/portkey.py:691-701)."""
        return None

    def _create_client_for_provider(self, provider_name: str):
        """/root/reference/src/kafka/v1.py:185 reaches into this private method to build the compaction
        client. There is no remote client here; returning None makes summarisation unavailable, and the
        construction code should pick TruncationCompactionProvider."""
        return None


class EchoLLMProvider(LLMProvider):
    """Returns the last user message. Host-plumbing stand-in for BASELINE config 1 (thread replay on
    CPU: history load / assembly / persistence without any model)."""

    def __init__(self, model: str = "echo", tool_provider: Optional[Any] = None):
        super().__init__(tool_provider=tool_provider)
        self.model = model

    async def stream_completion(self, messages, *, temperature=None, max_tokens=None, stop=None, **kwargs):
        validated = self.validate_messages(messages)
        model = kwargs.pop("model", None) or self.model
        cid = f"chatcmpl-{uuid.uuid4().hex[:24]}"
        last_user = next((m.content for m in reversed(validated) if m.role == "user" and m.content), "")
        yield StreamChunk(role="assistant", model=model, id=cid)
        if last_user:
            yield StreamChunk(content=last_user, model=model, id=cid)
        yield StreamChunk(finish_reason="stop", model=model, id=cid)

    async def completion(self, messages, *, temperature=None, max_tokens=None, stop=None, **kwargs):
        parts, finish, cid, model = [], None, None, None
        async for ch in self.stream_completion(messages, **kwargs):
            if ch.content:
                parts.append(ch.content)
            finish, cid, model = ch.finish_reason or finish, ch.id, ch.model
        text = "".join(parts)
        return CompletionResponse(content=text, finish_reason=finish, model=model, id=cid,
                                  usage={"prompt_tokens": 0, "completion_tokens": len(text), "total_tokens": len(text)})

    async def close(self) -> None:
        return None


class ScriptedToolCallProvider(LLMProvider):
    """Deterministic tool-calling policy layered on a real provider (BASELINE config 4).

    A random-init model never emits a parseable tool call, yet the agent loop's defining load pattern
    is "several provider calls per turn, each re-sending a conversation that grew by the previous
    output, a tool call and its result" (/root/reference/src/agents/base.py:210-433). This wrapper makes
    that pattern reproducible: while fewer than `tool_rounds` tool results follow the last user message
    it lets the inner provider generate `think_tokens` tokens of text and then appends one scripted
    call to `tool_name` (finish_reason "tool_calls"); afterwards it delegates unchanged, so the last
    round is an ordinary text completion. Every round runs real prefill + decode on the engine, and
    because each round's prompt extends the previous one, only the new messages are prefilled.
    """

    def __init__(self, inner: LLMProvider, tool_name: str, tool_args: Optional[Dict[str, Any]] = None,
                 tool_rounds: int = 2, think_tokens: int = 16):
        super().__init__(tool_provider=inner.tool_provider)
        self.inner, self.tool_name, self.tool_args = inner, tool_name, dict(tool_args or {})
        self.tool_rounds, self.think_tokens = tool_rounds, think_tokens
        self.model = getattr(inner, "model", None)

    @property
    def tool_provider(self):
        return self.inner.tool_provider

    @tool_provider.setter
    def tool_provider(self, provider) -> None:
        self._tool_provider = provider
        if hasattr(self, "inner"):
            self.inner.tool_provider = provider

    @staticmethod
    def _rounds_done(messages: List[Message]) -> int:
        n = 0
        for m in reversed(messages):
            if m.role == "user":
                break
            if m.role == "tool":
                n += 1
        return n

    async def stream_completion(self, messages: List[Message], *, temperature: Optional[float] = None,
                                max_tokens: Optional[int] = None, stop: Optional[List[str]] = None,
                                **kwargs: Any) -> AsyncGenerator[StreamChunk, None]:
        done = self._rounds_done(self.validate_messages(messages))
        if done >= self.tool_rounds:
            async for ch in self.inner.stream_completion(messages, temperature=temperature, max_tokens=max_tokens,
                                                         stop=stop, **kwargs):
                yield ch
            return
        last = None
        async for ch in self.inner.stream_completion(messages, temperature=temperature,
                                                     max_tokens=self.think_tokens, stop=stop, **kwargs):
            if ch.finish_reason is None:
                yield ch
            last = ch
        import json as _json
        call_id = f"call_{uuid.uuid4().hex[:12]}"
        yield StreamChunk(tool_calls=[{"index": 0, "id": call_id, "type": "function",
                                       "function": {"name": self.tool_name, "arguments": ""}}],
                          model=last.model if last else self.model, id=last.id if last else None)
        yield StreamChunk(tool_calls=[{"index": 0, "function": {"arguments": _json.dumps(self.tool_args)}}],
                          model=last.model if last else self.model, id=last.id if last else None)
        yield StreamChunk(finish_reason="tool_calls", model=last.model if last else self.model,
                          id=last.id if last else None)

    async def completion(self, messages: List[Message], **kwargs: Any) -> CompletionResponse:
        parts: List[str] = []
        calls: Dict[int, Dict[str, Any]] = {}
        finish = cid = model = None
        async for ch in self.stream_completion(messages, **kwargs):
            if ch.content:
                parts.append(ch.content)
            for tc in ch.tool_calls or []:
                slot = calls.setdefault(tc.get("index", 0), {"id": "", "type": "function",
                                                             "function": {"name": "", "arguments": ""}})
                fn = tc.get("function") or {}
                slot["id"] = tc.get("id") or slot["id"]
                slot["function"]["name"] = fn.get("name") or slot["function"]["name"]
                slot["function"]["arguments"] += fn.get("arguments") or ""
            finish, cid, model = ch.finish_reason or finish, ch.id, ch.model
        return CompletionResponse(content="".join(parts) or None, finish_reason=finish, model=model, id=cid,
                                  tool_calls=[calls[i] for i in sorted(calls)] or None)

    async def close(self) -> None:
        await self.inner.close()
