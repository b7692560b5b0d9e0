"""ORACLE — CPUOracleProvider: the fp32 CPU forward wrapped in the LLMProvider plugin interface.
TEST INFRASTRUCTURE ONLY (same import rule as oracle/llama_fp32.py; parity unpinned, see there).

BASELINE.json north_star: "an fp32 PyTorch CPU eager forward of the same random-init architecture
wrapped in the reference's own LLMProvider interface ... run inside the reference server on the box's
host cores ... is also the reported baseline".  This class is that provider: it subclasses this repo's
structural twin of the reference's ABC (registered as a virtual subclass of the real one by
kafka_b200.llm.register_with_reference() when /root/reference is importable, and accepted unchanged
by the reference's Agent / KafkaAgent, which only duck-type the provider), shares the product's tokenizer, chat
template and seeded weights, and decodes greedily (or with the same candidate set as the GPU sampler).

It deliberately keeps NO state between calls: every completion re-runs the whole conversation, like
the reference's remote provider does (/root/reference/src/agents/base.py:222-223). That makes it an
independent check of the GPU engine's KV reuse.
"""
from __future__ import annotations

import uuid
from typing import Any, List, Optional

import torch

from kafka_b200.llm.base import LLMProvider
from kafka_b200.llm.types import CompletionResponse, StreamChunk
from kafka_b200.template import ChatTemplate
from kafka_b200.tokenizer import ByteTokenizer
from kafka_b200.weights import generable_ids
from oracle.llama_fp32 import OracleLlama, masked_argmax


class CPUOracleProvider(LLMProvider):
    def __init__(self, cfg, weights, model: Optional[str] = None, tool_provider: Optional[Any] = None,
                 default_max_tokens: int = 32, generable_only: bool = True):
        super().__init__(tool_provider=tool_provider)
        self.cfg = cfg
        self.model = model or f"cpu-oracle/{cfg.name}"
        self.oracle = OracleLlama(cfg, weights)
        self.tokenizer = ByteTokenizer(cfg.vocab)
        self.template = ChatTemplate(self.tokenizer)
        self.default_max_tokens = default_max_tokens
        self.allowed = generable_ids(cfg) if generable_only else None
        self.tokens_generated = 0

    def generate_ids(self, prompt: List[int], max_tokens: int) -> (List[int], str):
        toks, _ = self.oracle.greedy(torch.tensor(prompt), max_tokens, allowed=self.allowed,
                                     eos=self.tokenizer.eot_id)
        self.tokens_generated += len(toks)
        if toks and toks[-1] == self.tokenizer.eot_id:
            return toks[:-1], "stop"
        return toks, "length"

    async def stream_completion(self, messages, *, temperature=None, max_tokens=None, stop=None, **kwargs):
        validated = self.validate_messages(messages)
        model = kwargs.pop("model", None) or self.model
        if temperature not in (None, 0, 0.0):
            raise ValueError("CPUOracleProvider is a greedy oracle: temperature must be 0")
        cid = f"chatcmpl-{uuid.uuid4().hex[:24]}"
        tools = await self.get_tools()
        prompt = self.template.render(validated, tools=tools or None)
        ids, reason = self.generate_ids(prompt, max_tokens or self.default_max_tokens)
        yield StreamChunk(role="assistant", model=model, id=cid)
        dec = self.tokenizer.stream_decoder()
        for t in ids:
            piece = dec.push(t)
            if piece:
                yield StreamChunk(content=piece, model=model, id=cid)
        tail = dec.flush()
        if tail:
            yield StreamChunk(content=tail, model=model, id=cid)
        yield StreamChunk(finish_reason=reason, model=model, id=cid)

    async def completion(self, messages, *, temperature=None, max_tokens=None, stop=None, **kwargs):
        parts, finish, cid, model = [], None, None, None
        async for ch in self.stream_completion(messages, temperature=temperature, max_tokens=max_tokens, **kwargs):
            if ch.content:
                parts.append(ch.content)
            finish, cid, model = ch.finish_reason or finish, ch.id, ch.model
        return CompletionResponse(content="".join(parts), finish_reason=finish, model=model, id=cid)

    async def close(self) -> None:
        return None


class BatchedOracle:
    """The same greedy oracle for MANY conversations at once (BASELINE config 2 at its stated size: 64
    threads x 8 turns x 32 tokens on TinyLlama-1.1B shapes would take the one-conversation-at-a-time provider
    above the better part of an hour on the host cores).  Two things differ, neither in the arithmetic per
    token: the new tokens of all conversations are packed for the projections (OracleLlama.forward_packed /
    decode_batch), and every conversation's fp32 K/V is kept between calls under a key, so a follow-up turn
    only runs the tokens after the longest common prefix with what the key has seen.  Same tokenizer, chat
    template, weights and candidate set as CPUOracleProvider; tests/test_oracle_and_abi.py checks that both
    produce the same text."""

    def __init__(self, cfg, weights, generable_only: bool = True):
        self.cfg = cfg
        self.oracle = OracleLlama(cfg, weights)
        self.tokenizer = ByteTokenizer(cfg.vocab)
        self.template = ChatTemplate(self.tokenizer)
        self.allowed = generable_ids(cfg) if generable_only else None
        self._state = {}          # key -> (tokens whose K/V is cached, KVCache)
        self.tokens_generated = 0
        self.tokens_prefilled = 0

    def generate_ids(self, keys: List[str], prompts: List[List[int]], max_tokens: int):
        """-> [(ids without a trailing eot, finish_reason)] in the order of `keys`."""
        caches, news = [], []
        for key, prompt in zip(keys, prompts):
            seen, cache = self._state.get(key, ([], None))
            if cache is None:
                cache = self.oracle.new_cache()
            n = 0
            lim = min(len(seen), len(prompt) - 1)      # at least one token must run to produce logits
            while n < lim and seen[n] == prompt[n]:
                n += 1
            cache.truncate(n)
            caches.append(cache)
            news.append(torch.tensor(prompt[n:]))
            self.tokens_prefilled += len(prompt) - n
        logits = self.oracle.forward_packed(news, caches)
        n = len(keys)
        out: List[List[int]] = [[] for _ in range(n)]
        reason = ["length"] * n
        active = list(range(n))
        eot = self.tokenizer.eot_id
        for step in range(max_tokens):
            nxt = []
            for row, i in enumerate(active):
                t = masked_argmax(logits[row], self.allowed)
                out[i].append(t)
                if t == eot:
                    reason[i] = "stop"
                else:
                    nxt.append(i)
            active = nxt
            if not active or step == max_tokens - 1:
                break
            logits = self.oracle.decode_batch(torch.tensor([out[i][-1] for i in active]), [caches[i] for i in active])
        res = []
        for i, (key, prompt) in enumerate(zip(keys, prompts)):
            self._state[key] = (list(prompt) + out[i][:-1], caches[i])   # the last id was sampled, never fed
            self.tokens_generated += len(out[i])
            ids = out[i][:-1] if reason[i] == "stop" else out[i]
            res.append((ids, reason[i]))
        return res

    def complete(self, keys: List[str], conversations, max_tokens: int, tools=None):
        """conversations: one message list per key -> [(text, finish_reason)]."""
        prompts = [self.template.render(m, tools=tools or None) for m in conversations]
        out = []
        for ids, reason in self.generate_ids(keys, prompts, max_tokens):
            dec = self.tokenizer.stream_decoder()
            text = "".join(dec.push(t) for t in ids) + dec.flush()
            out.append((text, reason))
        return out
