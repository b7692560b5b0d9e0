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
