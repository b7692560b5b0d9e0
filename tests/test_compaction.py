"""Offline compaction (kafka_b200/compaction.py) against the reference's TruncationCompactionProvider on the
same random conversations, and the agent's overflow -> compact -> retry path on the fake engine."""
import asyncio
import random

import pytest

import ref_harness
from fakes import FakeStepper
from kafka_b200.agent import StreamingAgent
from kafka_b200.compaction import TruncationCompactionProvider, drop_orphans, safe_cut
from kafka_b200.engine import Engine
from kafka_b200.llm import LLMProviderError, Message
from kafka_b200.provider import B200LLMProvider
from kafka_b200.tokenizer import ByteTokenizer


def _conversation(rnd, n):
    """system message, then user / assistant turns, some with tool-call groups (1-3 calls + their results)."""
    msgs = [{"role": "system", "content": "S"}]
    k = 0
    while len(msgs) < n:
        msgs.append({"role": "user", "content": f"u{len(msgs)}"})
        if rnd.random() < 0.5:
            ids = [f"call_{k + j}" for j in range(rnd.randint(1, 3))]
            k += len(ids)
            msgs.append({"role": "assistant", "content": None if rnd.random() < 0.5 else "thinking",
                         "tool_calls": [{"id": i, "type": "function", "function": {"name": "f", "arguments": "{}"}}
                                        for i in ids]})
            for i in ids:
                msgs.append({"role": "tool", "content": "r", "tool_call_id": i, "name": "f"})
        if rnd.random() < 0.1:
            msgs.append({"role": "assistant", "content": ""})          # an empty one: must be dropped
        else:
            msgs.append({"role": "assistant", "content": f"a{len(msgs)}"})
    return msgs


def test_safe_cut_never_splits_a_tool_group():
    rnd = random.Random(3)
    for _ in range(50):
        msgs = _conversation(rnd, 60)[1:]
        for target in range(len(msgs) + 1):
            c = safe_cut(msgs, target)
            assert 0 <= c <= target
            if 0 < c < len(msgs):
                assert msgs[c].get("role") != "tool"
                assert not (msgs[c - 1].get("role") == "assistant" and msgs[c - 1].get("tool_calls"))
            kept = drop_orphans(msgs[c:])
            assert [m for m in kept if m["role"] == "tool"] == [m for m in msgs[c:] if m["role"] == "tool"]


@pytest.mark.skipif(not ref_harness.available(), reason="reference tree not available")
def test_truncation_equals_the_reference_on_random_conversations():
    ref_harness.load()
    from src.llm.context_compaction.base import find_safe_split_point, validate_message_structure
    from src.llm.context_compaction.v1 import TruncationCompactionProvider as Ref
    rnd = random.Random(7)
    for trial in range(60):
        msgs = _conversation(rnd, rnd.randint(5, 140))
        keep = rnd.choice([5, 20, 50])
        ours = asyncio.run(TruncationCompactionProvider(keep).compact([dict(m) for m in msgs], "S", "m"))
        theirs = asyncio.run(Ref(keep_count=keep).compact([dict(m) for m in msgs], "S", "m"))
        assert ours == theirs, trial
        body = msgs[1:]
        for target in (0, 1, len(body) // 2, len(body) - 1, len(body)):
            assert safe_cut(body, target) == find_safe_split_point(body, target)
        assert drop_orphans(body[7:]) == validate_message_structure(body[7:])


def test_token_budget_keeps_cutting_until_the_prompt_fits():
    rnd = random.Random(11)
    msgs = _conversation(rnd, 80)
    count = lambda ms: sum(10 + len(m.get("content") or "") for m in ms)
    full = count(msgs)
    out = asyncio.run(TruncationCompactionProvider(50, count_tokens=count, token_budget=full // 4).compact(msgs, "S", "m"))
    last = [m for m in msgs if not (m["role"] == "assistant" and not m.get("content") and not m.get("tool_calls"))][-1]
    assert out[0]["role"] == "system" and count(out) <= full // 4 and out[-1] == last
    announced = {tc["id"] for m in out for tc in (m.get("tool_calls") or [])}
    assert all(m["tool_call_id"] in announced for m in out if m["role"] == "tool")


def test_agent_overflow_compacts_and_retries_on_the_engine():
    """The provider raises the context-length error the reference's trigger matches; the agent compacts with the
    truncation provider (token budget = 3/4 of the context) and the retry completes."""
    tok = ByteTokenizer(512)
    eng = Engine(FakeStepper(eos_id=tok.eot_id, n_blocks=64, max_pos=1024, max_tokens=1024))
    provider = B200LLMProvider(eng, tok, model="tiny-test", default_temperature=0.0)
    compactor = provider.compaction_provider(keep_count=50, fraction=0.75)       # budget: 768 tokens, tools included
    history = [Message(role="system", content="SYS")]
    for i in range(30):
        history += [Message(role="user", content=f"question {i} " + "x" * 40),
                    Message(role="assistant", content=f"answer {i} " + "y" * 40)]
    history.append(Message(role="user", content="and now?"))
    try:
        assert len(provider.render(history, None)) > 1024

        async def run(agent):
            return [ev async for ev in agent.run(history, model="tiny-test", temperature=0.0, max_tokens=8)]
        with pytest.raises(LLMProviderError, match="maximum context length"):
            asyncio.run(run(StreamingAgent(provider, None)))                       # no compaction: the error surfaces
        events = asyncio.run(run(StreamingAgent(provider, None, context_compaction_provider=compactor)))
        assert events[-1]["type"] == "agent_done" and len(events[-1]["final_content"]) == 8
        assert eng.sched.stats["prompt_tokens"] <= 768
    finally:
        eng.shutdown()
