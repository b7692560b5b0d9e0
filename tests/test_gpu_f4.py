"""SURVEY.md §8(f) rank 4 on the real engine: the context-overflow path (provider error -> offline compaction ->
retry) and model-emitted tool calls streamed as StreamChunk.tool_calls deltas.

Reference behaviour mirrored: /root/reference/src/agents/base.py:234-271 (compaction retry),
/root/reference/src/llm/context_compaction/v1.py:242 (truncation), src/agents/base.py:286-334 (tool-call deltas).
"""
import asyncio
import json

import pytest
import torch

pytestmark = pytest.mark.gpu


def _build(cuda, n_blocks=128, max_seqs=8):
    from kafka_b200.factory import build_engine
    from kafka_b200.model_config import get_config
    from kafka_b200.provider import B200LLMProvider
    from kafka_b200.tokenizer import ByteTokenizer
    from kafka_b200.weights import generate_weights
    from oracle.cpu_provider import CPUOracleProvider
    cfg = get_config("tiny-test")                       # max_pos 2048
    w = generate_weights(cfg, seed=0, init="margin")
    eng = build_engine(cfg, cuda.index or 0, n_blocks=n_blocks, max_seqs=max_seqs, max_tokens=2048, weights=w)
    gpu = B200LLMProvider(eng, ByteTokenizer(cfg.vocab), model="tiny-test", default_temperature=0.0)
    return cfg, eng, gpu, CPUOracleProvider(cfg, w)


def test_overflow_is_compacted_and_the_retry_matches_the_oracle(cuda):
    from kafka_b200.agent import StreamingAgent
    from kafka_b200.llm import LLMProviderError, Message
    cfg, eng, gpu, cpu = _build(cuda)
    history = [Message(role="system", content="You are terse.")]
    for i in range(40):
        history += [Message(role="user", content=f"question {i}: " + "x" * 30),
                    Message(role="assistant", content=f"answer {i}: " + "y" * 30)]
    history.append(Message(role="user", content="and the last one?"))
    try:
        async def run(agent):
            return [ev async for ev in agent.run(history, model="tiny-test", temperature=0.0, max_tokens=16,
                                                 thread_id="overflow-1")]
        plain = StreamingAgent(gpu, None)
        asyncio.run(plain.tool_provider.get_tools())
        with pytest.raises(LLMProviderError, match="maximum context length"):
            asyncio.run(run(plain))                                            # the trigger phrase, no compaction
        before = eng.stats()["prompt_tokens"]
        agent = StreamingAgent(gpu, None, context_compaction_provider=gpu.compaction_provider())
        events = asyncio.run(run(agent))
        assert events[-1]["type"] == "agent_done" and events[-1]["reason"] == "text_response"
        prompt_tokens = eng.stats()["prompt_tokens"] - before
        assert 0 < prompt_tokens <= int(0.75 * cfg.max_pos)
        # the same compaction applied by hand, answered by the fp32 oracle: identical text
        compacted = asyncio.run(gpu.compaction_provider().compact([m.to_dict() for m in history], "", "tiny-test"))
        assert 2 < len(compacted) < len(history) and compacted[0]["role"] == "system"
        cpu.tool_provider = agent.tool_provider                                # same tool schemas in the prompt
        want = asyncio.run(cpu.completion([Message(**d) for d in compacted], temperature=0.0, max_tokens=16))
        assert events[-1]["final_content"] == want.content
    finally:
        eng.shutdown()


def test_model_emitted_tool_call_block_becomes_tool_call_deltas(cuda):
    """A random-init model never writes the tool-call syntax on its own, so the sampler's candidate mask plays the
    script: every step runs the full forward pass and the sampling kernel on the GPU, restricted to ONE candidate
    id (kb_sampling.allowed_lo / allowed_hi). The provider must turn the streamed text into tool_calls deltas,
    the agent must run the tool and go on to a second completion."""
    from kafka_b200.agent import StreamingAgent
    from kafka_b200.gpu_stepper import GpuStepper
    from kafka_b200.llm import Message
    from kafka_b200.tools import LocalToolProvider, Tool
    cfg, eng, gpu, _ = _build(cuda)
    call = '[{"id":"call_9","type":"function","function":{"name":"add","arguments":"{\\"a\\": 2, \\"b\\": 40}"}}]'
    scripts = [list(("Let me add.\n<tool_calls>" + call + "</tool_calls>").encode()), list(b"It is 42.")]
    stepper = eng.stepper
    assert isinstance(stepper, GpuStepper)
    state = {"req": {}, "order": []}
    orig_run_plan = stepper.run_plan

    def scripted_run_plan(p, read_back=True):
        # one conversation at a time in this test: the k-th sampling call of request r emits scripts[r][k]
        rid = state["current"]
        k = state["req"].get(rid, 0)
        state["req"][rid] = k + 1
        s = scripts[state["order"].index(rid)]
        t = s[k] if k < len(s) else stepper.eos_id
        stepper.cand = (t, t + 1, -1, -1)
        return orig_run_plan(p, read_back)

    def prefill_launch(items, parity):
        rid = items[0].req.req_id
        if rid not in state["order"]:
            state["order"].append(rid)
        state["current"] = rid
        if not any(it.is_last_chunk for it in items):
            return orig_run_plan(stepper.plan_prefill(items, parity))
        return scripted_run_plan(stepper.plan_prefill(items, parity))

    def decode_launch(reqs, parity):
        state["current"] = reqs[0].req_id
        return scripted_run_plan(stepper.plan_decode(reqs, parity))
    stepper.prefill_launch, stepper.decode_launch = prefill_launch, decode_launch

    ran = []

    def add(a: int, b: int) -> str:
        ran.append((a, b))
        return str(a + b)
    tools = LocalToolProvider([Tool("add", "add two integers", {"type": "object", "properties": {
        "a": {"type": "integer"}, "b": {"type": "integer"}}}, add)])
    gpu.tool_provider = tools
    agent = StreamingAgent(gpu, tools)
    try:
        async def go():
            return [ev async for ev in agent.run([Message(role="user", content="what is 2 + 40?")], model="tiny-test",
                                                 temperature=0.0, max_tokens=300, thread_id="tool-1")]
        events = asyncio.run(go())
        finishes = [e["choices"][0]["finish_reason"] for e in events if e.get("choices") and e["choices"][0]["finish_reason"]]
        assert finishes == ["tool_calls", "stop"]
        deltas = [tc for e in events if e.get("choices") for tc in (e["choices"][0]["delta"].get("tool_calls") or [])]
        assert deltas[0]["id"] == "call_9" and deltas[0]["function"]["name"] == "add"
        assert json.loads("".join(d["function"].get("arguments", "") for d in deltas)) == {"a": 2, "b": 40}
        assert ran == [(2, 40)]
        results = [e for e in events if e.get("type") == "tool_result" and e["is_complete"]]
        assert len(results) == 1
        text = "".join(e["choices"][0]["delta"].get("content") or "" for e in events if e.get("choices"))
        # the text before the block streams as it is generated (its newline included); the block itself never does
        assert text == "Let me add.\nIt is 42." and events[-1]["final_content"] == "It is 42."
        # the second completion found the first round resident (prefix reuse across the tool round)
        assert eng.stats()["cached_tokens"] > 0
    finally:
        eng.shutdown()
