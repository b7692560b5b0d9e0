"""The shape of the reference's `examples/agent.py` (/root/reference/examples/agent.py:17-158: an agent with two
local tools, once stateless and once on a server-side thread) on this repo's local provider — no gateway, no
network. Run it from the repository root:

    python examples/agent.py                      # model "echo": no GPU, shows the plumbing
    python examples/agent.py --model llama-3-8b   # a B200: seeded random weights (text is noise, the path is real)
    KAFKA_B200_CHECKPOINTS=my-llama=/models/Meta-Llama-3-8B-Instruct python examples/agent.py --model my-llama

Stateless: `StreamingAgent.run(messages)`. Threaded: `ThreadRunner.run(new_messages)` loads the thread's history
from SQLite, runs the agent, stores the new messages — and on a GPU the thread's KV stays resident, so the second
turn prefills only its new message (printed as `cached` tokens)."""
import argparse
import asyncio
import os
import sys
import tempfile

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "kafka-llm-service_b200"))

from kafka_b200.agent import StreamingAgent                  # noqa: E402
from kafka_b200.db import ThreadStore                        # noqa: E402
from kafka_b200.factory import make_provider, shutdown_engines   # noqa: E402
from kafka_b200.llm import Message                           # noqa: E402
from kafka_b200.orchestrator import ThreadRunner             # noqa: E402
from kafka_b200.tools import LocalToolProvider, Tool         # noqa: E402


def get_weather(city: str = "Paris") -> str:
    return f"It is 21 degrees and sunny in {city}."


async def add(a: float = 0, b: float = 0) -> str:
    return str(a + b)


def make_agent(model: str, thread_id=None):
    tools = LocalToolProvider([
        Tool("get_weather", "Current weather of a city", {"type": "object", "properties": {"city": {"type": "string"}}},
             get_weather),
        Tool("add", "Add two numbers", {"type": "object", "properties": {"a": {"type": "number"}, "b": {"type": "number"}}},
             add)])
    provider = make_provider(model, tool_provider=tools, thread_id=thread_id)
    compaction = provider.compaction_provider() if hasattr(provider, "compaction_provider") else None
    return StreamingAgent(provider, tools, system_prompt="You are a helpful assistant.",
                          context_compaction_provider=compaction)


async def show(events) -> str:
    text = ""
    async for ev in events:
        for ch in ev.get("choices") or []:
            piece = (ch.get("delta") or {}).get("content")
            if piece:
                text += piece
                print(piece, end="", flush=True)
        if ev.get("type") == "tool_result":
            print(f"\n[tool {ev.get('tool_name')}: {ev.get('delta')!r}]", end="")
    print()
    return text


async def main(model: str, max_tokens: int) -> None:
    print("--- stateless run")
    agent = make_agent(model)
    await show(agent.run([Message(role="user", content="What is the weather in Paris?")], model=model, temperature=0.0,
                         max_tokens=max_tokens))
    print("--- two turns on a thread")
    store = ThreadStore(os.path.join(tempfile.mkdtemp(prefix="kb-example-"), "threads.db"))
    for turn, text in enumerate(["Hello, remember the number 42.", "Which number did I mention?"]):
        agent = make_agent(model, thread_id="example-thread")
        runner = ThreadRunner(agent, store, "example-thread", engine=getattr(agent.llm_provider, "engine", None))
        await show(runner.run([{"role": "user", "content": text}], model=model, temperature=0.0, max_tokens=max_tokens))
        usage = getattr(agent.llm_provider, "usage_total", None)
        if usage:
            print(f"    turn {turn + 1}: prompt {usage['prompt_tokens']} tokens, {usage['cached_tokens']} of them from resident KV")
    print("stored messages:", len(await store.get_thread_messages("example-thread")))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="echo")
    ap.add_argument("--max-tokens", type=int, default=48)
    a = ap.parse_args()
    try:
        asyncio.run(main(a.model, a.max_tokens))
    finally:
        shutdown_engines()
