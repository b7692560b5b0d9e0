"""Local tool provider — the slice of the reference's tool layer the completion path touches.

The provider calls exactly one thing besides the engine: `tool_provider.get_tools()`
(/root/reference/src/llm/base.py:153-163, used at /root/reference/src/llm/portkey.py:317), and the
agent loop streams results through `run_tool_stream` (/root/reference/src/tools/agent.py:677-803).
This module re-creates that surface for *local* Python tools (sync, async, async-generator handlers,
/root/reference/src/tools/types.py:187-219). MCP servers and sandbox tools need the network and are
out of scope.
"""
from __future__ import annotations

import asyncio
import inspect
from dataclasses import dataclass
from typing import Any, AsyncGenerator, Callable, Dict, List, Optional


@dataclass
class ToolResultChunk:
    tool_call_id: str
    tool_name: str
    delta: str
    is_complete: bool = False


class Tool:
    def __init__(self, name: str, description: str, parameters: Dict[str, Any],
                 handler: Optional[Callable[..., Any]] = None):
        self.name, self.description, self.parameters, self.handler = name, description, parameters, handler

    def to_openai(self) -> Dict[str, Any]:
        return {"type": "function",
                "function": {"name": self.name, "description": self.description, "parameters": self.parameters}}

    async def run_stream(self, arguments: Dict[str, Any]) -> AsyncGenerator[str, None]:
        if self.handler is None:
            raise RuntimeError(f"No handler registered for tool '{self.name}'")
        if inspect.isasyncgenfunction(self.handler):
            async for chunk in self.handler(**arguments):
                yield chunk
        elif asyncio.iscoroutinefunction(self.handler):
            yield str(await self.handler(**arguments))
        else:
            yield str(self.handler(**arguments))


class LocalToolProvider:
    """get_tools / add_tool / run_tool_stream for local tools, with the reference's chunk protocol:
    output chunks with is_complete=False, then one empty chunk with is_complete=True; failures become a
    single `Error: ...` chunk with is_complete=True (agent.py:797-803)."""

    def __init__(self, tools: Optional[List[Tool]] = None):
        self._tools: Dict[str, Tool] = {}
        for t in tools or []:
            self.add_tool(t)

    def add_tool(self, tool: Tool) -> None:
        self._tools[tool.name] = tool

    def get_tool(self, name: str) -> Optional[Tool]:
        return self._tools.get(name)

    async def connect(self) -> None:
        return None

    async def disconnect(self) -> None:
        return None

    async def get_tools(self) -> List[Dict[str, Any]]:
        return [t.to_openai() for t in self._tools.values()]

    async def run_tool_stream(self, name: str, arguments: Dict[str, Any],
                              tool_call_id: str) -> AsyncGenerator[ToolResultChunk, None]:
        tool = self._tools.get(name)
        if tool is None:
            yield ToolResultChunk(tool_call_id, name, f"Error: Tool not found: {name}", True)
            return
        try:
            async for chunk in tool.run_stream(arguments):
                yield ToolResultChunk(tool_call_id, name, chunk, False)
            yield ToolResultChunk(tool_call_id, name, "", True)
        except Exception as e:
            yield ToolResultChunk(tool_call_id, name, f"Error: {e}", True)
