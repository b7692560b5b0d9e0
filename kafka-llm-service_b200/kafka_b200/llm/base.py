"""LLMProvider — structural twin of the reference's plugin ABC (/root/reference/src/llm/base.py:67-312).

Same constructor (`tool_provider=None`), same `tool_provider` property, `has_tools`, `get_tools`,
abstract `stream_completion` / `completion` with identical signatures, `validate_messages` raising
ValueError on an empty list, and `get_model_info`. When the reference package is importable the
concrete providers of this repo are additionally registered as virtual subclasses of its ABC
(see kafka_b200.llm.register_with_reference).
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from typing import Any, AsyncGenerator, Dict, List, Optional

from .types import CompletionResponse, Message, StreamChunk


class LLMProvider(ABC):
    def __init__(self, tool_provider: Optional[Any] = None):
        self._tool_provider = tool_provider

    @property
    def tool_provider(self) -> Optional[Any]:
        return self._tool_provider

    @tool_provider.setter
    def tool_provider(self, provider: Optional[Any]) -> None:
        self._tool_provider = provider

    def has_tools(self) -> bool:
        return self._tool_provider is not None

    async def get_tools(self) -> List[Dict[str, Any]]:
        """Tool schemas in OpenAI function format; [] without a tool provider (base.py:153-163)."""
        if self._tool_provider is None:
            return []
        return await self._tool_provider.get_tools()

    @abstractmethod
    async def stream_completion(self, messages: List[Message], *, temperature: Optional[float] = None,
                                max_tokens: Optional[int] = None, stop: Optional[List[str]] = None,
                                **kwargs: Any) -> AsyncGenerator[StreamChunk, None]:
        """Yield >= 1 chunk; the last one carries finish_reason; safe to abandon mid-stream."""
        yield StreamChunk()  # pragma: no cover

    @abstractmethod
    async def completion(self, messages: List[Message], *, temperature: Optional[float] = None,
                         max_tokens: Optional[int] = None, stop: Optional[List[str]] = None,
                         **kwargs: Any) -> CompletionResponse:
        ...  # pragma: no cover

    def validate_messages(self, messages: List[Message]) -> List[Message]:
        if not messages:
            raise ValueError("Messages list cannot be empty")
        return messages

    def get_model_info(self) -> Dict[str, Any]:
        return {}
