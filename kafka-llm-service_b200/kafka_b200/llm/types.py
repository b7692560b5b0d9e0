"""Wire types of the provider plugin — the structural twin of /root/reference/src/llm/types.py.

Same names, fields, defaults and helper behaviour as the reference (Message :29-68, StreamChunk
:71-110, CompletionResponse :113-148, LLMProviderError :151-184) so that objects of either family can
be handed to code written against the other (the reference's Agent only reads attributes).
"""
from __future__ import annotations

from enum import Enum
from typing import Any, Dict, List, Optional

from pydantic import BaseModel


class Role(str, Enum):
    SYSTEM = "system"
    USER = "user"
    ASSISTANT = "assistant"
    TOOL = "tool"


class Message(BaseModel):
    """One conversation message, OpenAI-shaped. ``content`` may be None (assistant tool-call turns)."""
    role: str
    content: Optional[str] = None
    name: Optional[str] = None
    tool_calls: Optional[List[Dict[str, Any]]] = None
    tool_call_id: Optional[str] = None

    def to_dict(self) -> Dict[str, Any]:
        """Dict without the None fields (reference: types.py:54-68)."""
        out: Dict[str, Any] = {"role": self.role}
        for key in ("content", "name", "tool_calls", "tool_call_id"):
            val = getattr(self, key)
            if val is not None:
                out[key] = val
        return out


class StreamChunk(BaseModel):
    """One increment of a streamed completion. Absent fields are None, never ''."""
    content: Optional[str] = None
    role: Optional[str] = None
    tool_calls: Optional[List[Dict[str, Any]]] = None
    finish_reason: Optional[str] = None
    model: Optional[str] = None
    id: Optional[str] = None

    @property
    def delta(self) -> str:
        return self.content or ""

    @property
    def is_final(self) -> bool:
        return self.finish_reason is not None


class CompletionResponse(BaseModel):
    content: Optional[str] = None
    role: str = "assistant"
    finish_reason: Optional[str] = None
    model: Optional[str] = None
    id: Optional[str] = None
    usage: Optional[Dict[str, int]] = None
    tool_calls: Optional[List[Dict[str, Any]]] = None

    def to_message(self) -> Message:
        return Message(role=self.role, content=self.content, tool_calls=self.tool_calls)


class LLMProviderError(Exception):
    """Provider failure; ``str()`` renders ``[provider] message (status: N)`` like the reference."""

    def __init__(self, message: str, status_code: Optional[int] = None, provider: Optional[str] = None,
                 original_error: Optional[Exception] = None):
        super().__init__(message)
        self.message = message
        self.status_code = status_code
        self.provider = provider
        self.original_error = original_error

    def __str__(self) -> str:
        parts = []
        if self.provider:
            parts.append(f"[{self.provider}]")
        parts.append(self.message)
        if self.status_code:
            parts.append(f"(status: {self.status_code})")
        return " ".join(parts)


_osa = object.__setattr__


def content_chunk(content: str, model: Optional[str], id: Optional[str]) -> StreamChunk:
    """StreamChunk(content=..., model=..., id=...) without pydantic's validation machinery.

    One chunk is built per generated token per stream (tens of thousands per second); the values are
    produced right here and need no validation. Equivalent to ``StreamChunk.model_construct`` with every
    field given (tests/test_host_path.py checks equality with the validated constructor)."""
    m = StreamChunk.__new__(StreamChunk)
    _osa(m, "__dict__", {"content": content, "role": None, "tool_calls": None, "finish_reason": None,
                         "model": model, "id": id})
    _osa(m, "__pydantic_fields_set__", {"content", "model", "id"})
    _osa(m, "__pydantic_extra__", None)
    _osa(m, "__pydantic_private__", None)
    return m
