"""Provider plugin surface (mirror of /root/reference/src/llm/__init__.py)."""
from .types import Role, Message, StreamChunk, CompletionResponse, LLMProviderError
from .base import LLMProvider

__all__ = ["Role", "Message", "StreamChunk", "CompletionResponse", "LLMProviderError", "LLMProvider"]
