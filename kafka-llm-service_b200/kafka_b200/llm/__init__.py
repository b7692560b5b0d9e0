"""Provider plugin surface (mirror of /root/reference/src/llm/__init__.py)."""
from .types import Role, Message, StreamChunk, CompletionResponse, LLMProviderError
from .base import LLMProvider

__all__ = ["Role", "Message", "StreamChunk", "CompletionResponse", "LLMProviderError", "LLMProvider"]


def register_with_reference() -> bool:
    """If the reference package is importable (`src.llm.base`), register this repo's LLMProvider twin as
    a virtual subclass of the reference's ABC so `isinstance(p, src.llm.base.LLMProvider)` holds for
    every provider defined here. Returns False when the reference is not on sys.path."""
    try:
        from src.llm.base import LLMProvider as RefProvider  # type: ignore
    except Exception:
        return False
    RefProvider.register(LLMProvider)
    return True
