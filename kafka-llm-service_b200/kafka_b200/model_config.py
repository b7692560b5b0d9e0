"""Llama-family model shapes served by the B200 provider.

The reference has no notion of an architecture: ``model`` is a string forwarded to the gateway
(/root/reference/src/llm/portkey.py:267) and routed by substring
(/root/reference/src/llm/utils.py:11-29, where Llama names fall through to ``"unknown"``). The
names registered here are the ones the provider factory claims; everything else keeps going to the
reference's remote provider.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, asdict
from typing import Dict, List, Optional, Tuple


@dataclass(frozen=True)
class LlamaConfig:
    name: str
    hidden: int
    n_layers: int
    n_heads: int
    n_kv_heads: int
    head_dim: int
    ffn: int
    vocab: int
    rope_theta: float = 10000.0
    rms_eps: float = 1e-5
    max_pos: int = 8192
    # Llama-3.1's RoPE scaling rule: (factor, low_freq_factor, high_freq_factor, original_max_position_embeddings);
    # None = plain RoPE. The rule only reshapes the head_dim / 2 frequencies (inv_freq below); nothing else changes.
    rope_llama3: Optional[Tuple[float, float, float, int]] = None

    @property
    def group(self) -> int:
        return self.n_heads // self.n_kv_heads

    @property
    def q_dim(self) -> int:
        return self.n_heads * self.head_dim

    @property
    def kv_dim(self) -> int:
        return self.n_kv_heads * self.head_dim

    @property
    def qkv_dim(self) -> int:
        return self.q_dim + 2 * self.kv_dim

    @property
    def kv_bytes_per_token(self) -> int:
        """K+V bytes one token occupies over all layers (bf16)."""
        return 2 * self.n_layers * self.kv_dim * 2

    def matmul_params(self) -> int:
        """Parameters of the per-token projections (QKV, O, gate, up, down), all layers."""
        per_layer = self.hidden * self.qkv_dim + self.q_dim * self.hidden + 3 * self.hidden * self.ffn
        return per_layer * self.n_layers

    def lm_head_params(self) -> int:
        return self.vocab * self.hidden

    def weight_bytes(self) -> int:
        """All bf16 weights (embed + layers + norms + lm_head)."""
        norms = (2 * self.n_layers + 1) * self.hidden
        return 2 * (self.matmul_params() + 2 * self.vocab * self.hidden + norms)

    def to_dict(self) -> Dict:
        return asdict(self)


def inv_freq(cfg: LlamaConfig) -> List[float]:
    """The head_dim / 2 rotary frequencies, in float32 steps like `transformers` computes them
    (modeling_rope_utils: theta^(-2i/d), then for the "llama3" rule: wavelengths beyond original_max / low_freq_factor
    are slowed by `factor`, those below original_max / high_freq_factor are kept, the band between is blended)."""
    import torch
    d = cfg.head_dim
    f = 1.0 / (cfg.rope_theta ** (torch.arange(0, d, 2, dtype=torch.int64).to(torch.float32) / d))
    if cfg.rope_llama3 is not None:
        factor, low, high, orig = cfg.rope_llama3
        wavelen = 2 * math.pi / f
        scaled = torch.where(wavelen > orig / low, f / factor, f)
        smooth = (orig / wavelen - low) / (high - low)
        blended = (1 - smooth) * scaled / factor + smooth * scaled
        medium = ~(wavelen < orig / high) & ~(wavelen > orig / low)
        f = torch.where(medium, blended, scaled)
    return [float(x) for x in f]


REGISTRY: Dict[str, LlamaConfig] = {}


def register(cfg: LlamaConfig) -> LlamaConfig:
    REGISTRY[cfg.name] = cfg
    return cfg


# Small shapes for tests / smoke (vocab >= 264 so the byte tokenizer's ids fit).
register(LlamaConfig("tiny-test", hidden=256, n_layers=2, n_heads=4, n_kv_heads=2, head_dim=64, ffn=512,
                     vocab=512, rope_theta=10000.0, max_pos=2048))
register(LlamaConfig("tiny-test-hd128", hidden=512, n_layers=2, n_heads=4, n_kv_heads=1, head_dim=128, ffn=1024,
                     vocab=512, rope_theta=500000.0, max_pos=4096))
# Public architectures.
register(LlamaConfig("tinyllama-1.1b", hidden=2048, n_layers=22, n_heads=32, n_kv_heads=4, head_dim=64,
                     ffn=5632, vocab=32000, rope_theta=10000.0, max_pos=2048))
register(LlamaConfig("llama-3-8b", hidden=4096, n_layers=32, n_heads=32, n_kv_heads=8, head_dim=128,
                     ffn=14336, vocab=128256, rope_theta=500000.0, max_pos=8192))
register(LlamaConfig("llama-3-70b", hidden=8192, n_layers=80, n_heads=64, n_kv_heads=8, head_dim=128,
                     ffn=28672, vocab=128256, rope_theta=500000.0, max_pos=8192))


def get_config(name: str) -> LlamaConfig:
    key = name.lower()
    for prefix in ("b200/", "cpu-oracle/"):
        if key.startswith(prefix):
            key = key[len(prefix):]
    if key not in REGISTRY:
        raise KeyError(f"unknown local model {name!r}; known: {sorted(REGISTRY)}")
    return REGISTRY[key]


def is_local_model(name: str) -> bool:
    try:
        from . import checkpoint            # names given through $KAFKA_B200_CHECKPOINTS count as local models
        checkpoint.register_from_env()
        get_config(name)
        return True
    except KeyError:
        return False
