"""Provider factory keyed on the `model` string, and the process-wide engine registry.

The reference constructs its provider at one site (/root/reference/src/kafka/v1.py:177-181) and, on
the threaded agent endpoint, does so once per request (/root/reference/server.py:237-245). Engines are
therefore process-wide singletons behind cheap per-request provider objects: `make_provider` looks up
(or lazily builds) the engine for (model, devices) and wraps it.

Environment (same style as the reference's env-var configuration):
  KAFKA_B200_GPUS        comma-separated device indices (default: "0")
  KAFKA_B200_SEED        weight seed (default 0);  KAFKA_B200_INIT  "iid" | "margin" (default iid)
  KAFKA_B200_KV_FRACTION fraction of free HBM given to the KV pools (default 0.85)
  KAFKA_B200_MAX_SEQS    concurrent sequences per GPU (default 512)
  KAFKA_B200_MAX_TOKENS  packed tokens per prefill call (default 4096: the fastest chunk, tools/prefill_profile.py)
  KAFKA_B200_CHECKPOINTS comma-separated name=dir pairs: Hugging Face Llama checkpoints (safetensors) served under
                         `name` with their own weights and tokenizer instead of seeded random ones (checkpoint.py)
"""
from __future__ import annotations

import os
import threading
from typing import Any, Dict, List, Optional, Tuple

from .model_config import LlamaConfig, get_config, is_local_model  # noqa: F401  (re-exported)
from .tokenizer import ByteTokenizer

_lock = threading.Lock()
_engines: Dict[Tuple[str, Tuple[int, ...], int, str], Any] = {}
_tokenizers: Dict[str, Any] = {}


def tokenizer_for(cfg: LlamaConfig):
    """The model's tokenizer: its checkpoint's tokenizer.json when it is served from one, else the byte tokenizer."""
    from . import checkpoint
    tok = _tokenizers.get(cfg.name)
    if tok is None:
        tok = _tokenizers[cfg.name] = checkpoint.load_tokenizer(checkpoint.checkpoint_dir(cfg), cfg)
    return tok


def _config(model: str) -> LlamaConfig:
    from . import checkpoint
    checkpoint.register_from_env()
    return get_config(model)


def _devices() -> List[int]:
    return [int(x) for x in os.environ.get("KAFKA_B200_GPUS", "0").split(",") if x.strip() != ""]


def build_engine(cfg: LlamaConfig, device_index: int, seed: int = 0, init: str = "iid",
                 n_blocks: Optional[int] = None, max_seqs: Optional[int] = None,
                 max_tokens: Optional[int] = None, weights=None, generable_only: bool = True):
    """One replica on one GPU: weights (seeded) -> EngineCore -> GpuStepper -> Engine thread."""
    import torch
    from .engine import Engine
    from .engine_core import EngineCore
    from .gpu_stepper import GpuStepper
    from . import checkpoint
    from .weights import fuse_for_engine, generate_fused_on_device, generate_weights

    dev = torch.device("cuda", device_index)
    torch.cuda.set_device(dev)
    max_seqs = max_seqs or int(os.environ.get("KAFKA_B200_MAX_SEQS", "512"))
    max_tokens = max(max_tokens or int(os.environ.get("KAFKA_B200_MAX_TOKENS", "4096")), max_seqs)
    if weights is not None:
        fused = fuse_for_engine(cfg, weights, dev)
    elif checkpoint.checkpoint_dir(cfg):
        fused = checkpoint.load_fused(checkpoint.checkpoint_dir(cfg), cfg, dev)      # real weights, layer by layer
    elif cfg.weight_bytes() > (4 << 30) and init == "iid":
        fused = generate_fused_on_device(cfg, seed, dev)       # large shapes: draw on the GPU
    else:
        fused = fuse_for_engine(cfg, generate_weights(cfg, seed=seed, init=init), dev)
    if n_blocks is None:
        torch.cuda.synchronize()
        torch.cuda.empty_cache()
        free, _ = torch.cuda.mem_get_info(dev)
        frac = float(os.environ.get("KAFKA_B200_KV_FRACTION", "0.85"))
        n_blocks = max(8, int(free * frac) // EngineCore.kv_block_bytes(cfg))
    core = EngineCore(cfg, fused, dev, n_blocks=n_blocks, max_tokens=max_tokens, max_seqs=max_seqs)
    return Engine(GpuStepper(core, tokenizer_for(cfg), generable_only=generable_only),
                  name=f"kb-engine-gpu{device_index}")


def get_engine(model: str, **opts):
    """Process-wide engine for `model` on the configured devices (sharded when more than one)."""
    from .engine import ShardedEngine
    cfg = _config(model)
    devs = tuple(opts.pop("devices", None) or _devices())
    seed = int(opts.pop("seed", os.environ.get("KAFKA_B200_SEED", "0")))
    init = opts.pop("init", os.environ.get("KAFKA_B200_INIT", "iid"))
    key = (cfg.name, devs, seed, init)
    with _lock:
        eng = _engines.get(key)
        if eng is None:
            replicas = [build_engine(cfg, d, seed=seed, init=init, **opts) for d in devs]
            eng = replicas[0] if len(replicas) == 1 else ShardedEngine(replicas)
            _engines[key] = eng
        return eng


def shutdown_engines() -> None:
    with _lock:
        for eng in _engines.values():
            eng.shutdown()
        _engines.clear()
        _tokenizers.clear()


def make_provider(model: str, tool_provider: Optional[Any] = None, thread_id: Optional[str] = None,
                  default_temperature: float = 0.7, default_max_tokens: Optional[int] = None,
                  default_top_p: Optional[float] = None, **engine_opts):
    """The equivalent of `PortkeyLLMProvider(model=..., tool_provider=...)` at
    /root/reference/src/kafka/v1.py:177 for model names this repo serves locally:

        "echo"                      -> EchoLLMProvider (no model; host plumbing / BASELINE config 1)
        "llama-3-8b", "b200/<name>" -> B200LLMProvider on the process-wide engine

    Anything else is not ours: the caller keeps constructing the reference's remote provider.
    """
    from .provider import B200LLMProvider, EchoLLMProvider
    if model == "echo":
        return EchoLLMProvider(model=model, tool_provider=tool_provider)
    cfg = _config(model)  # KeyError for names that are not local
    engine = get_engine(model, **engine_opts)
    return B200LLMProvider(engine, tokenizer_for(cfg), model=model, tool_provider=tool_provider,
                           default_temperature=default_temperature, default_max_tokens=default_max_tokens,
                           thread_id=thread_id, default_top_p=default_top_p)
