"""Real weights: a Hugging Face Llama checkpoint directory (config.json + *.safetensors [+ index] [+ tokenizer.json])
-> this package's LlamaConfig, canonical / fused tensors and tokenizer.

The reference never holds weights — `model` is a string it forwards to a gateway
(/root/reference/src/llm/portkey.py:267) — so there is no counterpart; this is what a deployment needs in order to
serve an actual model through the same `model` field:

    KAFKA_B200_CHECKPOINTS="llama-3-8b-instruct=/models/Meta-Llama-3-8B-Instruct"   (comma-separated name=dir pairs)

registers each directory under its name (factory.py reads the variable; `register_checkpoint` does the same from code).
Tensor names follow `transformers`' LlamaForCausalLM; its q / k projections are already in the rotate-half layout this
engine and the oracle use (tests/test_checkpoint.py: the oracle on loaded weights equals `transformers`' own forward).

RoPE: plain, or Llama-3.1's "llama3" scaling rule (the frequencies are reshaped on the host, model_config.inv_freq, and
handed to the kernel as a table). Unsupported checkpoint features are refused by name instead of being ignored: other
RoPE scaling rules, attention / MLP biases, a head_dim other than 64 / 128, non-Llama architectures.
"""
from __future__ import annotations

import json
import os
from pathlib import Path
from typing import Dict, Iterable, Iterator, Optional, Tuple

from .model_config import REGISTRY, LlamaConfig, register

CHECKPOINTS: Dict[str, str] = {}        # registered model name -> directory


class CheckpointError(ValueError):
    pass


def config_from_hf(path: str, name: Optional[str] = None) -> LlamaConfig:
    p = Path(path)
    try:
        hf = json.loads((p / "config.json").read_text())
    except OSError as e:
        raise CheckpointError(f"{path}: no readable config.json ({e})") from e
    arch = hf.get("architectures") or []
    if hf.get("model_type") != "llama" and "LlamaForCausalLM" not in arch:
        raise CheckpointError(f"{path}: model_type {hf.get('model_type')!r} / architectures {arch} — only Llama is served")
    # RoPE: `rope_theta` + `rope_scaling` (transformers 4) or `rope_parameters` {rope_theta, rope_type, ...} (5)
    rp = hf.get("rope_parameters") or {}
    scaling = hf.get("rope_scaling") or ({k: v for k, v in rp.items() if k != "rope_theta"}
                                         if rp.get("rope_type", "default") != "default" else None)
    llama3 = None
    if scaling:
        kind = scaling.get("rope_type", scaling.get("type"))
        if kind != "llama3":
            raise CheckpointError(f"{path}: rope_scaling {scaling} is not implemented (plain RoPE and the llama3 rule only)")
        llama3 = (float(scaling["factor"]), float(scaling["low_freq_factor"]), float(scaling["high_freq_factor"]),
                  int(scaling["original_max_position_embeddings"]))
    theta = float(rp.get("rope_theta", hf.get("rope_theta", 10000.0)))
    for key in ("attention_bias", "mlp_bias"):
        if hf.get(key):
            raise CheckpointError(f"{path}: {key}=true is not implemented (the projections have no bias term)")
    if hf.get("hidden_act", "silu") != "silu":
        raise CheckpointError(f"{path}: hidden_act {hf['hidden_act']!r} is not implemented (SwiGLU with SiLU only)")
    hidden, heads = int(hf["hidden_size"]), int(hf["num_attention_heads"])
    head_dim = int(hf.get("head_dim") or hidden // heads)
    if head_dim not in (64, 128):
        raise CheckpointError(f"{path}: head_dim {head_dim} — the attention kernels take 64 or 128")
    cfg = LlamaConfig(name=(name or p.name).lower(), hidden=hidden, n_layers=int(hf["num_hidden_layers"]), n_heads=heads,
                      n_kv_heads=int(hf.get("num_key_value_heads") or heads), head_dim=head_dim,
                      ffn=int(hf["intermediate_size"]), vocab=int(hf["vocab_size"]),
                      rope_theta=theta, rms_eps=float(hf.get("rms_norm_eps", 1e-5)),
                      max_pos=int(hf.get("max_position_embeddings", 8192)), rope_llama3=llama3)
    if cfg.vocab % 8:
        raise CheckpointError(f"{path}: vocab_size {cfg.vocab} must be a multiple of 8")
    return cfg


def register_checkpoint(path: str, name: Optional[str] = None) -> LlamaConfig:
    """Make `name` (default: the directory's name) a local model backed by the checkpoint at `path`."""
    cfg = config_from_hf(path, name)
    register(cfg)
    CHECKPOINTS[cfg.name] = str(path)
    return cfg


def register_from_env(var: str = "KAFKA_B200_CHECKPOINTS") -> None:
    for item in os.environ.get(var, "").split(","):
        item = item.strip()
        if not item:
            continue
        name, sep, path = item.partition("=")
        if not sep:
            name, path = "", name
        key = (name or Path(path).name).lower()
        if key not in CHECKPOINTS or CHECKPOINTS[key] != path:
            register_checkpoint(path, name or None)


def checkpoint_dir(cfg: LlamaConfig) -> Optional[str]:
    return CHECKPOINTS.get(cfg.name)


# ---- tensors ----------------------------------------------------------------------------------------------------
_LAYER = {"input_layernorm.weight": "attn_norm", "post_attention_layernorm.weight": "mlp_norm",
          "self_attn.q_proj.weight": "wq", "self_attn.k_proj.weight": "wk", "self_attn.v_proj.weight": "wv",
          "self_attn.o_proj.weight": "wo", "mlp.gate_proj.weight": "w_gate", "mlp.up_proj.weight": "w_up",
          "mlp.down_proj.weight": "w_down"}
_TOP = {"model.embed_tokens.weight": "embed", "model.norm.weight": "final_norm", "lm_head.weight": "lm_head"}


def canonical_name(hf_name: str) -> Optional[str]:
    """`transformers` tensor name -> this package's canonical name (weights.generate_weights), None = not a weight
    of the forward pass (e.g. rotary inv_freq buffers of old checkpoints)."""
    if hf_name in _TOP:
        return _TOP[hf_name]
    if hf_name.startswith("model.layers."):
        idx, _, rest = hf_name[len("model.layers."):].partition(".")
        if idx.isdigit() and rest in _LAYER:
            return f"l{int(idx)}.{_LAYER[rest]}"
    return None


_FILES_CACHE: Dict[str, Dict[str, str]] = {}


def _files(path: str) -> Dict[str, str]:
    """hf tensor name -> safetensors file holding it (remembered per directory: load_fused asks once per layer)."""
    if str(path) not in _FILES_CACHE:
        _FILES_CACHE[str(path)] = _scan_files(path)
    return _FILES_CACHE[str(path)]


def _scan_files(path: str) -> Dict[str, str]:
    from safetensors import safe_open
    p = Path(path)
    index = p / "model.safetensors.index.json"
    if index.exists():
        wm = json.loads(index.read_text())["weight_map"]
        return {k: str(p / v) for k, v in wm.items()}
    files = sorted(p.glob("*.safetensors"))
    if not files:
        raise CheckpointError(f"{path}: no *.safetensors file (only the safetensors format is read)")
    out: Dict[str, str] = {}
    for f in files:
        with safe_open(str(f), framework="pt") as sf:
            for k in sf.keys():
                out[k] = str(f)
    return out


def _expected(cfg: LlamaConfig) -> Dict[str, Tuple[int, ...]]:
    H, F = cfg.hidden, cfg.ffn
    shapes = {"embed": (cfg.vocab, H), "final_norm": (H,), "lm_head": (cfg.vocab, H)}
    for i in range(cfg.n_layers):
        p = f"l{i}."
        shapes.update({p + "attn_norm": (H,), p + "mlp_norm": (H,), p + "wq": (cfg.q_dim, H), p + "wk": (cfg.kv_dim, H),
                       p + "wv": (cfg.kv_dim, H), p + "wo": (H, cfg.q_dim), p + "w_gate": (F, H), p + "w_up": (F, H),
                       p + "w_down": (H, F)})
    return shapes


def iter_canonical(path: str, cfg: LlamaConfig, names: Optional[Iterable[str]] = None) -> Iterator[Tuple[str, "object"]]:
    """(canonical name, bf16 CPU tensor) for every tensor of the forward pass (or just `names`), one file open at a
    time; shapes are checked against `cfg`. A checkpoint with tied embeddings yields `lm_head` = `embed`."""
    import torch
    from safetensors import safe_open
    files = _files(path)
    want = _expected(cfg)
    by_file: Dict[str, list] = {}
    have = set()
    for hf_name, f in files.items():
        cn = canonical_name(hf_name)
        if cn is None or cn not in want:
            continue
        have.add(cn)
        if names is None or cn in names:
            by_file.setdefault(f, []).append((hf_name, cn))
    tied = "lm_head" not in have and "embed" in have
    missing = sorted(set(want) - have - ({"lm_head"} if tied else set()))
    if missing:
        raise CheckpointError(f"{path}: {len(missing)} tensors missing, e.g. {missing[:4]}")
    for f, items in by_file.items():
        with safe_open(f, framework="pt") as sf:
            for hf_name, cn in items:
                t = sf.get_tensor(hf_name)
                if tuple(t.shape) != want[cn]:
                    raise CheckpointError(f"{path}: {hf_name} has shape {tuple(t.shape)}, the config says {want[cn]}")
                t = t.to(torch.bfloat16)
                yield cn, t
                if cn == "embed" and tied and (names is None or "lm_head" in names):
                    yield "lm_head", t


def load_canonical(path: str, cfg: LlamaConfig) -> Dict[str, "object"]:
    """All canonical tensors on the CPU (small models, tests, the oracle)."""
    return dict(iter_canonical(path, cfg))


def load_fused(path: str, cfg: LlamaConfig, device) -> Dict[str, "object"]:
    """The engine's layout (weights.fuse_for_engine) built layer by layer, so that the host never holds more than
    one layer of a large model."""
    from .weights import fuse_for_engine, fuse_layer
    _files(path)                  # fail early on a broken directory
    top = dict(iter_canonical(path, cfg, {"embed", "final_norm", "lm_head"}))
    out = fuse_for_engine(LlamaConfig(**{**cfg.to_dict(), "n_layers": 0}), top, device)
    del top
    layer_keys = ("attn_norm", "mlp_norm", "wq", "wk", "wv", "wo", "w_gate", "w_up", "w_down")
    for i in range(cfg.n_layers):
        src = dict(iter_canonical(path, cfg, {f"l{i}.{k}" for k in layer_keys}))
        out.update(fuse_layer(cfg, src, i, device))
    return out


def load_fused_shard(path: str, cfg: LlamaConfig, rank: int, tp: int, device, shard_lm_head: bool = False
                     ) -> Dict[str, "object"]:
    """Rank `rank`'s tensor-parallel shard (tp.shard_weights) in the engine's layout, cut while loading: one full
    layer passes through host memory at a time, so a 70B checkpoint needs ~2 GB of host memory per rank, not 141."""
    from .tp import lm_head_slice, local_config, shard_weights
    from .weights import fuse_for_engine, fuse_layer
    if tp <= 1:
        return load_fused(path, cfg, device)
    loc = local_config(cfg, tp)
    top = dict(iter_canonical(path, cfg, {"embed", "final_norm", "lm_head"}))
    sl = lm_head_slice(cfg, rank, tp) if shard_lm_head else None
    if sl is not None:
        top["lm_head"] = top["lm_head"][sl[0]:sl[0] + sl[1]]
    out = fuse_for_engine(LlamaConfig(**{**loc.to_dict(), "n_layers": 0}), top, device)
    del top
    layer_keys = ("attn_norm", "mlp_norm", "wq", "wk", "wv", "wo", "w_gate", "w_up", "w_down")
    one = LlamaConfig(**{**cfg.to_dict(), "n_layers": 1})
    for i in range(cfg.n_layers):
        src = dict(iter_canonical(path, cfg, {f"l{i}.{k}" for k in layer_keys}))
        full = {f"l0.{k}": src[f"l{i}.{k}"] for k in layer_keys}
        full.update({"embed": None, "final_norm": None, "lm_head": None})      # not touched per layer
        cut = shard_weights(one, full, rank, tp)
        out.update(fuse_layer(loc, cut, i, device, src_index=0))
    return out


def load_tokenizer(path: Optional[str], cfg: LlamaConfig):
    """The checkpoint's tokenizer (tokenizer.json with Llama-3's control tokens) or, without one, the byte tokenizer."""
    from .tokenizer import ByteTokenizer, HFTokenizer
    if path and (Path(path) / "tokenizer.json").exists():
        return HFTokenizer(str(Path(path) / "tokenizer.json"), cfg.vocab)
    return ByteTokenizer(cfg.vocab)


__all__ = ["CHECKPOINTS", "CheckpointError", "REGISTRY", "canonical_name", "checkpoint_dir", "config_from_hf",
           "iter_canonical", "load_canonical", "load_fused", "load_fused_shard", "load_tokenizer", "register_checkpoint", "register_from_env"]
