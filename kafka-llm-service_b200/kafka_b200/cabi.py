"""ctypes binding of libkafka_b200.so (the C ABI declared in include/kafka_b200.h).

This is the only place Python touches the CUDA path. Tensors are passed as raw device pointers
(``tensor.data_ptr()``) and the current torch CUDA stream as a ``cudaStream_t``; PyTorch is used for
allocation and streams only. There is deliberately no fallback: if the shared library is missing or
a call fails, ``KafkaB200Error`` is raised.

Reference counterpart: none — the reference's provider boundary is a Python ABC
(/root/reference/src/llm/base.py:67) whose implementation is an HTTPS call
(/root/reference/src/llm/portkey.py:379). INTEGRATION.md shows the stub a maintainer would add.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path
from typing import Optional

KB_BLOCK_TOKENS = 64
KB_MAX_SPLITS = 16
KB_ABI_VERSION = 2
PROF_CATS = ("embed", "rmsnorm", "gemm_qkv", "rope_append", "attention", "gemm_o", "gemm_gate_up", "gemm_down",
             "lm_head", "sample", "allreduce")
ALLREDUCE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p)

EPI_BF16, EPI_F32, EPI_RESID_F32, EPI_SWIGLU_BF16, EPI_SWIGLU_F16, EPI_F16 = 0, 1, 2, 3, 4, 5
GEMM_F16 = 0x100  # OR into the epilogue: both operands are fp16 instead of bf16 (formats must match)

_LIB_PATH = Path(__file__).resolve().parent / "_lib" / "libkafka_b200.so"


class KafkaB200Error(RuntimeError):
    """Raised when the CUDA library is missing or a C-ABI call returns non-zero."""


class ModelCfg(C.Structure):
    _fields_ = [
        ("hidden", C.c_int32), ("n_layers", C.c_int32), ("n_heads", C.c_int32), ("n_kv_heads", C.c_int32),
        ("head_dim", C.c_int32), ("ffn", C.c_int32), ("vocab", C.c_int32), ("max_pos", C.c_int32),
        ("rope_theta", C.c_float), ("rms_eps", C.c_float),
    ]


class LayerWeights(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("attn_norm", "wqkv", "wo", "mlp_norm", "w_gate_up", "w_down")]


class WeightPtrs(C.Structure):
    _fields_ = [("embed", C.c_void_p), ("final_norm", C.c_void_p), ("lm_head", C.c_void_p),
                ("layers", C.POINTER(LayerWeights))]


class Sampling(C.Structure):
    _fields_ = [("temperature", C.c_float), ("seed", C.c_uint64), ("step", C.c_uint64),
                ("banned_token", C.c_int32), ("allowed_lo", C.c_int32), ("allowed_hi", C.c_int32),
                ("extra_a", C.c_int32), ("extra_b", C.c_int32),
                ("seq_temperature", C.c_void_p), ("seq_seed", C.c_void_p),
                ("top_p", C.c_float), ("seq_top_p", C.c_void_p)]


TP_MAX_RANKS, TP_MAX_BLOCKS = 8, 128
TP_FLAG_BYTES = 2048


class TpPeers(C.Structure):
    _fields_ = [("peer_base", C.c_void_p * TP_MAX_RANKS), ("mc_base", C.c_void_p), ("bytes", C.c_size_t),
                ("flags_offset", C.c_size_t), ("epoch_counter", C.POINTER(C.c_uint32)),
                ("timeout_cycles", C.c_ulonglong)]


# name -> (restype, argtypes); kept in one table so tests can check every symbol the header declares.
_i32, _vp, _sz, _f = C.c_int32, C.c_void_p, C.c_size_t, C.c_float
SIGNATURES = {
    "kb_engine_create": (C.c_int, [C.c_int, C.POINTER(ModelCfg), C.POINTER(_vp)]),
    "kb_engine_destroy": (None, [_vp]),
    "kb_bind_weights": (C.c_int, [_vp, C.POINTER(WeightPtrs)]),
    "kb_bind_kv_pool": (C.c_int, [_vp, _vp, _vp, _i32]),
    "kb_workspace_bytes": (_sz, [C.POINTER(ModelCfg), _i32, _i32]),
    "kb_bind_workspace": (C.c_int, [_vp, _vp, _sz, _i32, _i32]),
    "kb_prefill": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, C.POINTER(Sampling), _vp, _vp, _vp]),
    "kb_decode": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _i32, _i32, C.POINTER(Sampling), _vp, _vp, _i32, _vp]),
    "kb_rmsnorm": (C.c_int, [_vp, _vp, _vp, _i32, _i32, _f, _i32, _vp]),
    "kb_embed": (C.c_int, [_vp, _vp, _vp, _i32, _i32, _i32, _vp]),
    "kb_gemm": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp]),
    "kb_gemm_ex": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _vp]),
    "kb_rope_append": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _f, _vp]),
    "kb_rope_append_freqs": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _vp, _vp]),
    "kb_set_rope_inv_freq": (C.c_int, [_vp, _vp]),
    "kb_attn_decode_scratch_bytes": (_sz, [_i32, _i32, _i32]),
    "kb_attn_decode": (C.c_int, [_vp, _i32, _vp, _vp, _vp, _i32, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32,
                                 _vp, _sz, _i32, _vp]),
    "kb_static_scratch_capacity": (_sz, []),
    "kb_attn_decode_static_scratch": (C.c_int, [_vp, _i32, _vp, _vp, _vp, _i32, _vp, _vp, _i32, _i32, _i32, _i32,
                                                _i32, _i32, _i32, _vp]),
    "kb_attn_prefill": (C.c_int, [_vp, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32,
                                  _i32, _i32, _vp]),
    "kb_sample": (C.c_int, [_vp, _i32, _i32, C.POINTER(Sampling), _vp, _vp]),
    "kb_sample_slice": (C.c_int, [_vp, _i32, _i32, _i32, _i32, C.POINTER(Sampling), _vp, _vp]),
    "kb_select_tokens": (C.c_int, [_vp, _vp, _vp, _i32, _i32, _vp]),
    "kb_hbm_read_probe": (C.c_int, [_vp, _sz, _vp, _sz, _vp]),
    "kb_set_tensor_parallel": (C.c_int, [_vp, _i32, _i32, ALLREDUCE_FN, _vp]),
    "kb_set_tensor_parallel_peers": (C.c_int, [_vp, _i32, _i32, C.POINTER(TpPeers), _i32, _i32]),
    "kb_allreduce_peers": (C.c_int, [C.POINTER(TpPeers), _i32, _i32, _sz, _sz, _i32, C.c_uint32, _i32, _vp]),
    "kb_set_lm_head_shard": (C.c_int, [_vp, _i32, _i32, _sz]),
    "kb_tp_argmax_pairs": (C.c_int, [C.POINTER(TpPeers), _i32, _i32, _sz, _i32, C.c_uint32, _vp, _vp]),
    "kb_tp_reduce_norm_gather": (C.c_int, [C.POINTER(TpPeers), _i32, _i32, _sz, _sz, _vp, _vp, _i32, _i32, _f,
                                           C.c_uint32, _i32, _vp]),
    "kb_profile_begin": (C.c_int, [_vp]),
    "kb_profile_end": (C.c_int, [_vp, C.POINTER(C.c_float), C.POINTER(C.c_int32)]),
    "kb_last_error": (C.c_char_p, []),
    "kb_abi_version": (C.c_int, []),
    "kb_launch_count": (C.c_ulonglong, []),
}

_lib: Optional[C.CDLL] = None


def lib_path() -> Path:
    return Path(os.environ.get("KAFKA_B200_LIB", str(_LIB_PATH)))


def load() -> C.CDLL:
    """Load the shared library once; raise loudly if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not path.exists():
        raise KafkaB200Error(
            f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). There is no CPU fallback for the completion path.")
    try:
        lib = C.CDLL(str(path))
    except OSError as e:  # missing libcudart etc.
        raise KafkaB200Error(f"cannot load {path}: {e}") from e
    for name, (res, args) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:
            raise KafkaB200Error(f"{path} does not export {name}") from e
        fn.restype = res
        fn.argtypes = args
    if lib.kb_abi_version() != KB_ABI_VERSION:
        raise KafkaB200Error(f"ABI mismatch: library {lib.kb_abi_version()} vs binding {KB_ABI_VERSION}")
    _lib = lib
    return lib


def last_error() -> str:
    return (load().kb_last_error() or b"").decode("utf-8", "replace")


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise KafkaB200Error(f"{what} failed (rc={rc}): {last_error()}")


def launch_count() -> int:
    return int(load().kb_launch_count())


def _ptr(t) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _stream() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream


# ------------------------------------------------------------------------------------------------
# thin tensor-level wrappers of the per-kernel entry points (used by tests, smoke and the engine)
# ------------------------------------------------------------------------------------------------
def rmsnorm(x, w, eps: float, out_f16: bool = False):
    import torch
    n, hidden = x.shape
    y = torch.empty((n, hidden), dtype=torch.float16 if out_f16 else torch.bfloat16, device=x.device)
    check(load().kb_rmsnorm(_ptr(x), _ptr(w), _ptr(y), n, hidden, eps, int(out_f16), _stream()), "kb_rmsnorm")
    return y


def embed(tokens, table):
    import torch
    n = tokens.numel()
    vocab, hidden = table.shape
    x = torch.empty((n, hidden), dtype=torch.float32, device=table.device)
    check(load().kb_embed(_ptr(tokens), _ptr(table), _ptr(x), n, hidden, vocab, _stream()), "kb_embed")
    return x


def gemm(a, w, epilogue: int = EPI_BF16, resid=None, out=None, tile_n: int = 0):
    """out = a[M,K] @ w[N,K]^T with the fused epilogue (see include/kafka_b200.h)."""
    import torch
    M, K = a.shape
    N, K2 = w.shape
    assert K == K2
    if a.dtype != w.dtype:
        raise KafkaB200Error(f"kb_gemm: operand formats must match (A {a.dtype}, W {w.dtype})")
    if a.dtype == torch.float16:
        epilogue |= GEMM_F16
    if out is None:
        epi = epilogue & 0xFF
        if epi == EPI_BF16:
            out = torch.empty((M, N), dtype=torch.bfloat16, device=a.device)
        elif epi == EPI_SWIGLU_BF16:
            out = torch.empty((M, N // 2), dtype=torch.bfloat16, device=a.device)
        elif epi == EPI_SWIGLU_F16:
            out = torch.empty((M, N // 2), dtype=torch.float16, device=a.device)
        elif epi == EPI_F16:
            out = torch.empty((M, N), dtype=torch.float16, device=a.device)
        else:
            out = torch.empty((M, N), dtype=torch.float32, device=a.device)
    check(load().kb_gemm_ex(_ptr(a), _ptr(w), _ptr(out), _ptr(resid), M, N, K, epilogue, tile_n, _stream()),
          "kb_gemm")
    return out


def _need_f16(what: str, *tensors) -> None:
    """q / k / v and the KV pools are fp16 (include/kafka_b200.h); a bf16 tensor would be read as garbage."""
    import torch
    for t in tensors:
        if t.dtype != torch.float16:
            raise KafkaB200Error(f"{what}: q / k / v and the KV pools must be float16, got {t.dtype}")


def rope_append(qkv, cu_seqlens, ctx_lens, block_tables, k_pool_layer, v_pool_layer, n_heads, n_kv_heads,
                head_dim, theta):
    _need_f16("kb_rope_append", qkv, k_pool_layer, v_pool_layer)
    n_tokens = qkv.shape[0]
    n_seqs, max_blocks = block_tables.shape
    check(load().kb_rope_append(_ptr(qkv), _ptr(cu_seqlens), _ptr(ctx_lens), _ptr(block_tables),
                                _ptr(k_pool_layer), _ptr(v_pool_layer), n_seqs, n_tokens, max_blocks, n_heads,
                                n_kv_heads, head_dim, theta, _stream()), "kb_rope_append")


def rope_append_freqs(qkv, cu_seqlens, ctx_lens, block_tables, k_pool_layer, v_pool_layer, n_heads, n_kv_heads,
                      head_dim, inv_freq):
    """rope_append with the head_dim / 2 rotary frequencies as a float32 device tensor (scaled RoPE)."""
    _need_f16("kb_rope_append_freqs", qkv, k_pool_layer, v_pool_layer)
    if inv_freq.dtype != __import__("torch").float32 or inv_freq.numel() != head_dim // 2 or not inv_freq.is_cuda:
        raise ValueError(f"inv_freq must be {head_dim // 2} float32 values on the device")
    n_tokens = qkv.shape[0]
    n_seqs, max_blocks = block_tables.shape
    check(load().kb_rope_append_freqs(_ptr(qkv), _ptr(cu_seqlens), _ptr(ctx_lens), _ptr(block_tables),
                                      _ptr(k_pool_layer), _ptr(v_pool_layer), n_seqs, n_tokens, max_blocks, n_heads,
                                      n_kv_heads, head_dim, _ptr(inv_freq), _stream()), "kb_rope_append_freqs")


def attn_decode(q, k_pool_layer, v_pool_layer, kv_lens, block_tables, n_heads, n_kv_heads, head_dim,
                kv_len_bias: int = 0, max_kv_len: int = 0, scratch=None, out_f16: bool = False):
    import torch
    _need_f16("kb_attn_decode", q, k_pool_layer, v_pool_layer)
    batch, max_blocks = block_tables.shape
    out = torch.empty((batch, n_heads * head_dim), dtype=torch.float16 if out_f16 else torch.bfloat16,
                      device=q.device)
    sb = 0 if scratch is None else scratch.numel() * scratch.element_size()
    check(load().kb_attn_decode(_ptr(q), q.stride(0), _ptr(k_pool_layer), _ptr(v_pool_layer), _ptr(kv_lens),
                                kv_len_bias, _ptr(block_tables), _ptr(out), batch, max_blocks, n_heads, n_kv_heads,
                                head_dim, max_kv_len, _ptr(scratch), sb, int(out_f16), _stream()), "kb_attn_decode")
    return out


def attn_decode_static_scratch(q, k_pool_layer, v_pool_layer, kv_lens, block_tables, n_heads, n_kv_heads, head_dim,
                               kv_len_bias: int = 0, max_kv_len: int = 0, out=None):
    """kb_attn_decode for hosts that cannot allocate: the split-KV partials go to the library's
    built-in 2 MiB device array instead of a caller buffer. Pass ``out`` to avoid any allocation on
    this side too (the default allocates the result with torch for convenience in tests)."""
    import torch
    _need_f16("kb_attn_decode_static_scratch", q, k_pool_layer, v_pool_layer)
    batch, max_blocks = block_tables.shape
    if out is None:
        out = torch.empty((batch, n_heads * head_dim), dtype=torch.bfloat16, device=q.device)
    check(load().kb_attn_decode_static_scratch(_ptr(q), q.stride(0), _ptr(k_pool_layer), _ptr(v_pool_layer),
                                               _ptr(kv_lens), kv_len_bias, _ptr(block_tables), _ptr(out), batch,
                                               max_blocks, n_heads, n_kv_heads, head_dim, max_kv_len,
                                               1 if out.dtype == torch.float16 else 0, _stream()),
          "kb_attn_decode_static_scratch")
    return out


def attn_decode_scratch_bytes(batch: int, n_heads: int, head_dim: int) -> int:
    return int(load().kb_attn_decode_scratch_bytes(batch, n_heads, head_dim))


def attn_prefill(q, k_pool_layer, v_pool_layer, cu_seqlens, ctx_lens, block_tables, max_new, n_heads, n_kv_heads,
                 head_dim, out_f16: bool = False):
    import torch
    _need_f16("kb_attn_prefill", q, k_pool_layer, v_pool_layer)
    n_tokens = q.shape[0]
    n_seqs, max_blocks = block_tables.shape
    out = torch.empty((n_tokens, n_heads * head_dim), dtype=torch.float16 if out_f16 else torch.bfloat16,
                      device=q.device)
    check(load().kb_attn_prefill(_ptr(q), q.stride(0), _ptr(k_pool_layer), _ptr(v_pool_layer), _ptr(cu_seqlens),
                                 _ptr(ctx_lens), _ptr(block_tables), _ptr(out), n_seqs, n_tokens, max_new,
                                 max_blocks, n_heads, n_kv_heads, head_dim, int(out_f16), _stream()),
          "kb_attn_prefill")
    return out


def make_sampling(temperature: float = 0.0, seed: int = 0, step: int = 0, banned_token: int = -1,
                  allowed_lo: int = 0, allowed_hi: int = 0, extra_a: int = -1, extra_b: int = -1,
                  seq_temperature=None, seq_seed=None, top_p: float = 0.0, seq_top_p=None) -> Sampling:
    """seq_temperature / seq_top_p (float32) and seq_seed (int64 bit patterns) are optional per-sequence device
    tensors. top_p in (0, 1): nucleus sampling (0 or >= 1: off)."""
    return Sampling(float(temperature), int(seed), int(step), int(banned_token), int(allowed_lo), int(allowed_hi),
                    int(extra_a), int(extra_b), _ptr(seq_temperature), _ptr(seq_seed), float(top_p), _ptr(seq_top_p))


def sample(logits, temperature: float = 0.0, seed: int = 0, step: int = 0, banned_token: int = -1,
           allowed_lo: int = 0, allowed_hi: int = 0, extra_a: int = -1, extra_b: int = -1, top_p: float = 0.0,
           seq_top_p=None):
    import torch
    n, vocab = logits.shape
    out = torch.empty((n,), dtype=torch.int32, device=logits.device)
    sp = make_sampling(temperature, seed, step, banned_token, allowed_lo, allowed_hi, extra_a, extra_b, top_p=top_p,
                       seq_top_p=seq_top_p)
    check(load().kb_sample(_ptr(logits), n, vocab, C.byref(sp), _ptr(out), _stream()), "kb_sample")
    return out


def sample_slice(logits_slice, vocab: int, vocab_offset: int, temperature: float = 0.0, seed: int = 0, step: int = 0,
                 banned_token: int = -1, allowed_lo: int = 0, allowed_hi: int = 0, extra_a: int = -1, extra_b: int = -1,
                 out_pairs=None):
    """One rank's half of vocabulary-sharded sampling: `logits_slice` [n, vocab_local] holds ids
    [vocab_offset, +vocab_local) of a `vocab`-wide vocabulary. Returns an int32 [n, 2] tensor of
    (fp32 bits of the winning value, global id); id 0x7fffffff when the slice holds no candidate."""
    import torch
    n, local = logits_slice.shape
    out = out_pairs if out_pairs is not None else torch.empty((n, 2), dtype=torch.int32, device=logits_slice.device)
    sp = make_sampling(temperature, seed, step, banned_token, allowed_lo, allowed_hi, extra_a, extra_b)
    check(load().kb_sample_slice(_ptr(logits_slice), n, vocab, vocab_offset, local, C.byref(sp), _ptr(out), _stream()),
          "kb_sample_slice")
    return out


def select_tokens(dst, prev_out, src_index, stream=None):
    """dst[i] = prev_out[src_index[i]] where src_index[i] >= 0 (device int32 tensors; in place)."""
    check(load().kb_select_tokens(_ptr(dst), _ptr(prev_out), _ptr(src_index), dst.numel(), prev_out.numel(),
                                  _stream() if stream is None else stream), "kb_select_tokens")
    return dst


def hbm_read_probe(buf, sink=None):
    """Enqueue one read-only pass over `buf` (a CUDA tensor); returns the bytes read. See kb_hbm_read_probe."""
    import torch
    if sink is None:
        sink = torch.zeros(4 * 160, dtype=torch.int32, device=buf.device)
    n = buf.numel() * buf.element_size()
    check(load().kb_hbm_read_probe(_ptr(buf), n, _ptr(sink), sink.numel() * 4, _stream()), "kb_hbm_read_probe")
    return n
