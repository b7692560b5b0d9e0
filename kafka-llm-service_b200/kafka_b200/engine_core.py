"""EngineCore — one GPU's model replica behind the C ABI.

Owns (through PyTorch, which is only the allocator here) the fused bf16 weights, the paged KV pools
and the activation workspace, binds them into a ``kb_engine`` and exposes the two hot-path calls
(``prefill`` / ``decode``) on device tensors. No arithmetic happens in Python.

Takes over what `/root/reference/src/llm/portkey.py:379` delegates to a remote service.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional

import torch

from . import cabi
from .cabi import KB_BLOCK_TOKENS, KafkaB200Error
from .model_config import LlamaConfig
from .kv_layout import KV_DTYPE, pool_shape


class EngineCore:
    def __init__(self, cfg: LlamaConfig, fused_weights: Dict[str, torch.Tensor], device: torch.device,
                 n_blocks: int, max_tokens: int, max_seqs: int):
        if not torch.cuda.is_available():
            raise KafkaB200Error("EngineCore needs a CUDA device (sm_100a); there is no CPU fallback")
        self.lib = cabi.load()
        self.cfg = cfg
        self.device = torch.device(device)
        self.n_blocks = int(n_blocks)
        self.max_tokens = int(max_tokens)
        self.max_seqs = int(max_seqs)
        torch.cuda.set_device(self.device)

        self.weights = fused_weights  # keep the tensors alive: the library only borrows pointers
        from .weights import GEMM_WEIGHT_DTYPE, GEMM_WEIGHT_NAMES
        for name, t in fused_weights.items():
            want = GEMM_WEIGHT_DTYPE if name.split(".")[-1] in GEMM_WEIGHT_NAMES else torch.bfloat16
            if t.device != self.device or t.dtype != want or not t.is_contiguous():
                raise KafkaB200Error(f"weight {name} must be a contiguous {want} tensor on {self.device}")

        self._cfg_c = cabi.ModelCfg(cfg.hidden, cfg.n_layers, cfg.n_heads, cfg.n_kv_heads, cfg.head_dim, cfg.ffn,
                                    cfg.vocab, cfg.max_pos, cfg.rope_theta, cfg.rms_eps)
        handle = C.c_void_p()
        cabi.check(self.lib.kb_engine_create(self.device.index or 0, C.byref(self._cfg_c), C.byref(handle)),
                   "kb_engine_create")
        self._h = handle

        layers = (cabi.LayerWeights * cfg.n_layers)()
        for i in range(cfg.n_layers):
            p = f"l{i}."
            layers[i] = cabi.LayerWeights(*(fused_weights[p + k].data_ptr() for k in
                                            ("attn_norm", "wqkv", "wo", "mlp_norm", "w_gate_up", "w_down")))
        self._layers_c = layers
        wp = cabi.WeightPtrs(fused_weights["embed"].data_ptr(), fused_weights["final_norm"].data_ptr(),
                             fused_weights["lm_head"].data_ptr(), layers)
        cabi.check(self.lib.kb_bind_weights(self._h, C.byref(wp)), "kb_bind_weights")

        # KV pools: zero-initialised so masked tail positions are finite (see kb_attn_decode's contract)
        shape = pool_shape(cfg.n_layers, self.n_blocks, cfg.n_kv_heads, cfg.head_dim)
        self.k_pool = torch.zeros(shape, dtype=KV_DTYPE, device=self.device)
        self.v_pool = torch.zeros(shape, dtype=KV_DTYPE, device=self.device)
        cabi.check(self.lib.kb_bind_kv_pool(self._h, self.k_pool.data_ptr(), self.v_pool.data_ptr(), self.n_blocks),
                   "kb_bind_kv_pool")

        ws_bytes = int(self.lib.kb_workspace_bytes(C.byref(self._cfg_c), self.max_tokens, self.max_seqs))
        if ws_bytes == 0:
            raise KafkaB200Error(f"kb_workspace_bytes rejected the configuration: {cabi.last_error()}")
        self.workspace = torch.empty(ws_bytes, dtype=torch.uint8, device=self.device)
        cabi.check(self.lib.kb_bind_workspace(self._h, self.workspace.data_ptr(), ws_bytes, self.max_tokens,
                                              self.max_seqs), "kb_bind_workspace")
        # scaled RoPE (Llama-3.1's rule): the frequencies are reshaped here and handed over as a table
        self.rope_inv_freq = None
        if cfg.rope_llama3 is not None:
            from .model_config import inv_freq
            self.rope_inv_freq = torch.tensor(inv_freq(cfg), dtype=torch.float32, device=self.device)
            cabi.check(self.lib.kb_set_rope_inv_freq(self._h, self.rope_inv_freq.data_ptr()), "kb_set_rope_inv_freq")

    # ---- sizing helpers ----------------------------------------------------------------------
    @staticmethod
    def kv_block_bytes(cfg: LlamaConfig) -> int:
        """Bytes one KV block occupies over all layers, K and V."""
        return 2 * cfg.n_layers * cfg.n_kv_heads * KB_BLOCK_TOKENS * cfg.head_dim * 2

    def close(self) -> None:
        if getattr(self, "_h", None):
            torch.cuda.synchronize(self.device)     # nothing enqueued may outlive the buffers it borrows
            self.lib.kb_engine_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- per-kernel timing (CUDA events between launches; bench.py) ------------------------------
    def profile_begin(self) -> None:
        cabi.check(self.lib.kb_profile_begin(self._h), "kb_profile_begin")

    def profile_end(self):
        """-> {category: (milliseconds, launches)} summed over the calls since profile_begin."""
        n = len(cabi.PROF_CATS)
        ms = (C.c_float * n)()
        cnt = (C.c_int32 * n)()
        cabi.check(self.lib.kb_profile_end(self._h, ms, cnt), "kb_profile_end")
        return {name: (float(ms[i]), int(cnt[i])) for i, name in enumerate(cabi.PROF_CATS)}

    lm_head_slice = None     # (first id, ids) when tp.build_tp_core bound a vocabulary slice of the lm_head

    def _check_logits(self, logits_out, n_seqs: int) -> None:
        if logits_out is None:
            return
        width = self.lm_head_slice[1] if self.lm_head_slice else self.cfg.vocab
        if logits_out.dim() != 2 or logits_out.shape[0] < n_seqs or logits_out.shape[1] != width:
            raise ValueError(f"logits_out must be [>= {n_seqs}, {width}] fp32 (this rank's vocabulary slice when the "
                             f"lm_head is sharded), got {tuple(logits_out.shape)}")

    # ---- hot path ----------------------------------------------------------------------------
    def prefill(self, tokens: torch.Tensor, cu_seqlens: torch.Tensor, ctx_lens: torch.Tensor,
                block_tables: torch.Tensor, max_new: int, sampling: Optional[cabi.Sampling] = None,
                logits_out: Optional[torch.Tensor] = None, out_tokens: Optional[torch.Tensor] = None) -> None:
        n_seqs, max_blocks = block_tables.shape
        self._check_logits(logits_out, n_seqs)
        cabi.check(self.lib.kb_prefill(self._h, tokens.data_ptr(), cu_seqlens.data_ptr(), ctx_lens.data_ptr(),
                                       block_tables.data_ptr(), n_seqs, tokens.numel(), max_blocks, int(max_new),
                                       C.byref(sampling) if sampling is not None else None,
                                       None if logits_out is None else logits_out.data_ptr(),
                                       None if out_tokens is None else out_tokens.data_ptr(),
                                       torch.cuda.current_stream().cuda_stream), "kb_prefill")

    def decode(self, last_tokens: torch.Tensor, ctx_lens: torch.Tensor, block_tables: torch.Tensor,
               sampling: cabi.Sampling, out_tokens: torch.Tensor, max_ctx: int = 0,
               logits_out: Optional[torch.Tensor] = None, advance_ctx: bool = False) -> None:
        batch, max_blocks = block_tables.shape
        self._check_logits(logits_out, batch)
        cabi.check(self.lib.kb_decode(self._h, last_tokens.data_ptr(), ctx_lens.data_ptr(),
                                      block_tables.data_ptr(), batch, max_blocks, int(max_ctx), C.byref(sampling),
                                      out_tokens.data_ptr(),
                                      None if logits_out is None else logits_out.data_ptr(),
                                      1 if advance_ctx else 0, torch.cuda.current_stream().cuda_stream),
                   "kb_decode")
