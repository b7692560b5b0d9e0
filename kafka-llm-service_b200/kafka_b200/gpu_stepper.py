"""GpuStepper — binds the scheduler's batches to the C ABI (kb_prefill / kb_decode).

Per call it stages the small integer arguments (token ids, cu_seqlens, ctx_lens, block tables,
per-sequence sampling parameters) in pinned host memory, copies them to persistent device buffers on
its own stream, launches the call, and brings the sampled ids back through pinned memory. PyTorch
is the allocator and stream provider only. Block tables and sampling arrays are re-uploaded only when
the batch membership changes (decode steps of a stable batch upload 2 x B ints).

Reference counterpart: none (the reference awaits a network call, /root/reference/src/llm/portkey.py:379).
"""
from __future__ import annotations

from typing import List, Optional, Sequence

from .engine import Stepper
from .scheduler import BLOCK_TOKENS, PrefillItem, Request


class _Staging:
    """One set of pinned + device buffers (the prefill and decode paths each own one so that their
    membership caches do not evict each other)."""

    def __init__(self, torch, dev, max_tokens: int, max_seqs: int, max_blocks: int):
        i32 = torch.int32
        self.torch = torch
        T, S = max_tokens, max_seqs

        def pair(n, dtype):
            return torch.zeros(n, dtype=dtype, pin_memory=True), torch.zeros(n, dtype=dtype, device=dev)

        self.h_tok, self.d_tok = pair(T, i32)
        self.h_cu, self.d_cu = pair(S + 1, i32)
        self.h_ctx, self.d_ctx = pair(S, i32)
        self.h_bt, self.d_bt = pair(S * max_blocks, i32)
        self.h_temp, self.d_temp = pair(S, torch.float32)
        self.h_seed, self.d_seed = pair(S, torch.int64)
        self.h_out, self.d_out = pair(S, i32)
        self.bt_key = None
        self.sp_key = None

    @staticmethod
    def _up(h, d, values, dtype, torch):
        n = len(values)
        h[:n] = torch.tensor(values, dtype=dtype)
        d[:n].copy_(h[:n], non_blocking=True)
        return d[:n]

    def stage(self, tokens: List[int], cu: Optional[List[int]], ctx: List[int], reqs: Sequence[Request]):
        torch = self.torch
        n = len(reqs)
        d_tok = self._up(self.h_tok, self.d_tok, tokens, torch.int32, torch)
        d_cu = self._up(self.h_cu, self.d_cu, cu, torch.int32, torch) if cu is not None else None
        d_ctx = self._up(self.h_ctx, self.d_ctx, ctx, torch.int32, torch)
        key = tuple((r.req_id, len(r.entry.blocks)) for r in reqs)
        mb = max(1, max(len(r.entry.blocks) for r in reqs))
        if key != self.bt_key:
            flat: List[int] = []
            for r in reqs:
                b = r.entry.blocks
                flat += b
                if len(b) < mb:
                    flat += [0] * (mb - len(b))
            self._up(self.h_bt, self.d_bt, flat, torch.int32, torch)
            self.bt_key = key
        if key != self.sp_key:
            self._up(self.h_temp, self.d_temp, [r.params.temperature for r in reqs], torch.float32, torch)
            self._up(self.h_seed, self.d_seed,
                     [(r.params.seed * 0x9E3779B97F4A7C15 + r.req_id) & 0x7FFFFFFFFFFFFFFF for r in reqs],
                     torch.int64, torch)
            self.sp_key = key
        return d_tok, d_cu, d_ctx, self.d_bt[:n * mb].view(n, mb)


class GpuStepper(Stepper):
    def __init__(self, core, tokenizer, generable_only: bool = True):
        import torch
        from . import cabi
        self.torch, self.cabi = torch, cabi
        self.core = core
        cfg = core.cfg
        self.max_seqs, self.max_tokens, self.n_blocks = core.max_seqs, core.max_tokens, core.n_blocks
        self.max_pos, self.eos_id = cfg.max_pos, tokenizer.eot_id
        self.max_blocks = (cfg.max_pos + BLOCK_TOKENS - 1) // BLOCK_TOKENS
        self.stream = torch.cuda.Stream(device=core.device)
        self.st_prefill = _Staging(torch, core.device, core.max_tokens, core.max_seqs, self.max_blocks)
        self.st_decode = _Staging(torch, core.device, core.max_seqs, core.max_seqs, self.max_blocks)
        self.step_no = 0
        lo, hi, nl, eot = tokenizer.generable_range()
        self.cand = (lo, hi, nl, eot) if generable_only else (0, 0, -1, -1)
        self.banned_token = -1   # e.g. the EOS id while benchmarking fixed-length turns
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def _sampling(self, st: _Staging):
        self.step_no += 1
        lo, hi, ea, eb = self.cand
        return self.cabi.make_sampling(temperature=0.0, seed=0, step=self.step_no, banned_token=self.banned_token,
                                       allowed_lo=lo,
                                       allowed_hi=hi, extra_a=ea, extra_b=eb, seq_temperature=st.d_temp,
                                       seq_seed=st.d_seed)

    def _fetch(self, st: _Staging, n: int) -> List[int]:
        st.h_out[:n].copy_(st.d_out[:n], non_blocking=True)
        self.stream.synchronize()
        self.d2h_bytes += 4 * n
        return st.h_out[:n].tolist()

    def prefill(self, items: List[PrefillItem]) -> List[Optional[int]]:
        st = self.st_prefill
        with self.torch.cuda.stream(self.stream):
            toks: List[int] = []
            cu = [0]
            for it in items:
                toks += it.tokens
                cu.append(len(toks))
            reqs = [it.req for it in items]
            d_tok, d_cu, d_ctx, d_bt = st.stage(toks, cu, [it.ctx_len for it in items], reqs)
            self.h2d_bytes += 4 * (len(toks) + 2 * len(items) + 1)
            need = any(it.is_last_chunk for it in items)
            self.core.prefill(d_tok, d_cu, d_ctx, d_bt, max(len(it.tokens) for it in items), self._sampling(st),
                              None, st.d_out[:len(items)] if need else None)
            if not need:
                self.stream.synchronize()
                return [None] * len(items)
            out = self._fetch(st, len(items))
        return [out[i] if it.is_last_chunk else None for i, it in enumerate(items)]

    def decode(self, reqs: List[Request]) -> List[int]:
        st = self.st_decode
        with self.torch.cuda.stream(self.stream):
            ctx = [r.ctx_len for r in reqs]
            d_tok, _, d_ctx, d_bt = st.stage([r.generated[-1] for r in reqs], None, ctx, reqs)
            self.h2d_bytes += 8 * len(reqs)
            self.core.decode(d_tok, d_ctx, d_bt, self._sampling(st), st.d_out[:len(reqs)], max_ctx=max(ctx))
            return self._fetch(st, len(reqs))

    def close(self) -> None:
        self.core.close()
