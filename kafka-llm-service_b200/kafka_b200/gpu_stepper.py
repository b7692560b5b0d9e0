"""GpuStepper — binds the scheduler's batches to the C ABI (kb_prefill / kb_decode).

Per call it stages the small integer arguments (token ids, cu_seqlens, ctx_lens, block tables,
per-sequence sampling parameters) in pinned host memory, copies them to persistent device buffers on
its own stream, enqueues the call, and brings the sampled ids back through pinned memory. PyTorch
is the allocator and stream provider only. Block tables and sampling arrays are re-uploaded only when
the batch membership changes (decode steps of a stable batch upload B ints).

Nothing here waits for the GPU except fetch(): *_launch only enqueues, and the ids a decode step needs
from a step that has not been read back yet are taken on the device from the result slots
(kb_select_tokens). There are two staging sets per path, alternated by the engine's parity, so the host
never rewrites pinned memory that an enqueued copy may still read.

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
        self.h_src, self.d_src = pair(S, i32)
        self.bt_key = None
        self.sp_key = None
        self.src_key = None
        self.event = torch.cuda.Event()

    @staticmethod
    def _up(h, d, values, dtype, torch):
        n = len(values)
        h[:n] = torch.tensor(values, dtype=dtype)
        d[:n].copy_(h[:n], non_blocking=True)
        return d[:n]

    def stage(self, tokens: Optional[List[int]], cu: Optional[List[int]], ctx: List[int], reqs: Sequence[Request]):
        torch = self.torch
        n = len(reqs)
        d_tok = self._up(self.h_tok, self.d_tok, tokens, torch.int32, torch) if tokens is not None else self.d_tok[:n]
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

    def stage_src(self, src: List[int]):
        """Result slots the in-flight conversations' ids come from (-1: the host staged the id)."""
        if src != self.src_key:
            self._up(self.h_src, self.d_src, src, self.torch.int32, self.torch)
            self.src_key = src
        return self.d_src[:len(src)]


class _Handle:
    __slots__ = ("event", "base", "n", "last")

    def __init__(self, event, base, n, last):
        self.event, self.base, self.n, self.last = event, base, n, last


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
        S = core.max_seqs
        self.st_prefill = [_Staging(torch, core.device, core.max_tokens, S, self.max_blocks) for _ in range(2)]
        self.st_decode = [_Staging(torch, core.device, S, S, self.max_blocks) for _ in range(2)]
        # result slots (Stepper.slot): [parity][decode | prefill][max_seqs] sampled ids
        self.d_res = torch.zeros(4 * S, dtype=torch.int32, device=core.device)
        self.h_res = torch.zeros(4 * S, dtype=torch.int32, pin_memory=True)
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

    def _finish_launch(self, st: _Staging, base: int, n: int, last) -> _Handle:
        if n:
            self.h_res[base:base + n].copy_(self.d_res[base:base + n], non_blocking=True)
            self.d2h_bytes += 4 * n
        st.event.record(self.stream)
        return _Handle(st.event, base, n, last)

    def fetch(self, h: _Handle) -> List[Optional[int]]:
        h.event.synchronize()
        if h.last is None:                      # decode step: one id per sequence
            return self.h_res[h.base:h.base + h.n].tolist()
        if not h.n:
            return [None] * len(h.last)
        out = self.h_res[h.base:h.base + h.n].tolist()
        return [out[i] if is_last else None for i, is_last in enumerate(h.last)]

    def prefill_launch(self, items: List[PrefillItem], parity: int) -> _Handle:
        st = self.st_prefill[parity]
        with self.torch.cuda.stream(self.stream):
            toks: List[int] = []
            cu = [0]
            for it in items:
                toks += it.tokens
                cu.append(len(toks))
            reqs = [it.req for it in items]
            d_tok, d_cu, d_ctx, d_bt = st.stage(toks, cu, [it.ctx_len for it in items], reqs)
            self.h2d_bytes += 4 * (len(toks) + 2 * len(items) + 1)
            last = [it.is_last_chunk for it in items]
            n = len(items) if any(last) else 0
            base = self.slot(parity, True, 0)
            self.core.prefill(d_tok, d_cu, d_ctx, d_bt, max(len(it.tokens) for it in items), self._sampling(st),
                              None, self.d_res[base:base + n] if n else None)
            return self._finish_launch(st, base, n, last)

    def decode_launch(self, reqs: List[Request], parity: int) -> _Handle:
        st = self.st_decode[parity]
        n = len(reqs)
        with self.torch.cuda.stream(self.stream):
            ctx = [r.ctx_len for r in reqs]
            src = [r.out_slot if r.inflight > 0 else -1 for r in reqs]
            n_dev = n - src.count(-1)
            toks = None if n_dev == n else [r.generated[-1] if r.inflight == 0 else 0 for r in reqs]
            d_tok, _, d_ctx, d_bt = st.stage(toks, None, ctx, reqs)
            self.h2d_bytes += 4 * n + (4 * n if toks is not None else 0)
            if n_dev:
                self.cabi.select_tokens(d_tok, self.d_res, st.stage_src(src))
            base = self.slot(parity, False, 0)
            self.core.decode(d_tok, d_ctx, d_bt, self._sampling(st), self.d_res[base:base + n], max_ctx=max(ctx))
            return self._finish_launch(st, base, n, None)

    def close(self) -> None:
        self.core.close()
