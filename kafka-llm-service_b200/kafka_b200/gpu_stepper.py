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
        self.h_top, self.d_top = pair(S, torch.float32)
        self.nucleus = False      # some sequence of the current membership samples with 0 < top_p < 1
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

    def plan(self, reqs: Sequence[Request]) -> dict:
        """Host-only: which per-membership arrays have to be uploaded again for this batch."""
        key = tuple((r.req_id, len(r.entry.blocks)) for r in reqs)
        mb = max(1, max(len(r.entry.blocks) for r in reqs))
        p = {"mb": mb}
        if key != self.bt_key:
            flat: List[int] = []
            for r in reqs:
                b = r.entry.blocks
                flat += b
                if len(b) < mb:
                    flat += [0] * (mb - len(b))
            p["bt"] = flat
            self.bt_key = key
        if key != self.sp_key:
            p["temp"] = [r.params.temperature for r in reqs]
            p["seed"] = [(r.params.seed * 0x9E3779B97F4A7C15 + r.req_id) & 0x7FFFFFFFFFFFFFFF for r in reqs]
            p["top_p"] = [r.params.top_p if 0.0 < r.params.top_p < 1.0 else 0.0 for r in reqs]
            self.nucleus = any(p["top_p"])
            self.sp_key = key
        p["nucleus"] = self.nucleus
        return p

    def plan_src(self, src: List[int]) -> Optional[List[int]]:
        """Result slots the in-flight conversations' ids come from (-1: the host staged the id); None
        when the device copy is still right."""
        if src == self.src_key:
            return None
        self.src_key = src
        return src

    def apply(self, n: int, tokens: Optional[List[int]], cu: Optional[List[int]], ctx: List[int], p: dict):
        """Upload what the plan carries; returns the device views the call needs."""
        torch = self.torch
        d_tok = self._up(self.h_tok, self.d_tok, tokens, torch.int32, torch) if tokens is not None else self.d_tok[:n]
        d_cu = self._up(self.h_cu, self.d_cu, cu, torch.int32, torch) if cu is not None else None
        d_ctx = self._up(self.h_ctx, self.d_ctx, ctx, torch.int32, torch)
        if "bt" in p:
            self._up(self.h_bt, self.d_bt, p["bt"], torch.int32, torch)
        if "temp" in p:
            self._up(self.h_temp, self.d_temp, p["temp"], torch.float32, torch)
            self._up(self.h_seed, self.d_seed, p["seed"], torch.int64, torch)
            self._up(self.h_top, self.d_top, p.get("top_p") or [0.0] * len(p["temp"]), torch.float32, torch)
        if p.get("src") is not None:
            self._up(self.h_src, self.d_src, p["src"], torch.int32, torch)
        mb = p["mb"]
        return d_tok, d_cu, d_ctx, self.d_bt[:n * mb].view(n, mb)


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

    def fetch(self, h: _Handle) -> List[Optional[int]]:
        h.event.synchronize()
        if h.last is None:                      # decode step: one id per sequence
            return self.h_res[h.base:h.base + h.n].tolist()
        if not h.n:
            return [None] * len(h.last)
        out = self.h_res[h.base:h.base + h.n].tolist()
        return [out[i] if is_last else None for i, is_last in enumerate(h.last)]

    # A launch is split in two: plan_*() turns the scheduler's objects into a plain dict (host only),
    # run_plan() uploads and enqueues it. With tensor parallelism the leader publishes the dict and every
    # rank runs the same plan (tp_serve.py); on one GPU `publish` stays None.
    publish = None

    def plan_prefill(self, items: List[PrefillItem], parity: int) -> dict:
        toks: List[int] = []
        cu = [0]
        for it in items:
            toks += it.tokens
            cu.append(len(toks))
        p = self.st_prefill[parity].plan([it.req for it in items])
        self.step_no += 1
        p.update(op="prefill", parity=parity, n=len(items), tokens=toks, cu=cu, ctx=[it.ctx_len for it in items],
                 last=[it.is_last_chunk for it in items], max_new=max(len(it.tokens) for it in items),
                 step=self.step_no, banned=self.banned_token)
        return p

    def plan_decode(self, reqs: List[Request], parity: int) -> dict:
        st = self.st_decode[parity]
        n = len(reqs)
        ctx = [r.ctx_len for r in reqs]
        src = [r.out_slot if r.inflight > 0 else -1 for r in reqs]
        n_dev = n - src.count(-1)
        p = st.plan(reqs)
        self.step_no += 1
        p.update(op="decode", parity=parity, n=n, ctx=ctx, select=n_dev > 0, src=st.plan_src(src) if n_dev else None,
                 tokens=None if n_dev == n else [r.generated[-1] if r.inflight == 0 else 0 for r in reqs],
                 step=self.step_no, banned=self.banned_token)
        return p

    def run_plan(self, p: dict, read_back: bool = True) -> _Handle:
        parity, n = p["parity"], p["n"]
        lo, hi, ea, eb = self.cand
        with self.torch.cuda.stream(self.stream):
            if p["op"] == "prefill":
                st = self.st_prefill[parity]
                d_tok, d_cu, d_ctx, d_bt = st.apply(n, p["tokens"], p["cu"], p["ctx"], p)
                self.h2d_bytes += 4 * (len(p["tokens"]) + 2 * n + 1)
                last = p["last"]
                n_out = n if any(last) else 0
                base = self.slot(parity, True, 0)
            else:
                st = self.st_decode[parity]
                d_tok, d_cu, d_ctx, d_bt = st.apply(n, p["tokens"], None, p["ctx"], p)
                self.h2d_bytes += 4 * n + (4 * n if p["tokens"] is not None else 0)
                if p["select"]:
                    self.cabi.select_tokens(d_tok, self.d_res, st.d_src[:n])
                last, n_out = None, n
                base = self.slot(parity, False, 0)
            sp = self.cabi.make_sampling(temperature=0.0, seed=0, step=p["step"], banned_token=p["banned"],
                                         allowed_lo=lo, allowed_hi=hi, extra_a=ea, extra_b=eb,
                                         seq_temperature=st.d_temp, seq_seed=st.d_seed,
                                         # nucleus sampling needs the whole vocabulary on this GPU: with a
                                         # vocabulary-sharded lm_head (tensor parallel) top_p is not applied
                                         seq_top_p=st.d_top if (p.get("nucleus") and not self.core.lm_head_slice)
                                         else None)
            out = self.d_res[base:base + n_out] if n_out else None
            if p["op"] == "prefill":
                self.core.prefill(d_tok, d_cu, d_ctx, d_bt, p["max_new"], sp, None, out)
            else:
                self.core.decode(d_tok, d_ctx, d_bt, sp, out, max_ctx=max(p["ctx"]))
            if n_out and read_back:
                self.h_res[base:base + n_out].copy_(self.d_res[base:base + n_out], non_blocking=True)
                self.d2h_bytes += 4 * n_out
            st.event.record(self.stream)
            return _Handle(st.event, base, n_out, last)

    def prefill_launch(self, items: List[PrefillItem], parity: int) -> _Handle:
        p = self.plan_prefill(items, parity)
        if self.publish is not None:
            self.publish(p)
        return self.run_plan(p)

    def decode_launch(self, reqs: List[Request], parity: int) -> _Handle:
        p = self.plan_decode(reqs, parity)
        if self.publish is not None:
            self.publish(p)
        return self.run_plan(p)

    def close(self) -> None:
        self.core.close()
