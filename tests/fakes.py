"""Test doubles for the host-side logic (no GPU, no torch).

FakeStepper models the *memory* semantics of the CUDA path, not its arithmetic: it keeps a paged
"KV" store of token ids, written at (block table, position) exactly where the kernels would write K/V,
and derives the next token from the sequence it READS BACK through the block table. Any bookkeeping
mistake in the scheduler (wrong ctx_len, stale or shared blocks, bad prefix reuse) therefore changes
the generated tokens, which tests compare with `expected_tokens` — a pure function of the token
sequence.
"""
from __future__ import annotations

import hashlib
import struct
from typing import List, Optional, Sequence

from kafka_b200.engine import Stepper
from kafka_b200.scheduler import BLOCK_TOKENS, PrefillItem, Request


def next_token(seq: Sequence[int], eos_id: int, eos_every: int = 0) -> int:
    h = int.from_bytes(hashlib.blake2b(struct.pack(f"<{len(seq)}i", *seq), digest_size=8).digest(), "little")
    if eos_every and h % eos_every == 0:
        return eos_id
    return 32 + (h >> 8) % 95  # printable ASCII


def expected_tokens(prompt: Sequence[int], n: int, eos_id: int, eos_every: int = 0) -> List[int]:
    seq, out = list(prompt), []
    for _ in range(n):
        t = next_token(seq, eos_id, eos_every)
        out.append(t)
        if t == eos_id:
            break
        seq.append(t)
    return out


class FakeStepper(Stepper):
    def __init__(self, n_blocks=64, max_seqs=8, max_tokens=256, max_pos=1024, eos_id=511, eos_every=0):
        self.n_blocks, self.max_seqs, self.max_tokens = n_blocks, max_seqs, max_tokens
        self.max_pos, self.eos_id, self.eos_every = max_pos, eos_id, eos_every
        self.kv = [[None] * BLOCK_TOKENS for _ in range(n_blocks)]
        self.prefill_calls: List[List[int]] = []   # tokens per item of every prefill call
        self.decode_batches: List[int] = []
        self.prefilled_tokens = 0
        self.res: dict = {}            # result slot -> sampled id
        self.outstanding = 0           # launched calls not fetched yet

    def _write(self, blocks, pos, tok):
        self.kv[blocks[pos // BLOCK_TOKENS]][pos % BLOCK_TOKENS] = tok

    def _read(self, blocks, n):
        seq = [self.kv[blocks[p // BLOCK_TOKENS]][p % BLOCK_TOKENS] for p in range(n)]
        assert None not in seq, "read of a KV position that was never written"
        return seq

    # Work is executed at launch time, in launch order — exactly what the CUDA stream does — and the
    # results are parked in the slots the engine named, to be handed out by fetch().
    def prefill_launch(self, items: List[PrefillItem], parity: int):
        assert sum(len(i.tokens) for i in items) <= self.max_tokens
        assert len(items) <= self.max_seqs
        self.prefill_calls.append([len(i.tokens) for i in items])
        slots: List[Optional[int]] = []
        for i, it in enumerate(items):
            blocks = it.req.entry.blocks
            for k, t in enumerate(it.tokens):
                self._write(blocks, it.ctx_len + k, t)
            self.prefilled_tokens += len(it.tokens)
            if it.is_last_chunk:
                s = self.slot(parity, True, i)
                self.res[s] = next_token(self._read(blocks, it.ctx_len + len(it.tokens)), self.eos_id, self.eos_every)
                slots.append(s)
            else:
                slots.append(None)
        self.outstanding += 1
        assert self.outstanding <= 4, "engine enqueued more than two iterations without reading back"
        return slots

    def decode_launch(self, reqs: List[Request], parity: int):
        assert len(reqs) <= self.max_seqs
        self.decode_batches.append(len(reqs))
        fed = []
        for r in reqs:   # gather every input before any output is written (kb_select_tokens runs first)
            fed.append(self.res[r.out_slot] if r.inflight > 0 else r.generated[-1])
        slots = []
        for i, (r, tok) in enumerate(zip(reqs, fed)):
            blocks = r.entry.blocks
            self._write(blocks, r.ctx_len, tok)
            s = self.slot(parity, False, i)
            self.res[s] = next_token(self._read(blocks, r.ctx_len + 1), self.eos_id, self.eos_every)
            slots.append(s)
        self.outstanding += 1
        assert self.outstanding <= 4, "engine enqueued more than two iterations without reading back"
        return slots

    def fetch(self, handle) -> List[Optional[int]]:
        self.outstanding -= 1
        return [None if s is None else self.res[s] for s in handle]
