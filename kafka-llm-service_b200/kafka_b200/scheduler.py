"""Host-side scheduling of one GPU replica: KV block allocation, thread -> KV-block residency,
longest-prefix reuse, continuous batching.  Pure Python, no torch, no CUDA — the arithmetic is
behind a `Stepper` (engine.py binds the C ABI; tests bind a deterministic fake).

Why this exists.  The reference re-sends the whole conversation on every call
(/root/reference/src/agents/base.py:222-223) and never tells the provider which thread it belongs to
(/root/reference/src/llm/base.py:166-174 — no thread_id in the signature).  KV reuse is therefore
keyed on the *token prefix*: a request reuses the resident entry that shares the longest prefix with
its prompt (an optional thread hint only narrows the search), re-prefilling from the first
divergent token.  That also covers the cases where the stored history legitimately differs from what
the model saw (tool rounds not persisted by /chat/completions, compaction, orphan-tool dropping —
SURVEY.md §0 items 8, §3E).
"""
from __future__ import annotations

import hashlib
import itertools
from array import array
import struct
import time
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence, Tuple

BLOCK_TOKENS = 64


def stable_hash64(data: bytes) -> int:
    """Process-independent 64-bit hash (never Python's salted hash())."""
    return int.from_bytes(hashlib.blake2b(data, digest_size=8).digest(), "little")


def shard_of(thread_id: Optional[str], prompt_ids: Sequence[int], n_shards: int) -> int:
    """Which replica owns a request: hash of the thread id, else of the prompt's first 256 tokens."""
    if n_shards <= 1:
        return 0
    if thread_id:
        return stable_hash64(thread_id.encode("utf-8")) % n_shards
    head = prompt_ids[:256]
    return stable_hash64(struct.pack(f"<{len(head)}i", *head)) % n_shards


class ContextLengthError(ValueError):
    """Prompt does not fit the model's context. The message deliberately contains the phrase the
    reference's compaction trigger looks for (/root/reference/src/llm/context_compaction/base.py:42)."""

    def __init__(self, n_tokens: int, max_pos: int):
        super().__init__(f"This model's maximum context length is {max_pos} tokens, however the messages "
                         f"resulted in {n_tokens} tokens (context_length_exceeded)")


# ------------------------------------------------------------------------------------------------
class BlockAllocator:
    def __init__(self, n_blocks: int):
        self.n_blocks = n_blocks
        self._free = list(range(n_blocks - 1, -1, -1))
        self.refs = [0] * n_blocks     # reference counts: a block shared by several conversations is read-only

    @property
    def n_free(self) -> int:
        return len(self._free)

    def alloc(self, n: int) -> Optional[List[int]]:
        if n > len(self._free):
            return None
        out = self._free[-n:][::-1] if n else []
        del self._free[len(self._free) - n:]
        for b in out:
            self.refs[b] = 1
        return out

    def share(self, blocks: Sequence[int]) -> None:
        """Another conversation starts reading these (already written, immutable) blocks."""
        for b in blocks:
            self.refs[b] += 1

    def is_shared(self, block: int) -> bool:
        return self.refs[block] > 1

    def free(self, blocks: Sequence[int]) -> None:
        """Drop one reference per block; a block returns to the pool when nobody reads it any more."""
        for b in blocks:
            self.refs[b] -= 1
            if self.refs[b] == 0:
                self._free.append(b)


@dataclass(eq=False)     # identity, not field-by-field: these sit in lists searched on the hot path
class KVEntry:
    """KV of one conversation resident in HBM: `tokens[i]` is the id whose K/V sits at position i."""
    entry_id: int
    tokens: List[int] = field(default_factory=list)
    blocks: List[int] = field(default_factory=list)
    thread_hint: Optional[str] = None
    busy: bool = False
    last_used: float = 0.0
    chain: List[int] = field(default_factory=list)  # chain[k] = hash of tokens[: (k+1)*BLOCK_TOKENS]


def _chain_hashes(tokens: Sequence[int], prev: Optional[List[int]] = None, start_block: int = 0) -> List[int]:
    """Rolling hash per full block: h_k = H(h_{k-1} || block_k)."""
    out = list(prev[:start_block]) if prev else []
    h = out[-1] if out else 0
    n_full = len(tokens) // BLOCK_TOKENS
    if n_full <= start_block:
        return out
    raw = array("i", tokens[start_block * BLOCK_TOKENS: n_full * BLOCK_TOKENS]).tobytes()  # one conversion
    step = 4 * BLOCK_TOKENS
    for k in range(n_full - start_block):
        h = int.from_bytes(hashlib.blake2b(h.to_bytes(8, "little") + raw[k * step:(k + 1) * step],
                                           digest_size=8).digest(), "little")
        out.append(h)
    return out


class PrefixIndex:
    """Finds the resident entry sharing the longest token prefix with a prompt.

    Every full block boundary of every entry is registered under its chain hash, so a lookup walks
    the prompt's own chain and stops at the deepest registered hash; the partial block after it is
    compared token by token. Entries shorter than one block are found through the thread hint or a
    scan of the (few) short entries."""

    def __init__(self):
        self.by_hash: Dict[int, Dict[int, KVEntry]] = {}
        self.by_hint: Dict[str, KVEntry] = {}
        self.entries: Dict[int, KVEntry] = {}

    def add(self, e: KVEntry, known: Optional[List[int]] = None, n_known: int = 0) -> None:
        """`known`: chain of a sequence that agrees with e.tokens on its first n_known tokens (the prompt's
        chain, already computed at admission) — only the blocks after that are hashed."""
        self.entries[e.entry_id] = e
        if e.thread_hint:
            self.by_hint[e.thread_hint] = e
        start = min(len(known), n_known // BLOCK_TOKENS, len(e.tokens) // BLOCK_TOKENS) if known else 0
        e.chain = _chain_hashes(e.tokens, known, start)
        for h in e.chain:
            self.by_hash.setdefault(h, {})[e.entry_id] = e

    def remove(self, e: KVEntry) -> None:
        self.entries.pop(e.entry_id, None)
        if e.thread_hint and self.by_hint.get(e.thread_hint) is e:
            del self.by_hint[e.thread_hint]
        for h in e.chain:
            d = self.by_hash.get(h)
            if d is not None:
                d.pop(e.entry_id, None)
                if not d:
                    del self.by_hash[h]
        e.chain = []

    @staticmethod
    def _lcp(a: Sequence[int], b: Sequence[int], start: int = 0) -> int:
        n = min(len(a), len(b))
        i = start
        while i + 256 <= n and a[i:i + 256] == b[i:i + 256]:   # slice compares run in C
            i += 256
        while i < n and a[i] == b[i]:
            i += 1
        return i

    def chain_of(self, prompt: Sequence[int], thread_hint: Optional[str]) -> List[int]:
        """The prompt's chain. A follow-up turn of a resident thread re-sends its whole conversation: the
        blocks it has in common with the thread's entry are taken from that entry's chain, not hashed again."""
        e = self.by_hint.get(thread_hint) if thread_hint else None
        if e is None or not e.chain:
            return _chain_hashes(prompt)
        n = self._lcp(e.tokens, prompt)
        return _chain_hashes(prompt, e.chain, min(len(e.chain), n // BLOCK_TOKENS))

    def donor(self, chain: Sequence[int], exclude: Optional[KVEntry] = None,
              prompt: Optional[Sequence[int]] = None) -> Tuple[Optional[KVEntry], int]:
        """Entry (idle OR busy) whose leading full blocks equal the prompt's, and how many. Those blocks
        are immutable — a conversation only ever appends beyond what it registered — so another
        conversation may read them (shared-prefix KV: a common system prompt is prefilled once).
        A 64-bit chain hash only nominates the candidate: with `prompt` given the shared tokens themselves
        are compared, so a hash collision can never hand one thread's context to another."""
        for k in range(len(chain) - 1, -1, -1):
            cands = self.by_hash.get(chain[k])
            if not cands:
                continue
            n = (k + 1) * BLOCK_TOKENS
            for e in cands.values():
                if e is not exclude and len(e.blocks) > k:
                    if prompt is not None and (len(e.tokens) < n or self._lcp(e.tokens, prompt[:n]) < n):
                        continue
                    return e, k + 1
        return None, 0

    def lookup(self, prompt: Sequence[int], thread_hint: Optional[str],
               chain: Optional[Sequence[int]] = None) -> Tuple[Optional[KVEntry], int]:
        """Idle entry to continue IN PLACE (exclusive reuse) and the length of the common prefix."""
        best, best_len = None, 0
        if thread_hint:
            e = self.by_hint.get(thread_hint)
            if e is not None and not e.busy:
                best, best_len = e, self._lcp(e.tokens, prompt)
        if chain is None:
            chain = _chain_hashes(prompt)
        for k in range(len(chain) - 1, -1, -1):
            if (k + 1) * BLOCK_TOKENS <= best_len:
                break
            cands = self.by_hash.get(chain[k])
            if not cands:
                continue
            for e in cands.values():
                if e.busy:
                    continue
                if thread_hint and e.thread_hint and e.thread_hint != thread_hint:
                    continue  # another thread's conversation: never taken over, only read (donor())
                n = self._lcp(e.tokens, prompt, (k + 1) * BLOCK_TOKENS)
                if n > best_len:
                    best, best_len = e, n
            if best_len >= (k + 1) * BLOCK_TOKENS:
                break
        return best, best_len


# ------------------------------------------------------------------------------------------------
@dataclass
class SamplingParams:
    temperature: float = 0.0
    seed: int = 0
    max_tokens: int = 128
    ignore_eos: bool = False
    top_p: float = 0.0          # nucleus sampling when 0 < top_p < 1 (and temperature > 0); 0 = off


@dataclass(eq=False)     # identity, not field-by-field: these sit in lists searched on the hot path
class Request:
    req_id: int
    prompt: List[int]
    params: SamplingParams
    thread_hint: Optional[str]
    on_event: Callable[[Tuple], None]     # ("token", id) | ("finish", reason, usage) | ("error", exc)
    entry: Optional[KVEntry] = None
    n_cached: int = 0                      # prompt tokens served from resident KV
    n_prefilled: int = 0                   # prompt tokens whose KV is known to be resident (>= n_cached)
    n_launched: int = 0                    # prompt tokens whose prefill has been enqueued (>= n_prefilled)
    inflight: int = 0                      # sampled ids enqueued on the GPU and not read back yet
    out_slot: int = -1                     # result slot (Stepper.slot) of the newest in-flight id
    generated: List[int] = field(default_factory=list)
    cancelled: bool = False
    done: bool = False                     # finished or released: results still in flight are dropped
    chain: Optional[List[int]] = None      # block chain hashes of `prompt` (computed once, at admission)
    t_submit: float = 0.0
    t_first: Optional[float] = None

    @property
    def ctx_len(self) -> int:
        """Position of the next token to feed = tokens whose KV is resident or enqueued to become so
        (every id still in flight will have been fed by the time a newly enqueued step runs)."""
        return self.n_launched + max(0, len(self.generated) + self.inflight - 1)


@dataclass(eq=False)     # identity, not field-by-field: these sit in lists searched on the hot path
class PrefillItem:
    req: Request
    tokens: List[int]
    ctx_len: int
    is_last_chunk: bool


class Scheduler:
    """Admission + batching policy. Single-threaded: only the engine thread calls it."""

    def __init__(self, n_blocks: int, max_pos: int, eos_id: int, max_seqs: int, max_prefill_tokens: int,
                 max_blocks_per_seq: Optional[int] = None):
        self.alloc = BlockAllocator(n_blocks)
        self.index = PrefixIndex()
        self.max_pos = max_pos
        self.eos_id = eos_id
        self.max_seqs = max_seqs
        self.max_prefill_tokens = max_prefill_tokens
        self.max_blocks_per_seq = max_blocks_per_seq or (max_pos + BLOCK_TOKENS - 1) // BLOCK_TOKENS
        self.waiting: List[Request] = []
        self.prefilling: List[Request] = []
        self.running: List[Request] = []
        self._ids = itertools.count(1)
        self.stats = {"requests": 0, "prompt_tokens": 0, "cached_tokens": 0, "generated_tokens": 0,
                      "evictions": 0, "prefill_calls": 0, "decode_steps": 0, "shared_tokens": 0}

    # ---- admission ---------------------------------------------------------------------------
    def add(self, req: Request) -> None:
        if len(req.prompt) + 1 > self.max_pos:
            raise ContextLengthError(len(req.prompt), self.max_pos)
        req.params.max_tokens = max(1, min(req.params.max_tokens, self.max_pos - len(req.prompt)))
        need = (len(req.prompt) + req.params.max_tokens + BLOCK_TOKENS - 1) // BLOCK_TOKENS
        if need > self.alloc.n_blocks:       # can never be resident, whatever gets evicted: fail now, not never
            raise MemoryError(f"request needs {need} KV blocks but the pool has {self.alloc.n_blocks}")
        self.waiting.append(req)
        self.stats["requests"] += 1

    def _evict_until(self, need_blocks: int, keep: Sequence[Optional[KVEntry]] = ()) -> bool:
        if self.alloc.n_free >= need_blocks:
            return True
        idle = sorted((e for e in self.index.entries.values() if not e.busy and all(e is not k for k in keep)),
                      key=lambda e: e.last_used)
        for e in idle:
            self.index.remove(e)
            self.alloc.free(e.blocks)
            e.blocks = []
            self.stats["evictions"] += 1
            if self.alloc.n_free >= need_blocks:
                return True
        return self.alloc.n_free >= need_blocks

    def admit(self) -> None:
        """Move waiting requests into the prefill set while sequence slots and KV blocks allow."""
        still = []
        for req in self.waiting:
            if req.cancelled:
                req.done = True          # dropped before it ever held a slot or a block: nothing to release
                continue
            if len(self.prefilling) + len(self.running) >= self.max_seqs:
                still.append(req)
                continue
            chain = req.chain = self.index.chain_of(req.prompt, req.thread_hint)
            entry, lcp = self.index.lookup(req.prompt, req.thread_hint, chain)
            lcp = min(lcp, len(req.prompt) - 1)  # at least one token must run to produce logits
            donor, d_blocks = self.index.donor(chain, exclude=entry, prompt=req.prompt)
            d_blocks = min(d_blocks, (len(req.prompt) - 1) // BLOCK_TOKENS)
            total = len(req.prompt) + req.params.max_tokens
            need_total = (total + BLOCK_TOKENS - 1) // BLOCK_TOKENS
            share = donor is not None and d_blocks * BLOCK_TOKENS > lcp
            if share:
                # read the donor's leading blocks, own everything after them
                entry, lcp = None, d_blocks * BLOCK_TOKENS
                extra, replace = need_total - d_blocks, []
            else:
                # continue this conversation's own entry in place. Blocks it shares with others are
                # read-only: if writing resumes inside one, fall back to its first token and take fresh
                # blocks from there on (copy-on-write by re-prefill of at most 63 tokens)
                blocks = entry.blocks if entry is not None else []
                first = lcp // BLOCK_TOKENS
                if first < len(blocks) and self.alloc.is_shared(blocks[first]):
                    lcp = first * BLOCK_TOKENS
                replace = [i for i in range(lcp // BLOCK_TOKENS, len(blocks)) if self.alloc.is_shared(blocks[i])]
                extra = max(0, need_total - len(blocks)) + len(replace)
            if not self._evict_until(extra, keep=(entry, donor)):
                if self.prefilling or self.running:
                    still.append(req)             # blocks come back when a running conversation finishes
                    continue
                # Nothing is running, so nothing will ever be freed by waiting: give up the reuse (the kept
                # entry / donor is what stands in the way) and take the request as a new conversation.
                entry, donor, share, lcp, replace, extra = None, None, False, 0, [], need_total
                if not self._evict_until(extra):
                    req.on_event(("error", MemoryError(
                        f"request needs {need_total} KV blocks but only {self.alloc.n_free} of "
                        f"{self.alloc.n_blocks} can be freed")))
                    continue
            if share:
                entry = KVEntry(entry_id=next(self._ids), thread_hint=req.thread_hint,
                                blocks=list(donor.blocks[:d_blocks]))
                self.alloc.share(entry.blocks)
                self.stats["shared_tokens"] += lcp
            elif entry is None:
                entry = KVEntry(entry_id=next(self._ids), thread_hint=req.thread_hint)
                lcp = 0
            else:
                self.index.remove(entry)          # re-registered (with its new tokens) when prefilled
                entry.thread_hint = req.thread_hint or entry.thread_hint
            fresh = self.alloc.alloc(extra)
            for i in replace:                     # swap shared blocks that would be overwritten
                self.alloc.free([entry.blocks[i]])
                entry.blocks[i] = fresh.pop()
            entry.blocks += fresh
            entry.tokens = list(req.prompt[:lcp])
            entry.busy = True
            req.entry = entry
            req.n_cached = req.n_prefilled = req.n_launched = lcp
            self.stats["prompt_tokens"] += len(req.prompt)
            self.stats["cached_tokens"] += lcp
            self.prefilling.append(req)
        self.waiting = still

    # ---- batching ----------------------------------------------------------------------------
    # The engine enqueues a step (next_*_batch -> Stepper.*_launch -> on_*_launched) and reads its
    # results back one iteration later (Stepper.fetch -> on_*_done), after the following step has been
    # enqueued. Everything the next launch needs (positions, membership) is therefore derived from
    # launch-time counters; sampled ids are only needed on the host to detect EOS / stop conditions.
    def next_prefill_batch(self) -> List[PrefillItem]:
        items: List[PrefillItem] = []
        budget = self.max_prefill_tokens
        for req in self.prefilling:
            if req.cancelled or budget <= 0:
                continue
            remaining = len(req.prompt) - req.n_launched
            n = min(remaining, budget)
            if n <= 0:
                continue
            items.append(PrefillItem(req, req.prompt[req.n_launched:req.n_launched + n], req.n_launched,
                                     n == remaining))
            budget -= n
        return items

    def on_prefill_launched(self, items: List[PrefillItem], slots: Sequence[int]) -> None:
        for it, slot in zip(items, slots):
            req = it.req
            req.n_launched += len(it.tokens)
            if it.is_last_chunk:                 # its first id is in flight: it can join decode steps now
                self.prefilling.remove(req)
                self.running.append(req)
                req.inflight += 1
                req.out_slot = slot

    def on_prefill_done(self, items: List[PrefillItem], sampled: Sequence[Optional[int]]) -> None:
        self.stats["prefill_calls"] += 1
        for it, tok in zip(items, sampled):
            req = it.req
            if req.done:
                continue
            req.n_prefilled += len(it.tokens)
            if it.is_last_chunk:
                req.inflight -= 1
                # the prompt's KV is complete and immutable from here on: let other conversations that
                # start with the same blocks (a common system prompt) read it while this one decodes
                if req.entry is not None:
                    req.entry.tokens = list(req.prompt)
                    self.index.add(req.entry, req.chain, len(req.prompt))
                if not req.cancelled:
                    self._on_token(req, int(tok))
        self._reap()

    def next_decode_batch(self) -> List[Request]:
        """Running conversations that still need a token beyond those already in flight."""
        self._reap()
        return [r for r in self.running if len(r.generated) + r.inflight < r.params.max_tokens]

    def on_decode_launched(self, reqs: List[Request], slots: Sequence[int]) -> None:
        for req, slot in zip(reqs, slots):
            req.inflight += 1
            req.out_slot = slot

    def on_decode_done(self, reqs: List[Request], sampled: Sequence[int]) -> None:
        self.stats["decode_steps"] += 1
        for req, tok in zip(reqs, sampled):
            if req.done:
                continue                          # finished on an earlier id: this step's id is discarded
            req.inflight -= 1
            if not req.cancelled:
                self._on_token(req, int(tok))
        self._reap()

    # ---- token handling / completion -----------------------------------------------------------
    def _on_token(self, req: Request, tok: int) -> None:
        if req.t_first is None:
            req.t_first = time.perf_counter()
        req.generated.append(tok)
        self.stats["generated_tokens"] += 1
        if tok == self.eos_id and not req.params.ignore_eos:
            self._finish(req, "stop")
            return
        req.on_event(("token", tok))
        if len(req.generated) >= req.params.max_tokens:
            self._finish(req, "length")

    def _finish(self, req: Request, reason: str) -> None:
        usage = {"prompt_tokens": len(req.prompt), "completion_tokens": len(req.generated),
                 "total_tokens": len(req.prompt) + len(req.generated), "cached_tokens": req.n_cached}
        req.on_event(("finish", reason, usage))
        self._release(req)

    def _release(self, req: Request) -> None:
        req.done = True
        if req in self.running:
            self.running.remove(req)
        if req in self.prefilling:
            self.prefilling.remove(req)
        e = req.entry
        if e is None:
            return
        req.entry = None
        self.index.remove(e)   # registered as a read-only donor while it was running; re-added below
        # KV is resident for the prefilled prompt part and for every generated token that was fed back
        # (steps still in flight only write beyond that, and anything enqueued later runs after them)
        e.tokens = req.prompt[:req.n_prefilled] + req.generated[:max(0, len(req.generated) - 1)]
        keep = (len(e.tokens) + BLOCK_TOKENS - 1) // BLOCK_TOKENS
        if keep < len(e.blocks):
            self.alloc.free(e.blocks[keep:])
            e.blocks = e.blocks[:keep]
        e.busy = False
        e.last_used = time.monotonic()
        if e.tokens:
            self.index.add(e, req.chain, min(req.n_prefilled, len(req.prompt)))
        else:
            self.alloc.free(e.blocks)
            e.blocks = []

    def _reap(self) -> None:
        for req in [r for r in self.running + self.prefilling if r.cancelled and r.entry is not None]:
            self._release(req)

    def idle(self) -> bool:
        return not (self.waiting or self.prefilling or self.running)

    def residency(self, thread_id: Optional[str] = None) -> List[Dict]:
        """Snapshot for the SQLite kv_residency table (db.py): all entries, or one thread's."""
        out = []
        if thread_id is not None:
            e = self.index.by_hint.get(thread_id)
            entries = [e] if e is not None else []
        else:
            entries = list(self.index.entries.values())
        for e in entries:
            out.append({"thread_id": e.thread_hint, "entry_id": e.entry_id, "n_tokens": len(e.tokens),
                        "prefix_hash": e.chain[-1] if e.chain else 0, "block_ids": list(e.blocks),
                        "last_used": e.last_used})
        return out
