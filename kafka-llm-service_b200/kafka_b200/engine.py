"""Engine — one GPU replica's serving loop: a dedicated host thread that owns the Scheduler and drives
the arithmetic (through a Stepper) with continuous batching.

    asyncio coroutines --submit()--> inbox --engine thread--> admit -> packed incremental prefill
        -> batched decode step -> sampled ids D2H (pinned) -> per-request callbacks

The loop is a two-stage pipeline: iteration k's prefill + decode are enqueued before iteration k-1's
sampled ids are read back, the ids a step needs from its predecessor being handed over on the device
(kb_select_tokens). A conversation that stops on EOS therefore costs at most two discarded steps.

Callbacks are invoked on the engine thread; the provider trampolines them onto its event loop with
``loop.call_soon_threadsafe``. The CUDA calls release the GIL (ctypes), so the event loop keeps
running while the GPU works.

Reference counterpart: the awaited network call of /root/reference/src/llm/portkey.py:379; the
reference has no engine, queue or batching of its own.
"""
from __future__ import annotations

import gc
import itertools
import os
import queue
import threading
import time
from typing import Callable, Dict, List, Optional, Sequence, Tuple

from .scheduler import ContextLengthError, PrefillItem, Request, SamplingParams, Scheduler, shard_of


def tune_gc() -> None:
    """The serving loop allocates a few thousand small containers per step (plans, events, chunks) next to
    millions of long-lived ones (the token lists of resident conversations): with CPython's default
    thresholds (700, 10, 10) the cyclic collector walks that heap several times per second, which cost
    ~40 % of the host time of a 512-thread turn (tools/host_profile.py). Collect 70x less often; nothing on
    this path builds reference cycles at a rate that matters. KB_GC_DEFAULT=1 leaves the interpreter alone."""
    if os.environ.get("KB_GC_DEFAULT") != "1" and gc.get_threshold()[0] < 50_000:
        gc.set_threshold(50_000, 20, 20)


class Stepper:
    """What the engine needs from the arithmetic side. gpu_stepper.GpuStepper binds the C ABI; tests
    bind a deterministic fake to exercise the host logic without a GPU.

    Work is ENQUEUED with *_launch and its sampled ids are read back later with fetch(), so that the
    engine can enqueue step N+1 while step N runs. Results live in numbered slots of a buffer the
    stepper owns (see slot()): a conversation whose newest id is still in flight (Request.inflight >
    0) is fed from Request.out_slot on the device, everything else from Request.generated[-1]."""
    max_seqs: int
    max_tokens: int
    n_blocks: int
    max_pos: int
    eos_id: int

    def slot(self, parity: int, is_prefill: bool, i: int) -> int:
        """Result slot of item i of the prefill / decode call enqueued with this parity (0 or 1; the
        engine alternates, so a slot is rewritten only after the step that read it has retired)."""
        return (parity * 2 + (1 if is_prefill else 0)) * self.max_seqs + i

    def prefill_launch(self, items: List[PrefillItem], parity: int):
        raise NotImplementedError

    def decode_launch(self, reqs: List[Request], parity: int):
        raise NotImplementedError

    def fetch(self, handle) -> List[Optional[int]]:
        """Wait for that call and return its ids (None for prefill items that are not a last chunk)."""
        raise NotImplementedError

    def prefill(self, items: List[PrefillItem]) -> List[Optional[int]]:
        return self.fetch(self.prefill_launch(items, 0))

    def decode(self, reqs: List[Request]) -> List[int]:
        return self.fetch(self.decode_launch(reqs, 0))

    def close(self) -> None:
        pass


class _Flight:
    """One loop iteration's enqueued work, retired (read back) during the next iteration."""
    __slots__ = ("items", "prefill_handle", "reqs", "decode_handle")

    def __init__(self):
        self.items: List[PrefillItem] = []
        self.prefill_handle = None
        self.reqs: List[Request] = []
        self.decode_handle = None


class Engine:
    """Serving loop of one replica. Thread-safe entry points: submit(), Request.cancelled = True,
    shutdown(), stats(). Everything else runs on the engine thread."""

    def __init__(self, stepper: Stepper, max_prefill_tokens: Optional[int] = None, name: str = "kb-engine",
                 pipeline: bool = True, prefill_first: Optional[bool] = None, max_prefill_streak: int = 16):
        tune_gc()
        self.stepper = stepper
        self.pipeline = pipeline   # False: read every step back before enqueueing the next (for A/B timing)
        # Prompts before tokens: while admitted prompts still wait for (the rest of) their prefill, an iteration
        # launches the prefill chunk only and leaves the running conversations' decode step out — a burst of
        # arrivals gets its first tokens after the prefill work alone instead of prefill + one decode step of an
        # ever larger batch per chunk (512 simultaneous turns: TTFT p50 ~0.37 s -> ~0.2 s), and the decode steps
        # that follow run at the full batch. Bounded: after `max_prefill_streak` such iterations in a row a decode
        # step is forced, so a steady stream of arrivals cannot starve the conversations that are generating
        # (at 4 096-token chunks a stall of at most ~0.7 s). KAFKA_B200_PREFILL_FIRST=0 interleaves as before.
        if prefill_first is None:
            prefill_first = os.environ.get("KAFKA_B200_PREFILL_FIRST", "1") != "0"
        self.prefill_first = prefill_first
        self.max_prefill_streak = max_prefill_streak
        self._prefill_streak = 0
        self.sched = Scheduler(n_blocks=stepper.n_blocks, max_pos=stepper.max_pos, eos_id=stepper.eos_id,
                               max_seqs=stepper.max_seqs,
                               max_prefill_tokens=max_prefill_tokens or stepper.max_tokens)
        self._inbox: "queue.Queue[Optional[Request]]" = queue.Queue()
        self._ids = itertools.count(1)
        self._stop = threading.Event()
        self._error: Optional[BaseException] = None
        self.timing = {"prefill_s": 0.0, "decode_s": 0.0, "prefill_tokens": 0, "decode_tokens": 0}
        self._thread = threading.Thread(target=self._run, name=name, daemon=True)
        self._thread.start()

    # ---- called from any thread ----------------------------------------------------------------
    def submit(self, prompt: Sequence[int], params: SamplingParams, on_event: Callable[[Tuple], None],
               thread_hint: Optional[str] = None) -> Request:
        if self._error is not None:
            raise RuntimeError(f"engine thread died: {self._error!r}")
        if not prompt:
            raise ValueError("empty prompt")
        if len(prompt) + 1 > self.stepper.max_pos:
            raise ContextLengthError(len(prompt), self.stepper.max_pos)
        req = Request(req_id=next(self._ids), prompt=list(prompt), params=params, thread_hint=thread_hint,
                      on_event=on_event, t_submit=time.perf_counter())
        self._inbox.put(req)
        return req

    def shutdown(self) -> None:
        self._stop.set()
        self._inbox.put(None)
        self._thread.join(timeout=30)
        self.stepper.close()

    def stats(self) -> Dict:
        s = dict(self.sched.stats)
        s.update(self.timing)
        s.update({"free_blocks": self.sched.alloc.n_free, "resident_entries": len(self.sched.index.entries),
                  "running": len(self.sched.running), "waiting": len(self.sched.waiting)})
        return s

    @property
    def max_pos(self) -> int:
        return self.stepper.max_pos

    def residency(self, thread_id: Optional[str] = None) -> List[Dict]:
        """Which threads' KV is resident (for the SQLite kv_residency directory); one thread's row when
        `thread_id` is given. Read from another thread while the engine mutates its index: retry if a
        dict changed size mid-iteration."""
        for _ in range(8):
            try:
                return self.sched.residency(thread_id)
            except RuntimeError:
                continue
        return []

    # ---- engine thread -------------------------------------------------------------------------
    def _drain(self, block: bool) -> None:
        try:
            item = self._inbox.get(timeout=0.2) if block else self._inbox.get_nowait()
        except queue.Empty:
            return
        while True:
            if item is not None:
                try:
                    self.sched.add(item)
                except Exception as e:
                    item.on_event(("error", e))
            try:
                item = self._inbox.get_nowait()
            except queue.Empty:
                return

    def _launch(self, parity: int) -> Optional[_Flight]:
        sched, st = self.sched, self.stepper
        f = _Flight()
        f.items = sched.next_prefill_batch()
        if f.items:
            t0 = time.perf_counter()
            f.prefill_handle = st.prefill_launch(f.items, parity)
            sched.on_prefill_launched(f.items, [st.slot(parity, True, i) for i in range(len(f.items))])
            self.timing["prefill_s"] += time.perf_counter() - t0
            self.timing["prefill_tokens"] += sum(len(i.tokens) for i in f.items)
        if self.prefill_first and f.items and sched.prefilling and self._prefill_streak < self.max_prefill_streak:
            self._prefill_streak += 1        # more prompt chunks are ready: they go before the next decode step
            return f
        self._prefill_streak = 0
        f.reqs = sched.next_decode_batch()   # includes conversations whose last prefill chunk was just enqueued
        if f.reqs:
            t0 = time.perf_counter()
            f.decode_handle = st.decode_launch(f.reqs, parity)
            sched.on_decode_launched(f.reqs, [st.slot(parity, False, i) for i in range(len(f.reqs))])
            self.timing["decode_s"] += time.perf_counter() - t0
            self.timing["decode_tokens"] += len(f.reqs)
        return f if (f.items or f.reqs) else None

    def _retire(self, f: _Flight) -> None:
        sched, st = self.sched, self.stepper
        if f.items:
            t0 = time.perf_counter()
            sampled = st.fetch(f.prefill_handle)
            self.timing["prefill_s"] += time.perf_counter() - t0
            sched.on_prefill_done(f.items, sampled)
        if f.reqs:
            t0 = time.perf_counter()
            sampled = st.fetch(f.decode_handle)
            self.timing["decode_s"] += time.perf_counter() - t0
            sched.on_decode_done(f.reqs, sampled)

    def _run(self) -> None:
        """Enqueue iteration k's work, then read iteration k-1's results back while the GPU runs k: the
        host-side token handling (callbacks, stop checks, admission) is hidden behind GPU time."""
        sched = self.sched
        pending: Optional[_Flight] = None
        parity = 0
        stalled = False     # requests are waiting but nothing could be launched: wait for the inbox, do not spin
        try:
            while not self._stop.is_set():
                self._drain(block=(sched.idle() or stalled) and pending is None)
                sched.admit()
                flight = self._launch(parity)
                if flight is not None:
                    parity ^= 1
                if pending is not None:
                    self._retire(pending)
                pending = flight
                stalled = flight is None
                if not self.pipeline and pending is not None:
                    self._retire(pending)
                    pending = None
        except BaseException as e:  # never die silently: fail every pending request
            self._error = e
            for r in sched.waiting + sched.prefilling + sched.running:
                try:
                    r.on_event(("error", e))
                except Exception:
                    pass


class ShardedEngine:
    """N replicas (one per GPU), requests routed by a stable hash of the thread id — no collective,
    nothing crosses NVLink."""

    def __init__(self, engines: List[Engine]):
        self.engines = engines

    def route(self, prompt: Sequence[int], thread_hint: Optional[str]) -> int:
        return shard_of(thread_hint, prompt, len(self.engines))

    def submit(self, prompt, params, on_event, thread_hint=None) -> Request:
        return self.engines[self.route(prompt, thread_hint)].submit(prompt, params, on_event, thread_hint)

    def shutdown(self) -> None:
        for e in self.engines:
            e.shutdown()

    def stats(self) -> Dict:
        return {f"gpu{i}": e.stats() for i, e in enumerate(self.engines)}

    def residency_by_gpu(self, thread_id: Optional[str] = None) -> Dict[int, List[Dict]]:
        return {i: e.residency(thread_id) for i, e in enumerate(self.engines)}

    @property
    def max_pos(self) -> int:
        return self.engines[0].stepper.max_pos
