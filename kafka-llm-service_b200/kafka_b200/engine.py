"""Engine — one GPU replica's serving loop: a dedicated host thread that owns the Scheduler and drives
the arithmetic (through a Stepper) with continuous batching.

    asyncio coroutines --submit()--> inbox --engine thread--> admit -> packed incremental prefill
        -> batched decode step -> sampled ids D2H (pinned) -> per-request callbacks

Callbacks are invoked on the engine thread; the provider trampolines them onto its event loop with
``loop.call_soon_threadsafe``. The CUDA calls release the GIL (ctypes), so the event loop keeps
running while the GPU works.

Reference counterpart: the awaited network call of /root/reference/src/llm/portkey.py:379; the
reference has no engine, queue or batching of its own.
"""
from __future__ import annotations

import itertools
import queue
import threading
import time
from typing import Callable, Dict, List, Optional, Sequence, Tuple

from .scheduler import ContextLengthError, PrefillItem, Request, SamplingParams, Scheduler, shard_of


class Stepper:
    """What the engine needs from the arithmetic side. gpu_stepper.GpuStepper binds the C ABI; tests
    bind a deterministic fake to exercise the host logic without a GPU."""
    max_seqs: int
    max_tokens: int
    n_blocks: int
    max_pos: int
    eos_id: int

    def prefill(self, items: List[PrefillItem]) -> List[Optional[int]]:
        raise NotImplementedError

    def decode(self, reqs: List[Request]) -> List[int]:
        raise NotImplementedError

    def close(self) -> None:
        pass


class Engine:
    """Serving loop of one replica. Thread-safe entry points: submit(), Request.cancelled = True,
    shutdown(), stats(). Everything else runs on the engine thread."""

    def __init__(self, stepper: Stepper, max_prefill_tokens: Optional[int] = None, name: str = "kb-engine"):
        self.stepper = stepper
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

    def _run(self) -> None:
        sched = self.sched
        try:
            while not self._stop.is_set():
                self._drain(block=sched.idle())
                sched.admit()
                items = sched.next_prefill_batch()
                if items:
                    t0 = time.perf_counter()
                    sampled = self.stepper.prefill(items)
                    self.timing["prefill_s"] += time.perf_counter() - t0
                    self.timing["prefill_tokens"] += sum(len(i.tokens) for i in items)
                    sched.on_prefill_done(items, sampled)
                reqs = sched.next_decode_batch()
                if reqs:
                    t0 = time.perf_counter()
                    sampled = self.stepper.decode(reqs)
                    self.timing["decode_s"] += time.perf_counter() - t0
                    self.timing["decode_tokens"] += len(reqs)
                    sched.on_decode_done(reqs, sampled)
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
