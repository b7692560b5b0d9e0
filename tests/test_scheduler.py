"""Host logic of one replica: block allocation, thread -> KV residency, longest-prefix reuse,
continuous batching, eviction, cancellation — on CPU against FakeStepper (tests/fakes.py)."""
import random
import threading
import time

import pytest

from fakes import FakeStepper, expected_tokens
from kafka_b200.engine import Engine, ShardedEngine
from kafka_b200.scheduler import (BLOCK_TOKENS, BlockAllocator, ContextLengthError, KVEntry, PrefixIndex,
                                  SamplingParams, shard_of, stable_hash64)

EOS = 511


class Collector:
    def __init__(self):
        self.tokens, self.finish, self.usage, self.error = [], None, None, None
        self.done = threading.Event()

    def __call__(self, ev):
        if ev[0] == "token":
            self.tokens.append(ev[1])
        elif ev[0] == "finish":
            self.finish, self.usage = ev[1], ev[2]
            self.done.set()
        else:
            self.error = ev[1]
            self.done.set()

    def wait(self, t=10):
        assert self.done.wait(t), "request did not finish"
        return self


def run(engine, prompt, n=8, hint=None, **kw):
    c = Collector()
    engine.submit(prompt, SamplingParams(max_tokens=n, **kw), c, thread_hint=hint)
    return c.wait()


@pytest.fixture
def engine():
    st = FakeStepper(n_blocks=64, max_seqs=8, max_tokens=256, max_pos=1024, eos_id=EOS)
    e = Engine(st, max_prefill_tokens=128)
    yield e
    e.shutdown()


def test_block_allocator():
    a = BlockAllocator(8)
    x = a.alloc(3)
    y = a.alloc(5)
    assert sorted(x + y) == list(range(8)) and a.alloc(1) is None and a.alloc(0) == []
    a.free(x)
    assert a.n_free == 3 and sorted(a.alloc(3)) == sorted(x)


def test_stable_hash_and_sharding():
    assert stable_hash64(b"thread-1") == stable_hash64(b"thread-1")
    assert stable_hash64(b"bench-0001") == 0xE1133E67972CA3C7   # pinned: must never depend on the process
    assert [shard_of(f"bench-{i:04d}", [], 8) for i in range(8)] == [6, 7, 7, 2, 4, 4, 5, 4]
    owners = [shard_of(f"bench-{i:04d}", [1, 2, 3], 8) for i in range(4096)]
    counts = [owners.count(g) for g in range(8)]
    assert min(counts) > 400 and max(counts) < 620            # 4096 threads spread over 8 GPUs
    assert shard_of("t", [1], 1) == 0
    assert shard_of(None, list(range(300)), 4) == shard_of(None, list(range(256)) + [9] * 44, 4)  # first 256 tokens


def test_prefix_index_longest_match():
    idx = PrefixIndex()
    base = list(range(200))
    e1 = KVEntry(1, tokens=base[:130], blocks=[0, 1, 2])
    e2 = KVEntry(2, tokens=base[:70] + [999] * 30, blocks=[3, 4])
    e3 = KVEntry(3, tokens=[7, 7, 7], blocks=[5], thread_hint="short")
    for e in (e1, e2, e3):
        idx.add(e)
    e, n = idx.lookup(base[:150], None)
    assert e is e1 and n == 130
    e, n = idx.lookup(base[:70] + [999] * 10 + [5], None)
    assert e is e2 and n == 80
    e, n = idx.lookup([7, 7, 8], "short")                      # shorter than a block: found through the hint
    assert e is e3 and n == 2
    e1.busy = True
    e, n = idx.lookup(base[:150], None)
    assert e is e2 and n == 70                                  # busy entries are never shared
    idx.remove(e2)
    e1.busy = False
    idx.remove(e1)
    assert idx.lookup(base[:150], None) == (None, 0)


def test_single_request_matches_pure_function(engine):
    prompt = list(range(40, 140))
    c = run(engine, prompt, n=12)
    assert c.tokens == expected_tokens(prompt, 12, EOS) and c.finish == "length"
    assert c.usage == {"prompt_tokens": 100, "completion_tokens": 12, "total_tokens": 112, "cached_tokens": 0}


def test_turns_reuse_resident_prefix(engine):
    """Second turn = first prompt + the generated reply + a new user message: only the new part is prefilled."""
    st = engine.stepper
    p1 = list(range(50, 250))                                    # 200 tokens
    c1 = run(engine, p1, n=10, hint="thread-A")
    before = st.prefilled_tokens
    p2 = p1 + c1.tokens + [EOS] + list(range(60, 90))            # reply is re-templated: tokens + <eot> + new msg
    c2 = run(engine, p2, n=10, hint="thread-A")
    assert c2.tokens == expected_tokens(p2, 10, EOS)
    # resident: p1 + first 9 generated tokens (the 10th was sampled but never fed back)
    assert c2.usage["cached_tokens"] == len(p1) + 9
    assert st.prefilled_tokens - before == len(p2) - (len(p1) + 9)
    # without the hint the prefix is still found (the reference never forwards thread ids)
    p3 = p2 + c2.tokens + [EOS, 65, 66]
    c3 = run(engine, p3, n=4)
    assert c3.usage["cached_tokens"] == len(p2) + 9 and c3.tokens == expected_tokens(p3, 4, EOS)


def test_divergent_history_reprefills_from_divergence(engine):
    p1 = list(range(30, 230))
    run(engine, p1, n=6, hint="t")
    p2 = p1[:150] + [400, 401, 402] + list(range(70, 100))       # history edited at position 150 (compaction etc.)
    c2 = run(engine, p2, n=6, hint="t")
    assert c2.usage["cached_tokens"] == 150 and c2.tokens == expected_tokens(p2, 6, EOS)


def test_identical_prompt_recomputes_last_token(engine):
    p = list(range(100, 190))
    a = run(engine, p, n=3, hint="x")
    b = run(engine, p, n=3, hint="x")                           # whole prompt resident: one token still runs
    assert a.tokens == b.tokens and b.usage["cached_tokens"] == len(p) - 1


def test_concurrent_requests_batch_and_stay_isolated(engine):
    prompts = [list(range(33 + i, 33 + i + 70 + 13 * i)) for i in range(8)]
    cols = [Collector() for _ in prompts]
    for p, c in zip(prompts, cols):
        engine.submit(p, SamplingParams(max_tokens=20), c, thread_hint=None)
    for p, c in zip(prompts, cols):
        c.wait()
        assert c.tokens == expected_tokens(p, 20, EOS)
    assert max(engine.stepper.decode_batches) > 1               # they really shared decode steps
    assert all(sum(call) <= 128 for call in engine.stepper.prefill_calls)   # chunked to the token budget


def test_more_requests_than_slots_queue_up(engine):
    prompts = [[40 + i] * 30 for i in range(20)]                 # 20 requests, 8 slots
    cols = [Collector() for _ in prompts]
    for p, c in zip(prompts, cols):
        engine.submit(p, SamplingParams(max_tokens=5), c)
    for p, c in zip(prompts, cols):
        assert c.wait().tokens == expected_tokens(p, 5, EOS)
    assert max(engine.stepper.decode_batches) <= 8


def test_eos_finishes_with_stop():
    st = FakeStepper(eos_id=EOS, eos_every=4)
    e = Engine(st)
    try:
        p = list(range(60, 100))
        want = expected_tokens(p, 64, EOS, eos_every=4)
        assert want[-1] == EOS
        c = run(e, p, n=64)
        assert c.finish == "stop" and c.tokens == want[:-1]      # the EOS itself is not streamed
        assert c.usage["completion_tokens"] == len(want)
    finally:
        e.shutdown()


def test_lru_eviction_when_pool_is_full():
    st = FakeStepper(n_blocks=8, max_seqs=2, max_tokens=256, max_pos=512, eos_id=EOS)
    e = Engine(st)
    try:
        pa, pb, pc = [50] * 150, [60] * 150, [70] * 150          # 3 blocks each (+ generation) -> only two fit
        run(e, pa, n=4, hint="a")
        run(e, pb, n=4, hint="b")
        run(e, pc, n=4, hint="c")                                # evicts "a" (least recently used)
        assert e.sched.stats["evictions"] >= 1
        hints = {x["thread_id"] for x in e.sched.residency()}
        assert "a" not in hints and {"b", "c"} <= hints
        c = run(e, pa + [1, 2], n=4, hint="a")                   # evicted thread: full re-prefill, same tokens
        assert c.usage["cached_tokens"] == 0 and c.tokens == expected_tokens(pa + [1, 2], 4, EOS)
    finally:
        e.shutdown()


def test_context_overflow_message_triggers_reference_compaction(engine):
    with pytest.raises(ContextLengthError) as ei:
        engine.submit(list(range(2000)), SamplingParams(), lambda ev: None)
    msg = str(ei.value).lower()
    assert "maximum context length" in msg and "context_length_exceeded" in msg


def test_max_tokens_clamped_to_context(engine):
    c = run(engine, [65] * 1000, n=500)
    assert c.usage["completion_tokens"] == 24                    # 1024 - 1000


def test_cancellation_frees_slot_and_keeps_kv(engine):
    got = []
    gate = threading.Event()

    def on_event(ev):
        got.append(ev)
        if len(got) == 3:
            gate.set()

    req = engine.submit([80] * 100, SamplingParams(max_tokens=500), on_event, thread_hint="c")
    assert gate.wait(10)
    req.cancelled = True
    deadline = time.time() + 10
    while engine.sched.running and time.time() < deadline:
        time.sleep(0.01)
    assert not engine.sched.running
    res = {x["thread_id"]: x for x in engine.sched.residency()}
    assert res["c"]["n_tokens"] >= 100                           # prompt + what was generated stays resident
    c = run(engine, [80] * 100 + [1], n=2, hint="c")
    assert c.usage["cached_tokens"] >= 100


def test_engine_failure_reaches_every_request():
    class Boom(FakeStepper):
        def decode_launch(self, reqs, parity):
            raise RuntimeError("device fault")
    e = Engine(Boom(eos_id=EOS))
    try:
        c = run(e, [70] * 10, n=5)
        assert isinstance(c.error, RuntimeError)
        with pytest.raises(RuntimeError, match="engine thread died"):
            e.submit([1, 2], SamplingParams(), lambda ev: None)
    finally:
        e.shutdown()


def test_sharded_engine_routes_by_thread():
    engines = [Engine(FakeStepper(eos_id=EOS)) for _ in range(4)]
    se = ShardedEngine(engines)
    try:
        for i in range(16):
            tid = f"bench-{i:04d}"
            c = run(se, [40 + i] * 70, n=3, hint=tid)
            assert c.tokens == expected_tokens([40 + i] * 70, 3, EOS)
        for i in range(16):
            tid = f"bench-{i:04d}"
            owner = shard_of(tid, [], 4)
            hints = {x["thread_id"] for x in engines[owner].sched.residency()}
            assert tid in hints
            for g in range(4):
                if g != owner:
                    assert tid not in {x["thread_id"] for x in engines[g].sched.residency()}
    finally:
        se.shutdown()


# ---- shared-prefix KV -----------------------------------------------------
def test_common_system_prompt_is_prefilled_once(engine):
    """The reference prepends one byte-identical system prompt to every thread: its
    full blocks are computed by the first thread and only read by the others."""
    st = engine.stepper
    system = [300 + (i % 50) for i in range(200)]                 # 3 full blocks + 8 tokens
    first = run(engine, system + [65, 66, 67], n=4, hint="t0")
    assert first.usage["cached_tokens"] == 0
    for i in range(1, 6):
        p = system + [70 + i] * (5 + i)
        before = st.prefilled_tokens
        c = run(engine, p, n=6, hint=f"t{i}")
        assert c.tokens == expected_tokens(p, 6, EOS)             # read back through the shared blocks
        assert c.usage["cached_tokens"] == 192                    # 3 shared blocks, nothing recomputed
        assert st.prefilled_tokens - before == len(p) - 192
    assert engine.sched.stats["shared_tokens"] == 5 * 192
    # every thread can still continue its own conversation in place
    c = run(engine, system + [65, 66, 67] + first.tokens + [EOS, 90, 91], n=3, hint="t0")
    assert c.usage["cached_tokens"] == 203 + 3 and c.tokens == expected_tokens(
        system + [65, 66, 67] + first.tokens + [EOS, 90, 91], 3, EOS)


def test_prefix_shared_from_a_conversation_that_is_still_decoding():
    st = FakeStepper(n_blocks=64, max_seqs=4, max_tokens=512, max_pos=2048, eos_id=EOS)
    e = Engine(st)
    try:
        system = [200 + (i % 31) for i in range(130)]             # 2 full blocks
        slow, fast = Collector(), Collector()
        e.submit(system + [65], SamplingParams(max_tokens=300), slow, thread_hint="slow")
        while len(slow.tokens) < 3:                                # prefill done, decoding
            time.sleep(0.005)
        e.submit(system + [66, 67], SamplingParams(max_tokens=5), fast, thread_hint="fast")
        fast.wait()
        assert fast.usage["cached_tokens"] == 128 and fast.tokens == expected_tokens(system + [66, 67], 5, EOS)
        slow.wait(30)
        assert slow.tokens == expected_tokens(system + [65], 300, EOS)   # the donor was not disturbed
    finally:
        e.shutdown()


def test_copy_on_write_when_a_shared_block_would_be_overwritten(engine):
    a_prompt = [400 + (i % 17) for i in range(200)]
    run(engine, a_prompt, n=4, hint="A")
    b_prompt = a_prompt[:192] + [77] * 20                         # shares A's first 3 blocks
    b = run(engine, b_prompt, n=4, hint="B")
    assert b.usage["cached_tokens"] == 192
    # A's history is edited at position 100, i.e. inside shared block 1: A must not write there
    a2 = a_prompt[:100] + [88] * 60
    c = run(engine, a2, n=5, hint="A")
    assert c.usage["cached_tokens"] == 64                         # resumed from the block boundary
    assert c.tokens == expected_tokens(a2, 5, EOS)
    # B still reads intact blocks
    b2 = b_prompt + b.tokens + [EOS, 65]
    d = run(engine, b2, n=4, hint="B")
    assert d.tokens == expected_tokens(b2, 4, EOS) and d.usage["cached_tokens"] == len(b_prompt) + 3


def test_reference_counts_return_every_block(engine):
    system = [250 + (i % 13) for i in range(150)]
    for i in range(5):
        run(engine, system + [65 + i] * 10, n=3, hint=f"r{i}")
    sched = engine.sched
    assert max(sched.alloc.refs) >= 2                             # something is shared
    for e in list(sched.index.entries.values()):                  # evict everything
        sched.index.remove(e)
        sched.alloc.free(e.blocks)
    assert sched.alloc.n_free == sched.alloc.n_blocks and set(sched.alloc.refs) == {0}


# ---- two-stage pipeline: step k+1 is enqueued before step k's ids are read back ---------------------
class _CountingStepper(FakeStepper):
    def __init__(self, **kw):
        super().__init__(**kw)
        self.max_outstanding = 0
        self.discarded = 0

    def decode_launch(self, reqs, parity):
        h = super().decode_launch(reqs, parity)
        self.max_outstanding = max(self.max_outstanding, self.outstanding)
        return h


@pytest.mark.parametrize("pipeline", [True, False])
def test_pipelined_loop_gives_the_same_tokens(pipeline):
    """Staggered arrivals, early EOS, max_tokens=1 and a follow-up turn on resident KV: ids must not
    depend on whether results are read back before or after the next step is enqueued."""
    st = _CountingStepper(n_blocks=128, max_seqs=8, max_tokens=256, max_pos=2048, eos_id=EOS, eos_every=9)
    e = Engine(st, max_prefill_tokens=96, pipeline=pipeline)
    try:
        prompts = [list(range(33 + 3 * i, 33 + 3 * i + 40 + 29 * i)) for i in range(12)]
        lens = [1, 2, 3, 17, 40, 5, 64, 9, 30, 2, 1, 25]
        cols = [Collector() for _ in prompts]
        for i, (p, c, n) in enumerate(zip(prompts, cols, lens)):
            e.submit(p, SamplingParams(max_tokens=n), c, thread_hint=f"t{i}")
            if i % 3 == 2:
                time.sleep(0.002)                                 # arrive while others are mid-decode
        outs = []
        for p, c, n in zip(prompts, cols, lens):
            c.wait()
            want = expected_tokens(p, n, EOS, eos_every=9)
            streamed = want[:-1] if want[-1] == EOS else want
            assert c.tokens == streamed and c.finish == ("stop" if want[-1] == EOS else "length")
            assert c.usage["completion_tokens"] == len(want)
            outs.append(want)
        # second turn of every thread continues on the KV the first one left behind
        for i, (p, want) in enumerate(zip(prompts, outs)):
            p2 = p + [t for t in want if t != EOS] + [35, 36, 37]
            c = run(e, p2, n=6, hint=f"t{i}")
            w2 = expected_tokens(p2, 6, EOS, eos_every=9)
            assert c.tokens == (w2[:-1] if w2[-1] == EOS else w2)
            assert c.usage["cached_tokens"] >= len(p) - 1
        assert st.max_outstanding >= (2 if pipeline else 1)
        if not pipeline:
            assert st.max_outstanding <= 2                        # at most one iteration (prefill + decode)
    finally:
        e.shutdown()


def test_cancel_with_a_step_in_flight_keeps_kv_consistent():
    st = FakeStepper(n_blocks=32, max_seqs=4, max_tokens=256, max_pos=1024, eos_id=EOS)
    e = Engine(st)
    try:
        p = list(range(50, 150))
        c = Collector()
        got = threading.Event()
        orig = c.__call__

        def on_event(ev):
            orig(ev)
            if len(c.tokens) == 5:
                got.set()
        r = e.submit(p, SamplingParams(max_tokens=200), on_event, thread_hint="c")
        assert got.wait(10)
        r.cancelled = True                                        # a step is in flight right now
        deadline = time.time() + 10
        while r.entry is not None and time.time() < deadline:
            time.sleep(0.001)
        assert r.entry is None and r.done
        n = len(c.tokens)
        p2 = p + c.tokens + [77]
        c2 = run(e, p2, n=8, hint="c")                            # reuses what the cancelled turn left
        assert c2.tokens == expected_tokens(p2, 8, EOS)
        assert c2.usage["cached_tokens"] >= len(p) + n - 2
    finally:
        e.shutdown()


def test_incremental_chain_hashes_equal_a_full_rehash(engine):
    """Chains are extended from what admission already hashed (the thread's resident entry, then the
    prompt's own chain): after turns that continue, diverge from and share conversations, every entry's
    chain must still be the chain of its tokens, and the index must hold exactly those hashes."""
    from kafka_b200.scheduler import _chain_hashes
    sys_prompt = list(range(40, 40 + 2 * BLOCK_TOKENS))
    convs = {}
    for turn in range(3):
        for t in range(3):
            hint = f"th-{t}"
            prompt = convs.get(hint, sys_prompt + [100 + t] * (20 + 30 * t))
            if turn == 2 and t == 1:
                prompt = prompt[:BLOCK_TOKENS + 9] + [77] * 70          # history edited: diverges inside block 1
            c = run(engine, prompt, n=37 + 11 * t, hint=hint)
            assert c.tokens == expected_tokens(prompt, 37 + 11 * t, EOS)
            convs[hint] = prompt + c.tokens + [200 + turn] * (5 + 64 * turn)
    time.sleep(0.05)
    idx = engine.sched.index
    assert len(idx.entries) >= 3
    registered = {}
    for e in idx.entries.values():
        assert e.chain == _chain_hashes(e.tokens), e.thread_hint
        for h in e.chain:
            registered.setdefault(h, set()).add(e.entry_id)
    assert {h: set(d) for h, d in idx.by_hash.items()} == registered


def test_engine_raises_gc_thresholds_unless_opted_out(monkeypatch):
    """engine.tune_gc: the cyclic collector runs ~70x less often under a serving loop (host time of a
    512-thread turn, tools/host_profile.py); KB_GC_DEFAULT=1 leaves the interpreter's setting alone."""
    import gc
    from kafka_b200 import engine as E
    old = gc.get_threshold()
    try:
        gc.set_threshold(700, 10, 10)
        monkeypatch.setenv("KB_GC_DEFAULT", "1")
        E.tune_gc()
        assert gc.get_threshold() == (700, 10, 10)
        monkeypatch.delenv("KB_GC_DEFAULT")
        E.tune_gc()
        assert gc.get_threshold()[0] >= 50_000
        gc.set_threshold(200_000, 5, 5)                     # a host that already tuned it higher keeps its value
        E.tune_gc()
        assert gc.get_threshold() == (200_000, 5, 5)
    finally:
        gc.set_threshold(*old)


def test_request_larger_than_the_pool_fails_at_once_and_engine_does_not_spin():
    """A follow-up of a resident thread that can never fit (8 blocks, pool of 4) must get an error event,
    not wait forever; and while a request waits with nothing in flight the engine thread sleeps on its inbox."""
    st = FakeStepper(n_blocks=4, max_seqs=4, max_tokens=256, max_pos=1024, eos_id=EOS)
    e = Engine(st, max_prefill_tokens=128)
    try:
        first = run(e, list(range(100, 160)), n=4, hint="t")          # 1 block, stays resident
        assert first.error is None and first.finish == "length"
        big = run(e, list(range(100, 160)) + first.tokens + [7] * 420, n=16, hint="t")   # 8 blocks > 4
        assert isinstance(big.error, MemoryError) and "pool has 4" in str(big.error)
        ok = run(e, list(range(100, 160)) + first.tokens + [9] * 20, n=4, hint="t")       # engine still serves
        assert ok.error is None and ok.tokens == expected_tokens(list(range(100, 160)) + first.tokens + [9] * 20, 4, EOS)
        assert e.stats()["waiting"] == 0
    finally:
        e.shutdown()


def test_reuse_is_given_up_when_the_kept_entry_blocks_admission():
    """Pool of 4: an idle conversation u holds 3 blocks. A new thread that shares u's first block and needs 3
    blocks would read that block from u (kept as donor) and then lacks one block that nobody will ever free —
    the scheduler must drop the sharing, evict u and serve the request as a new conversation."""
    st = FakeStepper(n_blocks=4, max_seqs=4, max_tokens=512, max_pos=1024, eos_id=EOS)
    e = Engine(st, max_prefill_tokens=512)
    try:
        common = list(range(300, 364))                                # exactly one block
        b = run(e, common + list(range(400, 500)), n=4, hint="u")     # 168 tokens -> 3 blocks
        assert b.error is None and e.stats()["free_blocks"] == 1
        prompt = common + [5] * 100                                   # 164 + 8 -> 3 blocks
        c = run(e, prompt, n=8, hint="t")
        assert c.error is None and c.tokens == expected_tokens(prompt, 8, EOS)
        assert e.stats()["evictions"] == 1 and e.stats()["waiting"] == 0
    finally:
        e.shutdown()


def test_donor_compares_tokens_not_only_the_chain_hash():
    """Two different conversations forced onto the same chain hash (a simulated 64-bit collision): the index
    must not offer one's blocks to the other."""
    idx = PrefixIndex()
    a = KVEntry(1, tokens=list(range(128)), blocks=[0, 1])
    idx.add(a)
    other = [7] * 128
    forged = list(a.chain)                      # pretend `other` hashes to the same chain
    e, n = idx.donor(forged, prompt=other)
    assert e is None and n == 0
    e, n = idx.donor(forged, prompt=list(range(128)) + [1, 2, 3])
    assert e is a and n == 2


@pytest.mark.parametrize("prefill_first,streak", [(True, 16), (True, 2), (False, 16)])
def test_prompts_go_before_tokens_but_cannot_starve_them(prefill_first, streak):
    """Engine(prefill_first=...): while admitted prompts still have chunks to prefill, iterations launch the chunk
    alone; after `max_prefill_streak` such iterations a decode step is forced. Whatever the policy, every request's
    tokens are the same."""
    log = []

    class Logged(FakeStepper):
        def prefill_launch(self, items, parity):
            log.append("P")
            return super().prefill_launch(items, parity)

        def decode_launch(self, reqs, parity):
            log.append("D")
            return super().decode_launch(reqs, parity)

    st = Logged(n_blocks=256, max_seqs=16, max_tokens=256, max_pos=1024, eos_id=EOS)
    e = Engine(st, max_prefill_tokens=64, prefill_first=prefill_first, max_prefill_streak=streak)
    try:
        rnd = random.Random(5)
        prompts = [[rnd.randrange(32, 127) for _ in range(100)] for _ in range(8)]     # 800 tokens = 13 chunks of 64
        gate = threading.Event()
        orig = st.prefill_launch

        def held(items, parity):                      # all eight are queued before the first chunk is planned
            gate.wait(5)
            return orig(items, parity)
        st.prefill_launch = held
        cols = [Collector() for _ in prompts]
        for p, c in zip(prompts, cols):
            e.submit(p, SamplingParams(max_tokens=6), c)
        time.sleep(0.2)
        gate.set()
        outs = [c.wait() for c in cols]
        for p, o in zip(prompts, outs):
            assert o.tokens == expected_tokens(p, 6, EOS)
        first_d = log.index("D")
        n_p = log.count("P")
        assert n_p >= 12
        if not prefill_first:
            assert first_d <= 2                       # a decode step rides along from the second iteration on
        elif streak >= n_p:
            assert first_d >= n_p - 1                 # every chunk but possibly the last went before any decode step
        else:
            runs = [len(r) for r in "".join(log[:log.index("D", first_d)]).split("D")]
            assert first_d == streak + 1 and max(runs) <= streak + 1     # forced decode after `streak` prompt-only iterations
    finally:
        e.shutdown()


@pytest.mark.parametrize("seed", list(range(1, 13)))
def test_random_traffic_soak_keeps_tokens_and_block_accounting_exact(seed):
    """A few hundred random operations on a pool far too small for them: threads with a common system prompt take
    multi-turn conversations, some edit their history, some cancel mid-generation, some stop on EOS, requests
    arrive in bursts larger than the batch; blocks are shared, copied on write, evicted and recycled. Every
    request that finishes must produce exactly the tokens that are a pure function of its prompt (FakeStepper reads
    them back through the block tables, so any stale, lost or wrongly shared block changes them), and in the end
    every block is either free or held by a resident entry, with the reference counts matching."""
    rnd = random.Random(seed)
    st = FakeStepper(n_blocks=40, max_seqs=6, max_tokens=192, max_pos=1024, eos_id=EOS, eos_every=23)
    e = Engine(st, max_prefill_tokens=96, prefill_first=bool(seed % 2), max_prefill_streak=3)
    try:
        system = [300 + (i % 29) for i in range(rnd.choice([70, 130, 200]))]
        convs = {f"th{i}": list(system) + [65 + i] * rnd.randrange(3, 40) for i in range(10)}
        finished = 0
        for burst in range(30):
            batch = []
            for tid in rnd.sample(sorted(convs), rnd.randrange(1, 9)):
                prompt = convs[tid]
                if rnd.random() < 0.15 and len(prompt) > len(system) + 10:      # the client edited its history
                    cut = rnd.randrange(len(system) // 2, len(prompt) - 1)
                    prompt = prompt[:cut] + [rnd.randrange(32, 127) for _ in range(rnd.randrange(1, 30))]
                if len(prompt) > 700:                                           # a fresh conversation on this thread
                    prompt = list(system) + [rnd.randrange(32, 127) for _ in range(5)]
                convs[tid] = prompt
                c = Collector()
                n = rnd.randrange(1, 24)
                req = e.submit(prompt, SamplingParams(max_tokens=n), c, thread_hint=tid)
                batch.append((tid, prompt, n, c, req, rnd.random() < 0.2))
            for tid, prompt, n, c, req, cancel in batch:
                if cancel:
                    time.sleep(rnd.random() * 0.003)
                    req.cancelled = True
            for tid, prompt, n, c, req, cancel in batch:
                # a cancelled request is dropped without a final event (its consumer is gone): wait for the
                # scheduler to let go of it; everything else must finish
                t_end = time.time() + 20
                while not c.done.is_set() and not (cancel and req.done):
                    assert time.time() < t_end, ("request neither finished nor was released", tid, burst)
                    time.sleep(0.001)
                assert c.error is None, c.error
                want = expected_tokens(prompt, n, EOS, eos_every=23)
                if want and want[-1] == EOS:
                    want = want[:-1]                                            # the EOS itself is not streamed
                if cancel and not c.done.is_set():
                    assert c.tokens == want[:len(c.tokens)]                     # whatever it got before the cancel
                else:
                    assert c.tokens == want, (tid, burst)
                    finished += 1
                # the thread continues with what it saw (like a chat client), plus a new user message
                seen = list(c.tokens)
                convs[tid] = prompt + seen + [EOS] + [rnd.randrange(32, 127) for _ in range(rnd.randrange(1, 25))]
        assert finished > 60
        sched = e.sched
        for _ in range(200):                                                    # let the loop retire the last flight
            if sched.idle():
                break
            time.sleep(0.005)
        assert sched.idle()
        held = {}
        for ent in sched.index.entries.values():
            assert not ent.busy
            for b in ent.blocks:
                held[b] = held.get(b, 0) + 1
        assert sched.alloc.n_free + len(held) == sched.alloc.n_blocks
        assert all(sched.alloc.refs[b] == n for b, n in held.items())
        assert sum(1 for r in sched.alloc.refs if r == 0) == sched.alloc.n_free
    finally:
        e.shutdown()
