"""Tensor-parallel sharding on CPU: world_size-2 `gloo` processes, each holding one shard
(kafka_b200.tp.shard_weights / local_config), must reproduce the full model's logits when the two
row-parallel projections are all-reduced — the decomposition the GPU engine relies on
(kb_set_tensor_parallel). Also covers the thread -> replica routing used by the data-parallel path."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from kafka_b200.model_config import get_config
from kafka_b200.tp import lm_head_slice, lm_pairs_bytes, local_config, residual_region_bytes, shard_weights
from kafka_b200.weights import fuse_for_engine, generate_weights
from oracle.llama_fp32 import OracleLlama


def _worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(2)
        cfg = get_config("tiny-test")
        w = generate_weights(cfg, seed=4)
        loc = local_config(cfg, world)

        def allreduce(t):
            t = t.clone()
            dist.all_reduce(t)
            return t
        shard = OracleLlama(loc, shard_weights(cfg, w, rank, world), allreduce=allreduce, tp_rank=rank)
        g = torch.Generator().manual_seed(9)
        prompt = torch.randint(0, 256, (70,), generator=g)
        cache = shard.new_cache()
        logits = [shard.forward(prompt, cache)]
        for t in (65, 66, 67):
            logits.append(shard.forward(torch.tensor([t]), cache))
        got = torch.stack(logits)
        if rank == 0:
            full = OracleLlama(cfg, w)
            c = full.new_cache()
            want = torch.stack([full.forward(prompt, c)] + [full.forward(torch.tensor([t]), c) for t in (65, 66, 67)])
            ret["err"] = (got - want).abs().max().item()
            ret["kv_heads_local"] = cache.k(0).shape[1]
        # every rank ends with the same logits (replicated lm_head on an identical residual stream)
        gathered = [torch.empty_like(got) for _ in range(world)]
        dist.all_gather(gathered, got)
        if rank == 0:
            ret["ranks_agree"] = all(torch.equal(gathered[0], x) for x in gathered)
        # the vocabulary-sharded sampler's decomposition (kb_set_lm_head_shard): every rank multiplies only ITS rows
        # of the lm_head, keeps the winner of its slice, and the largest of the ranks' (value, id) pairs — ties to
        # the lowest id — is the full model's arg-max
        v0, nv = lm_head_slice(cfg, rank, world)
        cut = shard_weights(cfg, w, rank, world, shard_lm_head=True)["lm_head"].float()
        assert cut.shape == (nv, cfg.hidden) and torch.equal(cut, w["lm_head"][v0:v0 + nv].float())
        xl = torch.linalg.lstsq(w["lm_head"].float(), got.T).solution.T        # some hidden rows with these logits
        local = xl @ cut.T
        val, idx = local.max(-1)
        pair = torch.stack([val, (idx + v0).float()], -1)
        pairs = [torch.empty_like(pair) for _ in range(world)]
        dist.all_gather(pairs, pair)
        best_v, best_i = pairs[0][:, 0], pairs[0][:, 1]
        for p in pairs[1:]:
            better = (p[:, 0] > best_v) | ((p[:, 0] == best_v) & (p[:, 1] < best_i))
            best_v, best_i = torch.where(better, p[:, 0], best_v), torch.where(better, p[:, 1], best_i)
        if rank == 0:
            ret["sharded_argmax_ok"] = bool(torch.equal(best_i.long(), (xl @ w["lm_head"].float().T).argmax(-1)))
    finally:
        dist.destroy_process_group()


def test_two_rank_tensor_parallel_equals_full_model():
    mgr = mp.Manager()
    ret = mgr.dict()
    port = 29500 + os.getpid() % 400
    mp.spawn(_worker, args=(2, port, ret), nprocs=2, join=True)
    assert ret["err"] < 1e-4, ret["err"]          # fp32 sums in a different order
    assert ret["kv_heads_local"] == 1 and ret["ranks_agree"] and ret["sharded_argmax_ok"]


def test_shard_shapes_and_fusion():
    cfg = get_config("llama-3-70b")
    loc = local_config(cfg, 8)
    assert (loc.n_heads, loc.n_kv_heads, loc.ffn, loc.hidden, loc.vocab) == (8, 1, 3584, 8192, 128256)
    assert loc.kv_bytes_per_token == 40 * 1024                      # 40 KiB/token/GPU at TP 8
    with pytest.raises(ValueError):
        local_config(get_config("tiny-test"), 4)                     # 2 KV heads cannot split 4 ways
    cfg = get_config("tiny-test")
    w = generate_weights(cfg, seed=1)
    parts = [shard_weights(cfg, w, r, 2) for r in range(2)]
    assert torch.equal(torch.cat([p["l0.wq"] for p in parts]), w["l0.wq"])
    assert torch.equal(torch.cat([p["l1.w_down"] for p in parts], dim=1), w["l1.w_down"])
    # lm_head by vocabulary: the slices partition the rows; a vocabulary that does not cut evenly stays whole
    cut = [shard_weights(cfg, w, r, 2, shard_lm_head=True)["lm_head"] for r in range(2)]
    assert torch.equal(torch.cat(cut), w["lm_head"]) and parts[0]["lm_head"] is w["lm_head"]
    assert lm_head_slice(get_config("llama-3-70b"), 3, 8) == (3 * 16032, 16032)
    import dataclasses
    assert lm_head_slice(dataclasses.replace(cfg, vocab=520), 0, 2) is None and lm_head_slice(cfg, 0, 1) is None
    assert residual_region_bytes(cfg, 64, 16) == residual_region_bytes(cfg, 64) + lm_pairs_bytes(16) == \
        2 * 64 * cfg.hidden * 4 + 256 + 512
    f = fuse_for_engine(local_config(cfg, 2), parts[1], "cpu")
    assert f["l0.wqkv"].shape == (2 * 64 + 2 * 64, 256) and f["l0.w_gate_up"].shape == (512, 256)
    assert f["l0.wo"].shape == (256, 128) and f["l0.w_down"].shape == (256, 256)


# ---- serving a tensor-parallel model: the leader's launch plans mirrored on a follower (tp_serve.py) ----
def _serve_worker(rank, world, port, ret):
    import hashlib
    import pickle
    import sys
    sys.path.insert(0, os.path.dirname(__file__))
    from fakes import FakeStepper, expected_tokens
    from kafka_b200.scheduler import SamplingParams
    from kafka_b200.tp_serve import PlanChannel, follow, lead

    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        chan = PlanChannel(dist.new_group(backend="gloo"))

        class PlanFake(FakeStepper):
            """FakeStepper with the plan / run_plan split of GpuStepper: the plan carries what a shard
            needs (tokens, positions, block ids), run_plan executes it on this rank's own fake KV."""
            publish = None

            def __init__(self, **kw):
                super().__init__(**kw)
                self.digest = hashlib.blake2b(digest_size=8)
                self.n_plans = 0

            def _plan(self, op, parity, rows):
                return {"op": op, "parity": parity, "rows": rows}

            def prefill_launch(self, items, parity):
                p = self._plan("prefill", parity, [(list(it.req.entry.blocks), it.ctx_len, list(it.tokens),
                                                     it.is_last_chunk) for it in items])
                if self.publish:
                    self.publish(p)
                return self.run_plan(p)

            def decode_launch(self, reqs, parity):
                p = self._plan("decode", parity, [(list(r.entry.blocks), r.ctx_len,
                                                    r.out_slot if r.inflight > 0 else None,
                                                    None if r.inflight > 0 else r.generated[-1]) for r in reqs])
                if self.publish:
                    self.publish(p)
                return self.run_plan(p)

            def run_plan(self, p, read_back=True):
                self.n_plans += 1
                self.digest.update(pickle.dumps(p))
                slots = []
                if p["op"] == "prefill":
                    for i, (blocks, ctx, toks, last) in enumerate(p["rows"]):
                        for k, t in enumerate(toks):
                            self._write(blocks, ctx + k, t)
                        if last:
                            s = self.slot(p["parity"], True, i)
                            from fakes import next_token
                            self.res[s] = next_token(self._read(blocks, ctx + len(toks)), self.eos_id, self.eos_every)
                            slots.append(s)
                        else:
                            slots.append(None)
                else:
                    from fakes import next_token
                    fed = [self.res[src] if src is not None else tok for _, _, src, tok in p["rows"]]
                    for i, ((blocks, ctx, _, _), tok) in enumerate(zip(p["rows"], fed)):
                        self._write(blocks, ctx, tok)
                        s = self.slot(p["parity"], False, i)
                        self.res[s] = next_token(self._read(blocks, ctx + 1), self.eos_id, self.eos_every)
                        slots.append(s)
                return slots

            def fetch(self, handle):
                return [None if s is None else self.res[s] for s in handle]

        st = PlanFake(n_blocks=64, max_seqs=4, max_tokens=256, max_pos=1024, eos_id=511)
        if rank != 0:
            n = follow(st, chan)
            ret["follower_plans"], ret["follower_digest"] = n, st.digest.hexdigest()
            # the follower's own KV produced the same ids as the leader's (it never read any back)
            ret["follower_res"] = dict(st.res)
            return
        eng = lead(st, chan, max_prefill_tokens=96)
        import threading
        prompts = [list(range(40 + i, 40 + i + 50 + 37 * i)) for i in range(6)]
        done = [threading.Event() for _ in prompts]
        outs = [[] for _ in prompts]

        def cb(i):
            def on_event(ev):
                if ev[0] == "token":
                    outs[i].append(ev[1])
                elif ev[0] in ("finish", "error"):
                    done[i].set()
            return on_event
        for i, p in enumerate(prompts):
            eng.submit(p, SamplingParams(max_tokens=9 + i), cb(i), thread_hint=f"t{i}")
        for d in done:
            assert d.wait(20)
        ret["leader_ok"] = all(o == expected_tokens(p, 9 + i, 511) for i, (p, o) in enumerate(zip(prompts, outs)))
        eng.shutdown()
        ret["leader_plans"], ret["leader_digest"], ret["leader_res"] = st.n_plans, st.digest.hexdigest(), dict(st.res)
    finally:
        dist.destroy_process_group()


def test_leader_plans_are_mirrored_on_the_follower():
    mgr = mp.Manager()
    ret = mgr.dict()
    port = 29500 + (os.getpid() + 7) % 400
    mp.spawn(_serve_worker, args=(2, port, ret), nprocs=2, join=True)
    assert ret["leader_ok"]
    assert ret["leader_plans"] == ret["follower_plans"] > 10
    assert ret["leader_digest"] == ret["follower_digest"]      # same plans, same order
    assert ret["leader_res"] == ret["follower_res"]            # and the same ids on the follower's "device"
