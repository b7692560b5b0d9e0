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
from kafka_b200.tp import local_config, shard_weights
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
    finally:
        dist.destroy_process_group()


def test_two_rank_tensor_parallel_equals_full_model():
    mgr = mp.Manager()
    ret = mgr.dict()
    port = 29500 + os.getpid() % 400
    mp.spawn(_worker, args=(2, port, ret), nprocs=2, join=True)
    assert ret["err"] < 1e-4, ret["err"]          # fp32 sums in a different order
    assert ret["kv_heads_local"] == 1 and ret["ranks_agree"]


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
    f = fuse_for_engine(local_config(cfg, 2), parts[1], "cpu")
    assert f["l0.wqkv"].shape == (2 * 64 + 2 * 64, 256) and f["l0.w_gate_up"].shape == (512, 256)
    assert f["l0.wo"].shape == (256, 128) and f["l0.w_down"].shape == (256, 256)
