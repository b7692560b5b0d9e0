"""Tensor parallelism on real GPUs (needs >= 2; skipped on the 1-GPU box): two ranks, each with the
local-shape engine of its shard and NCCL all-reduce through the kb_set_tensor_parallel callback, must
match the fp32 oracle of the FULL model within the logit tolerance (max-abs 2e-2), and sample
identical tokens on every rank.  Run with:  gpurun --gpus 2 -- python -m pytest tests/test_gpu_tp.py -m gpu
"""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, ret):
    import torch.distributed as dist
    from kafka_b200 import cabi
    from kafka_b200.model_config import get_config
    from kafka_b200.tp import build_tp_core
    from kafka_b200.weights import generate_weights
    from oracle.llama_fp32 import OracleLlama

    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        cfg = get_config("tiny-test")
        w = generate_weights(cfg, seed=0)
        core, grp = build_tp_core(cfg, rank, world, dev, n_blocks=16, max_tokens=512, max_seqs=4, weights=w)
        assert core.cfg.n_heads == cfg.n_heads // world and core.k_pool.shape[2] == cfg.n_kv_heads // world
        g = torch.Generator().manual_seed(1)
        lens = [5, 70, 129]
        prompts = [torch.randint(0, 256, (n,), generator=g) for n in lens]
        B, mb = len(lens), 4
        bt = torch.arange(B * mb, dtype=torch.int32).view(B, mb).to(dev)
        i32 = lambda x: torch.tensor(x, dtype=torch.int32, device=dev)
        logits = torch.empty(B, cfg.vocab, dtype=torch.float32, device=dev)
        out = torch.empty(B, dtype=torch.int32, device=dev)
        sp = cabi.make_sampling()
        core.prefill(torch.cat(prompts).int().to(dev), i32([0, 5, 75, 204]), i32([0, 0, 0]), bt, max(lens), sp, logits, out)
        torch.cuda.synchronize()
        oracle = OracleLlama(cfg, w)
        caches = [oracle.new_cache() for _ in prompts]
        ref = torch.stack([oracle.forward(p, c) for p, c in zip(prompts, caches)])
        errs = [(logits.cpu() - ref).abs().max().item()]
        toks = [out.cpu().clone()]
        ctx = i32(lens)
        cur = ref.argmax(-1)
        for step in range(4):
            core.decode(cur.int().to(dev), ctx, bt, sp, out, max_ctx=max(lens) + step, logits_out=logits,
                        advance_ctx=True)
            torch.cuda.synchronize()
            ref = oracle.decode_batch(cur, caches)
            errs.append((logits.cpu() - ref).abs().max().item())
            toks.append(out.cpu().clone())
            cur = ref.argmax(-1)
        mine = torch.stack(toks).to(dev)
        gathered = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(gathered, mine)
        if rank == 0:
            ret["max_err"] = max(errs)
            ret["ranks_agree"] = all(torch.equal(gathered[0], x) for x in gathered)
            ret["allreduce_calls"] = grp.calls
        core.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (gpurun --gpus 2)")
def test_two_gpu_tensor_parallel_matches_full_oracle():
    import torch.multiprocessing as mp
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(2, 29650 + os.getpid() % 300, ret), nprocs=2, join=True)
    assert ret["max_err"] <= 2e-2, ret["max_err"]
    assert ret["ranks_agree"]
    assert ret["allreduce_calls"] == 5 * 2 * 2        # 5 calls x 2 layers x 2 row-parallel projections
