"""Serving a tensor-parallel model: one process per GPU, the scheduler on rank 0 only.

    rank 0      Engine (scheduler, continuous batching, pipelined loop) -> GpuStepper; every launch plan
                (a plain dict of token ids, positions, block-table updates, sampling arrays) is published
                to the other ranks before rank 0 enqueues it on its own GPU
    rank 1..n   follow(): receive a plan, enqueue the same call on the local shard. No scheduler, no
                read-back: every rank ends a step with the same sampled ids on its device (the lm_head is cut
                by vocabulary and the ranks' winners are compared through the peer region, or replicated with
                NCCL), so the next plan can refer to them by slot (kb_select_tokens).

The plans travel over a CPU (gloo) process group — a few hundred bytes per decode step of a stable batch —
and the GPUs meet in the collectives of the forward pass (csrc/kb_collective.cu), which also keeps the
ranks within one kernel of each other. The provider, the HTTP server and the thread store run on rank 0
exactly as on one GPU.

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 -m kafka_b200.tp_serve --model llama-3-70b --port 8081

Reference counterpart: none (the reference's models sit behind a gateway, /root/reference/src/llm/portkey.py:379).
"""
from __future__ import annotations

import argparse
import os
import pickle
from typing import Optional

import torch


class PlanChannel:
    """Leader -> followers broadcast of small Python objects over a gloo group (length, then payload)."""

    def __init__(self, group, src: int = 0):
        import torch.distributed as dist
        self.dist, self.group, self.src = dist, group, src
        self._len = torch.zeros(1, dtype=torch.int64)

    def send(self, obj) -> None:
        data = pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL)
        self._len[0] = len(data)
        self.dist.broadcast(self._len, self.src, group=self.group)
        buf = torch.frombuffer(bytearray(data), dtype=torch.uint8)
        self.dist.broadcast(buf, self.src, group=self.group)

    def recv(self):
        self.dist.broadcast(self._len, self.src, group=self.group)
        buf = torch.empty(int(self._len[0]), dtype=torch.uint8)
        self.dist.broadcast(buf, self.src, group=self.group)
        return pickle.loads(buf.numpy().tobytes())


def build_tp_stepper(cfg, device, n_blocks: int, max_tokens: int, max_seqs: int, weights=None, seed: int = 0,
                     allreduce: Optional[str] = None, generable_only: bool = True):
    """This rank's shard + stepper + plan channel. Collective call on all ranks of the default group."""
    import torch.distributed as dist
    from .gpu_stepper import GpuStepper
    from .tp import build_tp_core
    rank, world = dist.get_rank(), dist.get_world_size()
    core, grp = build_tp_core(cfg, rank, world, device, n_blocks=n_blocks, max_tokens=max_tokens, max_seqs=max_seqs,
                              weights=weights, seed=seed, allreduce=allreduce)
    from .factory import tokenizer_for
    stepper = GpuStepper(core, tokenizer_for(cfg), generable_only=generable_only)
    stepper.tp_group = grp                      # keeps the symmetric region / callback alive
    chan = PlanChannel(dist.new_group(backend="gloo"))
    return stepper, chan


def lead(stepper, chan, max_prefill_tokens: Optional[int] = None):
    """Rank 0: the Engine whose launches are mirrored on the followers. Engine.shutdown() releases them."""
    from .engine import Engine
    stepper.publish = chan.send
    inner_close = stepper.close

    def close():
        chan.send({"op": "stop"})
        inner_close()
    stepper.close = close
    return Engine(stepper, max_prefill_tokens=max_prefill_tokens, name="kb-engine-tp-leader")


def follow(stepper, chan) -> int:
    """Ranks > 0: run the leader's plans on the local shard until it stops. Returns the number of plans."""
    n = 0
    while True:
        p = chan.recv()
        if p.get("op") == "stop":
            break
        stepper.run_plan(p, read_back=False)
        n += 1
    stepper.close()       # EngineCore.close() synchronises the device before releasing the shard
    return n


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama-3-70b")
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--port", type=int, default=8081)
    ap.add_argument("--db", default="threads.db")
    ap.add_argument("--max-seqs", type=int, default=int(os.environ.get("KAFKA_B200_MAX_SEQS", "256")))
    ap.add_argument("--max-tokens", type=int, default=int(os.environ.get("KAFKA_B200_MAX_TOKENS", "4096")))
    ap.add_argument("--kv-fraction", type=float, default=float(os.environ.get("KAFKA_B200_KV_FRACTION", "0.85")))
    args = ap.parse_args()

    import torch.distributed as dist
    from .engine_core import EngineCore
    from .model_config import get_config
    from .tp import local_config
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
    from . import checkpoint
    checkpoint.register_from_env()              # --model may name a checkpoint of $KAFKA_B200_CHECKPOINTS
    cfg = get_config(args.model)
    loc = local_config(cfg, dist.get_world_size())
    free, _ = torch.cuda.mem_get_info(dev)
    w_bytes = 2 * (loc.matmul_params() + 2 * cfg.vocab * cfg.hidden)
    fit = torch.tensor([int((free - w_bytes) * args.kv_fraction) // EngineCore.kv_block_bytes(loc)], device=dev)
    dist.all_reduce(fit, op=dist.ReduceOp.MIN)          # every rank must own the same number of blocks
    stepper, chan = build_tp_stepper(cfg, dev, n_blocks=max(8, int(fit.item())), max_tokens=args.max_tokens,
                                     max_seqs=args.max_seqs)
    try:
        if dist.get_rank() != 0:
            follow(stepper, chan)
            return
        import uvicorn
        from .db import ThreadStore
        from .provider import B200LLMProvider
        from .server import create_app
        from .factory import tokenizer_for
        engine = lead(stepper, chan)
        tok = tokenizer_for(cfg)

        def provider_factory(model, tool_provider=None, thread_id=None, **kw):
            return B200LLMProvider(engine, tok, model=cfg.name, tool_provider=tool_provider, thread_id=thread_id, **kw)
        try:
            uvicorn.run(create_app(provider_factory, ThreadStore(args.db), models=[cfg.name]), host=args.host,
                        port=args.port, log_level="info")
        finally:
            engine.shutdown()
    finally:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
