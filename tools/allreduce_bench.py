"""All-reduce micro-benchmark for the tensor-parallel path: our in-kernel collective over peer memory
(csrc/kb_collective.cu; multicast and plain peer-to-peer) beside NCCL, on the message the engine sends
(T x hidden fp32, in place), timed with CUDA events over back-to-back calls, max over ranks.

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 tools/allreduce_bench.py [--tokens 512] [--hidden 8192]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "kafka-llm-service_b200"))

import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tokens", type=int, nargs="*", default=[16, 128, 512, 2048])
    ap.add_argument("--hidden", type=int, default=8192)
    ap.add_argument("--iters", type=int, default=200)
    ap.add_argument("--blocks", type=int, nargs="*", default=[8, 16, 32, 64])
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    os.environ.setdefault("NCCL_DEBUG", "WARN")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    from kafka_b200.tp import PeerRegion

    n_max = max(args.tokens) * args.hidden
    out = {"world": world, "hidden": args.hidden, "iters": args.iters, "rows": []}
    regions = {}
    for name, mc in (("peers-multicast", True), ("peers-p2p", False)):
        try:
            reg = PeerRegion(4 * n_max, dev, use_multicast=mc)
            if mc and not reg.multicast:
                out.setdefault("notes", []).append("no multicast mapping on this box")
                continue
            regions[name] = reg
        except Exception as e:
            out.setdefault("notes", []).append(f"{name}: {e!r}")
    nccl_buf = torch.zeros(n_max, dtype=torch.float32, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def timed(fn):
        for _ in range(10):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        e0.record()
        for _ in range(args.iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / args.iters * 1e3], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for T in args.tokens:
        n = T * args.hidden
        row = {"tokens": T, "bytes": 4 * n}
        view = nccl_buf[:n]
        row["nccl_us"] = timed(lambda: dist.all_reduce(view))
        for name, reg in regions.items():
            for b in args.blocks:
                row[f"{name}_b{b}_us"] = timed(lambda: reg.allreduce_(0, n, b))
        ours = [(v, k) for k, v in row.items() if k.startswith("peers")]
        for name, reg in regions.items():                # the same T x hidden message as fp16 (half the bytes)
            row[f"f16_{name}_us"] = min(timed(lambda: reg.allreduce_(0, n, b, f16=True)) for b in args.blocks[-2:])
        # reduce-scatter + residual add + RMSNorm + all-gather in one kernel (the engine's default mode)
        if args.hidden <= 8192 and "peers-multicast" in regions:
            import ctypes as C
            from kafka_b200 import cabi
            reg = regions["peers-multicast"]
            xres = torch.zeros(n, dtype=torch.float32, device=dev)
            wn = torch.ones(args.hidden, dtype=torch.bfloat16, device=dev)
            xn_off = (2 * n + 255) & ~255

            def fused(b):
                cabi.check(cabi.load().kb_tp_reduce_norm_gather(C.byref(reg.peers), reg.rank, reg.size, 0, xn_off,
                                                                xres.data_ptr(), wn.data_ptr(), T, args.hidden, 1e-5,
                                                                reg.epoch, b, torch.cuda.current_stream().cuda_stream),
                           "fused")
                reg.epoch += 2
            for b in args.blocks:
                row[f"fused_b{b}_us"] = timed(lambda: fused(b))
        half = nccl_buf[:n // 2].view(torch.float16)
        row["f16_nccl_us"] = timed(lambda: dist.all_reduce(half))
        if ours:
            best = min(ours)
            row["best"] = best[1]
            row["best_busbw_GBps"] = 2 * (world - 1) / world * 4 * n / (best[0] * 1e-6) / 1e9
        row["nccl_busbw_GBps"] = 2 * (world - 1) / world * 4 * n / (row["nccl_us"] * 1e-6) / 1e9
        out["rows"].append(row)
    if rank == 0:
        print(json.dumps(out))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
