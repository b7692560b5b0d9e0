"""How much of a decode step is launch gap?  Times the same device-resident decode loop (kb_decode with
device-side token feedback) launched kernel by kernel on a stream and replayed from a CUDA graph that
captured two steps, for several thread counts.

    python tools/graph_probe.py [--model llama-3-8b] [--threads 1 8 64 128 512] [--history 2048]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "kafka-llm-service_b200"))

import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama-3-8b")
    ap.add_argument("--threads", type=int, nargs="*", default=[1, 8, 64, 128, 512])
    ap.add_argument("--history", type=int, default=2048)
    ap.add_argument("--steps", type=int, default=32)
    args = ap.parse_args()
    from kafka_b200 import cabi
    from kafka_b200.engine_core import EngineCore
    from kafka_b200.model_config import get_config
    from kafka_b200.weights import generate_fused_on_device
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    cfg = get_config(args.model)
    weights = generate_fused_on_device(cfg, seed=0, device=dev)
    rows = []
    for B in args.threads:
        bpt = (args.history + 6 * args.steps + 64 + 63) // 64
        eng = EngineCore(cfg, weights, dev, n_blocks=B * bpt, max_tokens=max(B, 64), max_seqs=B)
        for L in range(cfg.n_layers):
            eng.k_pool[L].normal_(0, 1)
            eng.v_pool[L].normal_(0, 1)
        g = torch.Generator(device="cpu").manual_seed(1)
        bt = torch.randperm(B * bpt, generator=g).int().view(B, bpt).to(dev)
        ctx = torch.full((B,), args.history, dtype=torch.int32, device=dev)
        sp = cabi.make_sampling(temperature=0.0, banned_token=cfg.vocab - 1)
        bufs = [torch.randint(32, 127, (B,), dtype=torch.int32, device=dev), torch.zeros(B, dtype=torch.int32, device=dev)]
        max_ctx = args.history + 6 * args.steps     # constant split choice for both variants
        stream = torch.cuda.Stream()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

        def two_steps():
            eng.decode(bufs[0], ctx, bt, sp, bufs[1], max_ctx=max_ctx, advance_ctx=True)
            eng.decode(bufs[1], ctx, bt, sp, bufs[0], max_ctx=max_ctx, advance_ctx=True)

        with torch.cuda.stream(stream):
            for _ in range(3):
                two_steps()
            stream.synchronize()
            e0.record()
            for _ in range(args.steps // 2):
                two_steps()
            e1.record()
            stream.synchronize()
            ms_stream = e0.elapsed_time(e1) / (args.steps // 2 * 2)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=stream):
                two_steps()
            for _ in range(2):
                graph.replay()
            stream.synchronize()
            e0.record()
            for _ in range(args.steps // 2):
                graph.replay()
            e1.record()
            stream.synchronize()
            ms_graph = e0.elapsed_time(e1) / (args.steps // 2 * 2)
        rows.append({"threads": B, "ms_per_step_stream": ms_stream, "ms_per_step_graph": ms_graph,
                     "gap_ms": ms_stream - ms_graph})
        eng.close()
        del eng
        torch.cuda.empty_cache()
    print(json.dumps({"model": cfg.name, "history": args.history, "rows": rows}))


if __name__ == "__main__":
    main()
