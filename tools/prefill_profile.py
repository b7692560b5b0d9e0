"""Per-kernel time of the turn prefill (U new tokens per thread on a resident history) and of a cold
prefill (whole prompt), through kb_prefill with the engine's event profile.

    python tools/prefill_profile.py [--threads 512] [--history 2048] [--user-tokens 64] [--chunk 4096]
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
    ap.add_argument("--threads", type=int, default=512)
    ap.add_argument("--history", type=int, default=2048)
    ap.add_argument("--user-tokens", type=int, default=64)
    ap.add_argument("--chunk", type=int, default=4096)
    ap.add_argument("--only", default=None, choices=["turn", "cold"])
    ap.add_argument("--cold-seqs", type=int, default=32)
    args = ap.parse_args()
    from kafka_b200 import cabi
    from kafka_b200.engine_core import EngineCore
    from kafka_b200.model_config import get_config
    from kafka_b200.weights import generate_fused_on_device
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    cfg = get_config(args.model)
    weights = generate_fused_on_device(cfg, seed=0, device=dev)
    B, U, H = args.threads, args.user_tokens, args.history
    bpt = (H + U + 64 + 63) // 64
    eng = EngineCore(cfg, weights, dev, n_blocks=B * bpt, max_tokens=max(args.chunk, B), max_seqs=B)
    for L in range(cfg.n_layers):
        eng.k_pool[L].normal_(0, 1)
        eng.v_pool[L].normal_(0, 1)
    g = torch.Generator(device="cpu").manual_seed(1)
    bt = torch.randperm(B * bpt, generator=g).int().view(B, bpt).to(dev)
    sp = cabi.make_sampling(temperature=0.0, banned_token=cfg.vocab - 1)
    out = {}

    def run(name, n_seqs, new, ctx0):
        ctx = torch.full((n_seqs,), ctx0, dtype=torch.int32, device=dev)
        toks = torch.randint(0, 256, (n_seqs, new), generator=g, dtype=torch.int32).to(dev)
        first = torch.empty(n_seqs, dtype=torch.int32, device=dev)
        per = max(1, args.chunk // new)

        def once():
            for s0 in range(0, n_seqs, per):
                s1 = min(n_seqs, s0 + per)
                cu = torch.arange(0, (s1 - s0 + 1) * new, new, dtype=torch.int32, device=dev)
                eng.prefill(toks[s0:s1].reshape(-1), cu, ctx[s0:s1], bt[s0:s1], new, sp, None, first[s0:s1])
        once()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        once()
        e1.record()
        torch.cuda.synchronize()
        eng.profile_begin()
        once()
        prof = eng.profile_end()
        total = e0.elapsed_time(e1)
        flops = n_seqs * (new * 2 * cfg.matmul_params() + 2 * cfg.lm_head_params()
                          + 4 * cfg.head_dim * cfg.n_heads * cfg.n_layers * (ctx0 * new + new * new / 2))
        out[name] = {"seqs": n_seqs, "new_tokens": new, "ctx": ctx0, "ms": total, "tokens_per_s": n_seqs * new / total * 1e3,
                     "tflops": flops / total / 1e9,
                     "kernels_ms": {k: round(v[0], 3) for k, v in prof.items() if v[1]},
                     "launches": {k: v[1] for k, v in prof.items() if v[1]}}

    if args.only != "cold":
        run("turn", B, U, H)
    if args.only != "turn":
        run("cold", min(B, args.cold_seqs), H, 0)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
