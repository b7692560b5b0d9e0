#!/usr/bin/env python
"""Can the HBM-bound decode attention and the tensor-bound projections share the GPU?

The decode step runs them one after the other: attention streams 4.5 GB per layer at the HBM ceiling while the tensor
pipes idle, then the GEMMs compute while HBM idles (~4 ms of a 28 ms step). Running them side by side (two halves of
the batch, one layer apart) only pays if attention keeps its bandwidth on FEWER SMs. This probe measures exactly that:
decode attention (512 sequences x 2 112 tokens, Llama-3-8B heads) alone, and while a persistent GEMM occupies k SMs
(one 192 KB CTA per SM, so no attention CTA fits beside it) on a second stream.

    python tools/overlap_probe.py [--blocked 0 32 48 64]
"""
import argparse
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "kafka-llm-service_b200"))

import torch  # noqa: E402

from kafka_b200 import cabi  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--blocked", type=int, nargs="+", default=[0, 32, 48, 64])
    ap.add_argument("--batch", type=int, default=512)
    ap.add_argument("--ctx", type=int, default=2112)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    hd, n_heads, n_kv = 128, 32, 8
    B, ctx = args.batch, args.ctx
    mb = (ctx + 63) // 64 + 1
    g = torch.Generator(device="cpu").manual_seed(0)
    kp = torch.randn(B * mb, n_kv, 64 * hd, device=dev).half()
    vp = torch.randn(B * mb, n_kv, 64 * hd, device=dev).half()
    bt = torch.randperm(B * mb, generator=g).int().view(B, mb).to(dev)
    q = torch.randn(B, (n_heads + 2 * n_kv) * hd, device=dev).half()
    lens = torch.full((B,), ctx, dtype=torch.int32, device=dev)
    attn_bytes = B * ctx * n_kv * hd * 2 * 2
    s_attn, s_gemm = torch.cuda.Stream(), torch.cuda.Stream()
    out = []
    for k in args.blocked:
        gemm_ms = None
        if k > 0:
            M, K = 128, 32768
            a = torch.randn(M, K, device=dev).half()
            w = (torch.randn(256 * k, K, device=dev) * 0.02).half()      # k tiles of 128 x 256 -> k persistent CTAs
            o = torch.empty(M, 256 * k, dtype=torch.float16, device=dev)
            with torch.cuda.stream(s_gemm):
                cabi.gemm(a, w, cabi.EPI_F16, out=o, tile_n=256)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(4):
                    cabi.gemm(a, w, cabi.EPI_F16, out=o, tile_n=256)
                e1.record()
            torch.cuda.synchronize()
            gemm_alone = e0.elapsed_time(e1) / 4
        with torch.cuda.stream(s_attn):
            for _ in range(3):
                cabi.attn_decode(q, kp, vp, lens, bt, n_heads, n_kv, hd, max_kv_len=ctx, out_f16=True)
        torch.cuda.synchronize()
        n_attn = 12
        if k > 0:
            n_gemm = max(4, int(n_attn * 0.7 / max(gemm_alone, 1e-3)) + 4)      # keep the SMs blocked throughout
            with torch.cuda.stream(s_gemm):
                g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                g0.record()
                for _ in range(n_gemm):
                    cabi.gemm(a, w, cabi.EPI_F16, out=o, tile_n=256)
                g1.record()
        with torch.cuda.stream(s_attn):
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record()
            for _ in range(n_attn):
                cabi.attn_decode(q, kp, vp, lens, bt, n_heads, n_kv, hd, max_kv_len=ctx, out_f16=True)
            a1.record()
        torch.cuda.synchronize()
        us = a0.elapsed_time(a1) / n_attn * 1e3
        row = {"sms_blocked": k, "attn_us": round(us, 1), "attn_GBps": round(attn_bytes / (us * 1e-6) / 1e9, 1)}
        if k > 0:
            gemm_ms = g0.elapsed_time(g1) / n_gemm
            flops = 2.0 * M * 256 * k * K
            row.update(gemm_alone_ms=round(gemm_alone, 3), gemm_concurrent_ms=round(gemm_ms, 3),
                       gemm_alone_tflops=round(flops / gemm_alone / 1e9, 1),
                       gemm_concurrent_tflops=round(flops / gemm_ms / 1e9, 1))
            del a, w, o
        out.append(row)
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
