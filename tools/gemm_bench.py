#!/usr/bin/env python
"""Micro-benchmark of kb_gemm on the decode/prefill shapes of Llama-3-8B (CUDA events, cold weights:
every launch reads a different weight buffer so that L2 does not hold it, as in the real step where
the attention kernel streams 4.5 GB between GEMMs).

    python tools/gemm_bench.py [--m 512] [--iters 20]
"""
import argparse
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "kafka-llm-service_b200"))

import torch  # noqa: E402

from kafka_b200 import cabi  # noqa: E402

SHAPES = {  # name: (N, K, epilogue)
    "qkv": (6144, 4096, cabi.EPI_F16),
    "o": (4096, 4096, cabi.EPI_RESID_F32),
    "gate_up": (28672, 4096, cabi.EPI_SWIGLU_F16),
    "down": (4096, 14336, cabi.EPI_RESID_F32),
    "lm_head": (128256, 4096, cabi.EPI_F32),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--m", type=int, nargs="+", default=[512])
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--tiles", type=int, nargs="+", default=[0, 64, 128, 256],
                    help="0 = heuristic; 64/128/256 = 1-CTA tile; negative multiple of 16 = CTA-pair tile width; "
                         "-(1024 * s + bn) = CTA-pair tile of width bn with split-K in s parts (o / down only)")
    ap.add_argument("--only", default=None)
    ap.add_argument("--ksweep", type=int, nargs="*", default=None,
                    help="override K with these values (time = fixed overhead + slope x K)")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    out = []
    for M in args.m:
        shapes = [(name, N, K, epi) for name, (N, K, epi) in SHAPES.items()]
        if args.ksweep:
            shapes = [(name, N, k, epi) for name, N, _, epi in shapes for k in args.ksweep]
        for name, N, K, epi in shapes:
            if args.only and name != args.only:
                continue
            nbuf = max(2, int(300e6 // (N * K * 2)) + 1)           # > 2 x L2 of distinct weights
            ws = [(torch.randn(N, K, device=dev) * 0.02).half() for _ in range(nbuf)]
            a = torch.randn(M, K, device=dev).half()
            resid = torch.zeros(M, N, device=dev) if epi == cabi.EPI_RESID_F32 else None
            for tile in args.tiles:
                if tile <= -1024 and epi != cabi.EPI_RESID_F32:
                    continue                                       # split-K exists for the in-place residual epilogue only
                # allocate the output once; the residual projections run in place, as in the engine
                o = resid if resid is not None else cabi.gemm(a, ws[0], epi, resid=resid, tile_n=tile)
                for i in range(3):
                    cabi.gemm(a, ws[i % nbuf], epi, resid=resid, out=o, tile_n=tile)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize()
                e0.record()
                for i in range(args.iters):
                    cabi.gemm(a, ws[i % nbuf], epi, resid=resid, out=o, tile_n=tile)
                e1.record()
                torch.cuda.synchronize()
                us = e0.elapsed_time(e1) / args.iters * 1e3
                tf = 2.0 * M * N * K / (us * 1e-6) / 1e12
                gbs = N * K * 2 / (us * 1e-6) / 1e9
                out.append({"M": M, "gemm": name, "N": N, "K": K, "tile_n": tile, "us": round(us, 2),
                            "tflops": round(tf, 1), "weight_GBps": round(gbs, 1)})
                print(json.dumps(out[-1]), flush=True)
            del ws
    return out


if __name__ == "__main__":
    main()
