"""Per-launch time of the small kernels of a decode step (RMSNorm, RoPE + KV append, sampling) at the bench's
shape (512 rows, Llama-3-8B), measured as the step runs them: back to back on one stream with programmatic
dependent launch, CUDA events around 200 launches. bench.py's per-kernel profile puts an event between every two
kernels, which adds the launch latency each of them hides in the real step.

    python tools/elementwise_bench.py [--rows 512]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "kafka-llm-service_b200"))
import torch  # noqa: E402
from kafka_b200 import cabi  # noqa: E402


def timed(fn, n=200, warm=20):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=512)
    ap.add_argument("--hidden", type=int, default=4096)
    ap.add_argument("--vocab", type=int, default=128256)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    R, H, V = args.rows, args.hidden, args.vocab
    out = {"rows": R, "hidden": H, "env": {k: v for k, v in os.environ.items() if k.startswith("KB_")}}

    x = torch.randn(R, H, device=dev)
    w = torch.ones(H, device=dev).bfloat16()
    y = torch.empty(R, H, device=dev, dtype=torch.float16)
    lib = cabi.load()
    st = cabi._stream()
    out["rmsnorm_us"] = round(timed(lambda: lib.kb_rmsnorm(cabi._ptr(x), cabi._ptr(w), cabi._ptr(y), R, H, 1e-5, 1, st)), 2)
    out["rmsnorm_GBps"] = round(R * H * 6 / out["rmsnorm_us"] / 1e3, 1)

    logits = torch.randn(R, V, device=dev)
    toks = torch.empty(R, dtype=torch.int32, device=dev)
    sp = cabi.make_sampling()
    import ctypes as C
    out["sample_full_vocab_us"] = round(timed(lambda: lib.kb_sample(cabi._ptr(logits), R, V, C.byref(sp), cabi._ptr(toks),
                                                                    st), n=50, warm=5), 2)
    out["sample_GBps"] = round(R * V * 4 / out["sample_full_vocab_us"] / 1e3, 1)
    sp2 = cabi.make_sampling(allowed_lo=32, allowed_hi=127, extra_a=10, extra_b=V - 1)
    out["sample_byte_range_us"] = round(timed(lambda: lib.kb_sample(cabi._ptr(logits), R, V, C.byref(sp2),
                                                                    cabi._ptr(toks), st), n=50, warm=5), 2)
    # temperature sampling over the whole vocabulary, without and with a nucleus cut (two histogram passes + the
    # row maximum before the sampling pass: four reads of the row instead of one)
    sp3 = cabi.make_sampling(temperature=0.8, seed=1, step=1)
    out["sample_temperature_full_vocab_us"] = round(timed(lambda: lib.kb_sample(cabi._ptr(logits), R, V, C.byref(sp3),
                                                                                cabi._ptr(toks), st), n=50, warm=5), 2)
    sp4 = cabi.make_sampling(temperature=0.8, seed=1, step=1, top_p=0.9)
    out["sample_top_p_full_vocab_us"] = round(timed(lambda: lib.kb_sample(cabi._ptr(logits), R, V, C.byref(sp4),
                                                                          cabi._ptr(toks), st), n=50, warm=5), 2)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
