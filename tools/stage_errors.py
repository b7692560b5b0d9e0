#!/usr/bin/env python
"""Where does the logit error of one decode step come from?  Runs a 1- or 2-layer model with Llama-3-8B's
layer geometry on the B200 and on the fp32 oracle, then compares the engine's workspace buffers of the LAST
layer (normed activations, rotated q/k/v, attention output, MLP inner, residual stream, final-normed rows,
logits) with the oracle's intermediates of the same step, stage by stage: max-abs and rms error, absolute and
relative to the stage's own rms.  Used to set the error budget of DESIGN.md §5 (numerical format).

    python tools/stage_errors.py [--layers 2] [--batch 40] [--model llama-3-8b]
"""
import argparse
import dataclasses
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "kafka-llm-service_b200")]

import torch  # noqa: E402


def align256(v):
    return (v + 255) & ~255


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=2)
    ap.add_argument("--batch", type=int, default=40)
    ap.add_argument("--model", default="llama-3-8b")
    ap.add_argument("--long", type=int, default=2200)
    args = ap.parse_args()
    from kafka_b200 import cabi
    from kafka_b200.engine_core import EngineCore
    from kafka_b200.model_config import get_config
    from kafka_b200.weights import fuse_for_engine, generate_weights
    from oracle.llama_fp32 import OracleLlama
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    cfg = dataclasses.replace(get_config(args.model), name="stage", n_layers=args.layers)
    w = generate_weights(cfg, seed=11, init="iid")
    T, S = 4096, 256
    eng = EngineCore(cfg, fuse_for_engine(cfg, w, dev), dev, n_blocks=800, max_tokens=T, max_seqs=S)
    oracle = OracleLlama(cfg, w)
    B = args.batch
    g = torch.Generator().manual_seed(22)
    lens = [int(x) for x in torch.randint(40, 120, (B,), generator=g)]
    lens[0] = args.long
    prompts = [torch.randint(0, 256, (n,), generator=g) for n in lens]
    mb = (max(lens) + 8 + 63) // 64
    tables = torch.zeros(B, mb, dtype=torch.int32)
    nxt = 0
    for i in range(B):
        need = (lens[i] + 8 + 63) // 64
        tables[i, :need] = torch.arange(nxt, nxt + need)
        nxt += need
    tables = tables.to(dev)
    i32 = lambda x: torch.tensor(x, dtype=torch.int32, device=dev)
    pos = 0
    while pos < B:
        idx, tok = [], 0
        while pos < B and tok + lens[pos] <= T:
            idx.append(pos); tok += lens[pos]; pos += 1
        cu = [0]
        for i in idx:
            cu.append(cu[-1] + lens[i])
        eng.prefill(torch.cat([prompts[i] for i in idx]).int().to(dev), i32(cu), i32([0] * len(idx)), tables[idx],
                    max(lens[i] for i in idx), None, None, None)
    caches = [oracle.new_cache() for _ in range(B)]
    cur = oracle.forward_packed(prompts, caches).argmax(-1)
    taps = {}
    ref = oracle.decode_batch(cur, caches, taps=taps)
    logits = torch.empty(B, cfg.vocab, dtype=torch.float32, device=dev)
    out = torch.empty(B, dtype=torch.int32, device=dev)
    eng.decode(cur.int().to(dev), i32(lens), tables, cabi.make_sampling(), out, max_ctx=max(lens), logits_out=logits)
    torch.cuda.synchronize()

    # the engine's workspace (kb_engine.cu layout_workspace): x, xr fp32 | xn, qkv, attn, h 16-bit | xl | logits
    ws = eng.workspace
    off = 0

    def take(nbytes, dtype, shape):
        nonlocal off
        t = ws[off:off + nbytes].view(dtype).view(shape)
        off += align256(nbytes)
        return t
    H, Q, KV, F = cfg.hidden, cfg.q_dim, cfg.kv_dim, cfg.ffn
    x = take(T * H * 4, torch.float32, (T, H))
    take(T * H * 4, torch.float32, (T, H))
    xn = take(T * H * 2, torch.float16, (T, H))
    qkv = take(T * cfg.qkv_dim * 2, torch.float16, (T, cfg.qkv_dim))
    attn = take(T * Q * 2, torch.float16, (T, Q))
    h = take(T * F * 2, torch.float16, (T, F))
    xl = take(S * H * 2, torch.float16, (S, H))

    def report(name, got, want):
        got, want = got.float().cpu(), want.float()
        e = got - want
        rms = want.pow(2).mean().sqrt().item()
        row = e.abs().amax(dim=1)
        print(f"{name:10s} rms(ref) {rms:9.4f}  max-abs err {e.abs().max().item():.3e}  rms err {e.pow(2).mean().sqrt().item():.3e}"
              f"  rel rms {e.pow(2).mean().sqrt().item() / rms:.3e}   worst row {int(row.argmax())} (ctx {lens[int(row.argmax())]})"
              f"  row0 (ctx {lens[0]}) max {row[0].item():.3e}  median row max {row.median().item():.3e}")
    print(f"{cfg.n_layers} layer(s), batch {B}, contexts {min(lens)}..{max(lens)}")
    report("q", qkv[:B, :Q], taps["q"])
    report("k", qkv[:B, Q:Q + KV], taps["k"])       # k in the qkv buffer is NOT rotated (the pool holds the rotated one)
    report("v", qkv[:B, Q + KV:], taps["v"])
    report("attn", attn[:B], taps["attn"])
    report("xn_mlp", xn[:B], taps["xn_mlp"])
    report("inner", h[:B], taps["inner"])
    report("x_final", x[:B], taps["x_final"])
    report("xl", xl[:B], taps["xl"])
    report("logits", logits, ref)
    # lm_head alone: the engine's own xl through an fp32 matmul vs its logits
    lm = oracle.w["lm_head"]
    report("lm_head", logits, xl[:B].float().cpu() @ lm.T)
    eng.close()


if __name__ == "__main__":
    main()
