"""Beyond the 8 192 positions the Llama-3-8B configuration stops at (a Llama-3.1 checkpoint declares 131 072): one
conversation grown to ~20 000 tokens on a head_dim-128 model — chunked incremental prefill on a prefix that reaches
16 k tokens (the tcgen05 prefill kernel's tile loop, 313 KV blocks through a scattered block table, RoPE angles up to
20 000 rad of the fastest pair), then decode steps that split the KV range many ways — against the fp32 oracle, plain
RoPE and Llama-3.1's scaled frequencies. Tolerance: max-abs 2e-2 on the logits."""
import dataclasses

import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 2e-2


@pytest.mark.parametrize("rope", ["plain", "llama3"])
def test_twenty_thousand_token_conversation_matches_the_oracle(cuda, rope):
    from kafka_b200 import cabi
    from kafka_b200.engine_core import EngineCore
    from kafka_b200.model_config import get_config
    from kafka_b200.weights import fuse_for_engine, generate_weights
    from oracle.llama_fp32 import OracleLlama
    cfg = dataclasses.replace(get_config("tiny-test-hd128"), name=f"tiny-hd128-long-{rope}", max_pos=32768,
                              rope_llama3=(8.0, 1.0, 4.0, 8192) if rope == "llama3" else None)
    w = generate_weights(cfg, seed=3)
    eng = EngineCore(cfg, fuse_for_engine(cfg, w, cuda), cuda, n_blocks=400, max_tokens=4096, max_seqs=4)
    oracle = OracleLlama(cfg, w)
    try:
        g = torch.Generator().manual_seed(8)
        n_total, chunk = 20000, 4096
        toks = torch.randint(0, 256, (n_total,), generator=g)
        n_blk = (n_total + 16 + 63) // 64
        table = torch.randperm(400, generator=g)[:n_blk].int()[None].contiguous().to(cuda)   # scattered physical blocks
        i32 = lambda x: torch.tensor(x, dtype=torch.int32, device=cuda)
        cache = oracle.new_cache()
        logits = torch.empty(1, cfg.vocab, dtype=torch.float32, device=cuda)
        out = torch.empty(1, dtype=torch.int32, device=cuda)
        worst = 0.0
        for lo in range(0, n_total, chunk):
            part = toks[lo:lo + chunk]
            eng.prefill(part.int().to(cuda), i32([0, len(part)]), i32([lo]), table, len(part), cabi.make_sampling(),
                        logits, out)
            torch.cuda.synchronize()
            ref = oracle.forward(part, cache)
            err = (logits[0].cpu() - ref).abs().max().item()
            worst = max(worst, err)
            assert err <= TOL, f"prefill chunk at {lo} ({rope} RoPE): logits max-abs err {err}"
        cur, ctx = ref.argmax(-1).view(1), n_total
        for step in range(4):
            eng.decode(cur.int().to(cuda), i32([ctx]), table, cabi.make_sampling(), out, max_ctx=ctx, logits_out=logits)
            torch.cuda.synchronize()
            ref = oracle.forward(cur, cache)
            err = (logits[0].cpu() - ref).abs().max().item()
            worst = max(worst, err)
            assert err <= TOL, f"decode at context {ctx} ({rope} RoPE): logits max-abs err {err}"
            cur, ctx = ref.argmax(-1).view(1), ctx + 1
        print(f"20 000-token conversation, {rope} RoPE: worst logit error {worst:.2e}")
    finally:
        eng.close()
