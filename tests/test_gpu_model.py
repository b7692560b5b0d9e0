"""Model-level parity on the B200: the full forward (prefill, incremental prefill, decode) through
kb_prefill / kb_decode against the fp32 CPU oracle on the same seeded weights.

Tolerance: BASELINE.json north_star — bf16 logits within max-abs 2e-2 of fp32.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

LOGIT_TOL = 2e-2


def _setup(name, cuda, init="iid", n_blocks=64, max_tokens=512, max_seqs=16):
    from kafka_b200.model_config import get_config
    from kafka_b200.weights import generate_weights, fuse_for_engine
    from kafka_b200.engine_core import EngineCore
    from oracle.llama_fp32 import OracleLlama
    cfg = get_config(name)
    w = generate_weights(cfg, seed=0, init=init)
    eng = EngineCore(cfg, fuse_for_engine(cfg, w, cuda), cuda, n_blocks=n_blocks, max_tokens=max_tokens,
                     max_seqs=max_seqs)
    return cfg, eng, OracleLlama(cfg, w)


def _i32(x, cuda):
    return torch.tensor(x, dtype=torch.int32, device=cuda)


@pytest.mark.parametrize("name", ["tiny-test", "tiny-test-hd128"])
def test_prefill_then_decode_logits(cuda, name):
    from kafka_b200 import cabi
    cfg, eng, oracle = _setup(name, cuda)
    g = torch.Generator().manual_seed(1)
    lens = [5, 70, 129]                      # ragged; crosses one and two block boundaries
    prompts = [torch.randint(0, 256, (n,), generator=g) for n in lens]
    B, max_blocks = len(lens), 4
    bt = torch.arange(B * max_blocks, dtype=torch.int32).view(B, max_blocks).flip(0).contiguous().to(cuda)
    tokens = torch.cat(prompts).int().to(cuda)
    cu = _i32([0] + list(torch.tensor(lens).cumsum(0)), cuda)
    ctx = _i32([0] * B, cuda)
    logits = torch.empty(B, cfg.vocab, dtype=torch.float32, device=cuda)
    out_tok = torch.empty(B, dtype=torch.int32, device=cuda)
    sp = cabi.make_sampling()
    eng.prefill(tokens, cu, ctx, bt, max(lens), sp, logits, out_tok)
    torch.cuda.synchronize()

    caches, ref = [], []
    for p in prompts:
        c = oracle.new_cache()
        ref.append(oracle.forward(p, c))
        caches.append(c)
    ref = torch.stack(ref)
    err = (logits.cpu() - ref).abs().max().item()
    assert err <= LOGIT_TOL, f"prefill logits max-abs err {err}"

    # teacher-forced decode: feed the ORACLE's greedy token to both sides for 8 steps
    ctx_lens = _i32(lens, cuda)
    cur = ref.argmax(-1)
    for step in range(8):
        eng.decode(cur.int().to(cuda), ctx_lens, bt, sp, out_tok, max_ctx=max(lens) + step, logits_out=logits,
                   advance_ctx=True)
        torch.cuda.synchronize()
        ref = torch.stack([oracle.forward(cur[i:i + 1], caches[i]) for i in range(B)])
        err = (logits.cpu() - ref).abs().max().item()
        assert err <= LOGIT_TOL, f"decode step {step}: logits max-abs err {err}"
        cur = ref.argmax(-1)
    assert ctx_lens.cpu().tolist() == [n + 8 for n in lens]  # advance_ctx

    # incremental prefill: 20 more tokens on top of the cached prefix of every sequence
    more = [torch.randint(0, 256, (20,), generator=g) for _ in range(B)]
    tokens = torch.cat(more).int().to(cuda)
    cu = _i32([0, 20, 40, 60], cuda)
    eng.prefill(tokens, cu, ctx_lens, bt, 20, sp, logits, out_tok)
    torch.cuda.synchronize()
    ref = torch.stack([oracle.forward(more[i], caches[i]) for i in range(B)])
    err = (logits.cpu() - ref).abs().max().item()
    assert err <= LOGIT_TOL, f"incremental prefill logits max-abs err {err}"
    assert torch.equal(out_tok.cpu(), logits.argmax(-1).int().cpu())
    eng.close()


def test_greedy_agreement_margin_init(cuda):
    """Free-running greedy decode, 32 steps, GPU feeding back its own device-side tokens, vs the oracle.
    BASELINE north_star: tokens agree over 32 steps on >= 99% of threads."""
    from kafka_b200 import cabi
    from kafka_b200.weights import generable_ids
    cfg, eng, oracle = _setup("tiny-test", cuda, init="margin", n_blocks=16 * 3, max_tokens=2048, max_seqs=16)
    g = torch.Generator().manual_seed(5)
    B, max_blocks = 16, 3
    lens = [int(x) for x in torch.randint(4, 90, (B,), generator=g)]
    prompts = [torch.randint(32, 127, (n,), generator=g) for n in lens]
    bt = torch.arange(B * max_blocks, dtype=torch.int32).view(B, max_blocks).to(cuda)
    allowed = generable_ids(cfg)
    sp = cabi.make_sampling(allowed_lo=32, allowed_hi=127, extra_a=10, extra_b=cfg.vocab - 1)
    out_tok = torch.empty(B, dtype=torch.int32, device=cuda)
    eng.prefill(torch.cat(prompts).int().to(cuda), _i32([0] + list(torch.tensor(lens).cumsum(0)), cuda),
                _i32([0] * B, cuda), bt, max(lens), sp, None, out_tok)
    ctx_lens = _i32(lens, cuda)
    got = [out_tok.clone()]
    for step in range(31):
        nxt = torch.empty_like(out_tok)
        eng.decode(got[-1], ctx_lens, bt, sp, nxt, max_ctx=max(lens) + step, advance_ctx=True)
        got.append(nxt)
    torch.cuda.synchronize()
    got = torch.stack(got, 1).cpu()  # [B, 32]
    agree = 0
    for i in range(B):
        toks, _ = oracle.greedy(prompts[i], 32, allowed=allowed)
        agree += int(toks == got[i].tolist())
    assert agree / B >= 0.99, f"greedy agreement {agree}/{B}"
    eng.close()
