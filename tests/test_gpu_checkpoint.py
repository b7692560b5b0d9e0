"""A real checkpoint end to end on the GPU: a `transformers` Llama written to disk (safetensors + tokenizer.json),
registered under a model name, loaded layer by layer into the engine — logits against `transformers`' own fp32
forward (an implementation that is not ours), then a completion through the provider with the checkpoint's
tokenizer, every generated token checked against the fp32 oracle on the loaded weights."""
import asyncio
import sys
import os

import pytest
import torch

pytestmark = pytest.mark.gpu
transformers = pytest.importorskip("transformers")
sys.path.insert(0, os.path.dirname(__file__))

TOL = 2e-2


@pytest.fixture(scope="module", params=["plain_rope", "llama3_rope"])
def ckpt(tmp_path_factory, request):
    from test_checkpoint import _bpe_tokenizer_json
    from kafka_b200 import checkpoint
    from kafka_b200.model_config import REGISTRY
    d = tmp_path_factory.mktemp("hf-ckpt")
    # head_dim 128 (the tcgen05 prefill-attention path), GQA group 4, tied-free, 3 layers
    # llama3_rope: Llama-3.1's scaled RoPE — the engine then rotates with a frequency table (kb_set_rope_inv_freq)
    rope = ({"rope_parameters": {"rope_theta": 500000.0, "rope_type": "llama3", "factor": 8.0, "low_freq_factor": 1.0,
                                 "high_freq_factor": 4.0, "original_max_position_embeddings": 128}}
            if request.param == "llama3_rope" else {"rope_theta": 500000.0})
    hf_cfg = transformers.LlamaConfig(hidden_size=512, intermediate_size=1024, num_hidden_layers=3, num_attention_heads=4,
                                      num_key_value_heads=1, head_dim=128, vocab_size=512, max_position_embeddings=4096,
                                      rms_norm_eps=1e-5, tie_word_embeddings=False, **rope)
    torch.manual_seed(7)
    m = transformers.LlamaForCausalLM(hf_cfg)
    with torch.no_grad():
        for n, p in m.named_parameters():
            if "norm" in n:
                p.add_(torch.randn_like(p) * 0.1)
            p.copy_(p.bfloat16().float())
    m.eval().save_pretrained(d, safe_serialization=True, max_shard_size="2MB")
    _bpe_tokenizer_json(d / "tokenizer.json")
    cfg = checkpoint.register_checkpoint(str(d), "hf-tiny-ckpt")
    assert (cfg.rope_llama3 is not None) == (request.param == "llama3_rope")
    yield cfg, m, str(d)
    REGISTRY.pop(cfg.name, None)
    checkpoint.CHECKPOINTS.pop(cfg.name, None)


def test_engine_logits_on_loaded_weights_match_transformers(cuda, ckpt):
    from kafka_b200 import cabi, checkpoint
    from kafka_b200.engine_core import EngineCore
    cfg, m, d = ckpt
    eng = EngineCore(cfg, checkpoint.load_fused(d, cfg, cuda), cuda, n_blocks=32, max_tokens=1024, max_seqs=8)
    try:
        g = torch.Generator().manual_seed(2)
        lens = [150, 64, 9]
        prompts = [torch.randint(0, 400, (n,), generator=g) for n in lens]
        cu = [0]
        for n in lens:
            cu.append(cu[-1] + n)
        i32 = lambda x: torch.tensor(x, dtype=torch.int32, device=cuda)
        tables = torch.tensor([[3, 9, 1], [4, 0, 0], [7, 0, 0]], dtype=torch.int32, device=cuda)
        logits = torch.empty(3, cfg.vocab, dtype=torch.float32, device=cuda)
        out = torch.empty(3, dtype=torch.int32, device=cuda)
        eng.prefill(torch.cat(prompts).int().to(cuda), i32(cu), i32([0, 0, 0]), tables, max(lens), cabi.make_sampling(),
                    logits, out)
        torch.cuda.synchronize()
        with torch.no_grad():
            want = torch.stack([m(p[None]).logits[0, -1] for p in prompts])
        err = (logits.cpu() - want).abs().max().item()
        assert want.abs().max() > 0.5 and err <= TOL, f"logits vs transformers: max-abs err {err}"
        # one decode step on the resident KV: transformers recomputes the whole sequence
        nxt = want.argmax(-1)
        eng.decode(nxt.int().to(cuda), i32(lens), tables, cabi.make_sampling(), out, max_ctx=max(lens), logits_out=logits)
        torch.cuda.synchronize()
        with torch.no_grad():
            want = torch.stack([m(torch.cat([p, nxt[i:i + 1]])[None]).logits[0, -1] for i, p in enumerate(prompts)])
        err = (logits.cpu() - want).abs().max().item()
        assert err <= TOL, f"decode logits vs transformers: max-abs err {err}"
    finally:
        eng.close()


def test_completion_through_the_provider_with_the_checkpoints_tokenizer(cuda, ckpt):
    from kafka_b200 import checkpoint, factory
    from kafka_b200.llm import Message
    from kafka_b200.tokenizer import HFTokenizer
    from oracle.llama_fp32 import OracleLlama
    cfg, m, d = ckpt
    provider = factory.make_provider("b200/HF-Tiny-Ckpt", default_temperature=0.0, devices=[0], n_blocks=64, max_seqs=8,
                                     max_tokens=1024)
    try:
        tok = provider.tokenizer
        assert isinstance(tok, HFTokenizer) and provider.engine.stepper.eos_id == tok.eot_id
        msgs = [Message(role="system", content="be brief"), Message(role="user", content="hello world, naïve café 😀")]
        resp = asyncio.run(provider.completion(msgs, temperature=0.0, max_tokens=12))
        assert isinstance(resp.content, str) and resp.usage["completion_tokens"] >= 1
        # the same conversation on the fp32 oracle over the loaded weights, teacher-forced with the provider's text:
        # at every position the emitted token must be (within 2 x the logit tolerance) the best allowed candidate
        prompt = provider.render(msgs, None)
        gen = tok.encode(resp.content)
        oracle = OracleLlama(cfg, checkpoint.load_canonical(d, cfg))
        lo, hi, ea, _ = tok.generable_range()
        allowed = torch.zeros(cfg.vocab, dtype=torch.bool)
        allowed[lo:hi] = True
        allowed[ea] = True
        all_logits = oracle.forward(torch.tensor(prompt + gen), oracle.new_cache(), all_logits=True)
        if tok.decode(gen) == resp.content and len(gen) == resp.usage["completion_tokens"]:   # canonical re-tokenisation
            for k, t in enumerate(gen):
                row = all_logits[len(prompt) - 1 + k].masked_fill(~allowed, float("-inf"))
                assert row[t] >= row.max() - 4e-2, f"token {k}: logit {row[t]:.4f} vs best {row.max():.4f}"
        # a second turn re-sends the conversation: the resident prefix is found through the HF tokenizer's ids
        msgs2 = msgs + [resp.to_message(), Message(role="user", content="again")]
        resp2 = asyncio.run(provider.completion(msgs2, temperature=0.0, max_tokens=4))
        assert resp2.usage.get("cached_tokens", 0) >= len(prompt) - 64
    finally:
        factory.shutdown_engines()
