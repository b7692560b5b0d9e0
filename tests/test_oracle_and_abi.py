"""CPU checks of the test infrastructure and of the boundary:
  * the fp32 oracle against an independent implementation (transformers.LlamaForCausalLM) and against
    the frozen fixtures of tests/golden/oracle_tiny.json;
  * the C-ABI library loads and exports every symbol include/kafka_b200.h declares (no compute calls —
    there is no GPU here), and the Python binding covers them all;
  * the KV layout permutation used by tests equals the kernels' swizzle rule.
"""
import ctypes
import json
import re
from pathlib import Path

import pytest
import torch

from kafka_b200 import cabi
from kafka_b200.model_config import LlamaConfig, get_config, is_local_model
from kafka_b200.weights import fuse_for_engine, generable_ids, generate_weights
from oracle.llama_fp32 import OracleLlama

ROOT = Path(__file__).resolve().parent.parent
GOLD = Path(__file__).parent / "golden"


# ---- oracle ------------------------------------------------------------------------------------
def test_oracle_matches_transformers_llama():
    """Pins RoPE (rotate-half), GQA head grouping, RMSNorm placement and SwiGLU against HF's code."""
    transformers = pytest.importorskip("transformers")
    cfg = LlamaConfig("hf-check", hidden=128, n_layers=2, n_heads=4, n_kv_heads=2, head_dim=32, ffn=256, vocab=320,
                      rope_theta=10000.0, max_pos=256)
    w = generate_weights(cfg, seed=3)
    hf_cfg = transformers.LlamaConfig(
        hidden_size=cfg.hidden, num_hidden_layers=cfg.n_layers, num_attention_heads=cfg.n_heads,
        num_key_value_heads=cfg.n_kv_heads, head_dim=cfg.head_dim, intermediate_size=cfg.ffn, vocab_size=cfg.vocab,
        rope_theta=cfg.rope_theta, rms_norm_eps=cfg.rms_eps, max_position_embeddings=cfg.max_pos,
        attention_bias=False, mlp_bias=False, tie_word_embeddings=False, hidden_act="silu")
    hf_cfg._attn_implementation = "eager"
    model = transformers.LlamaForCausalLM(hf_cfg).to(torch.float32).eval()
    sd = {"model.embed_tokens.weight": w["embed"], "model.norm.weight": w["final_norm"], "lm_head.weight": w["lm_head"]}
    for i in range(cfg.n_layers):
        p, q = f"l{i}.", f"model.layers.{i}."
        sd.update({q + "input_layernorm.weight": w[p + "attn_norm"], q + "self_attn.q_proj.weight": w[p + "wq"],
                   q + "self_attn.k_proj.weight": w[p + "wk"], q + "self_attn.v_proj.weight": w[p + "wv"],
                   q + "self_attn.o_proj.weight": w[p + "wo"], q + "post_attention_layernorm.weight": w[p + "mlp_norm"],
                   q + "mlp.gate_proj.weight": w[p + "w_gate"], q + "mlp.up_proj.weight": w[p + "w_up"],
                   q + "mlp.down_proj.weight": w[p + "w_down"]})
    missing, unexpected = model.load_state_dict({k: v.float() for k, v in sd.items()}, strict=False)
    assert not [m for m in missing if "rotary" not in m] and not unexpected
    g = torch.Generator().manual_seed(0)
    toks = torch.randint(0, cfg.vocab, (37,), generator=g)
    with torch.no_grad():
        ref = model(toks[None]).logits[0]
    oracle = OracleLlama(cfg, w)
    got = oracle.forward(toks, oracle.new_cache(), all_logits=True)
    assert (got - ref).abs().max().item() < 2e-4
    # incremental: prefill 30, then 7 single-token steps, must equal the one-shot forward
    cache = oracle.new_cache()
    steps = [oracle.forward(toks[:30], cache)] + [oracle.forward(toks[30 + i: 31 + i], cache) for i in range(7)]
    assert (torch.stack(steps) - ref[29:]).abs().max().item() < 2e-4


def test_oracle_golden_fixtures():
    gold = json.loads((GOLD / "oracle_tiny.json").read_text())
    for key, g in gold.items():
        name, init = key.split("/")
        cfg = get_config(name)
        o = OracleLlama(cfg, generate_weights(cfg, seed=0, init=init))
        toks, lg = o.greedy(torch.tensor(g["prompt"]), 8, allowed=generable_ids(cfg))
        assert toks == g["tokens"], key
        head = torch.tensor(g["logits_head"])
        assert (lg[:, :8] - head).abs().max().item() < 1e-3, key
        for s, want in zip(lg, g["logits_sum"]):
            assert abs(float(s.double().sum()) - want) < 5e-2, key


def test_weights_are_bf16_exact_and_fusion_layout():
    cfg = get_config("tiny-test")
    w = generate_weights(cfg, seed=1)
    assert all(t.dtype == torch.bfloat16 for t in w.values())
    assert torch.equal(generate_weights(cfg, seed=1)["l1.wo"], w["l1.wo"])           # deterministic
    assert not torch.equal(generate_weights(cfg, seed=2)["l1.wo"], w["l1.wo"])
    f = fuse_for_engine(cfg, w, "cpu")
    assert f["l0.wqkv"].shape == (cfg.qkv_dim, cfg.hidden)
    # projection weights are bound as fp16 re-encodings of the bf16 values (KB_GEMM_F16): same numbers
    assert f["l0.wqkv"].dtype == torch.float16 and f["lm_head"].dtype == torch.float16
    assert f["embed"].dtype == torch.bfloat16 and f["l0.attn_norm"].dtype == torch.bfloat16
    def same(a, b):   # identical except values below 2^-14, which land on fp16's 6e-8 subnormal grid
        return (a.float() - b.float()).abs().max().item() < 6e-8
    assert same(f["l0.wqkv"][cfg.q_dim: cfg.q_dim + cfg.kv_dim], w["l0.wk"])
    assert same(f["l0.w_gate_up"][0::2], w["l0.w_gate"]) and same(f["l0.w_gate_up"][1::2], w["l0.w_up"])
    big = get_config("tinyllama-1.1b")
    wd = generate_weights(big, seed=0, layers=[0])["l0.w_down"]
    lossy = (wd.half().float() != wd.float()).float().mean().item()
    assert lossy < 0.01 and (wd.half().float() - wd.float()).abs().max().item() < 6e-8   # only sub-2^-14 values move
    m = generate_weights(cfg, seed=1, init="margin")
    ids = generable_ids(cfg)
    changed = (m["lm_head"] != w["lm_head"]).any(dim=1).nonzero().flatten()
    assert set(changed.tolist()) <= set(ids.tolist())


def test_model_registry():
    c = get_config("llama-3-8b")
    assert c.kv_bytes_per_token == 131072 and c.matmul_params() == 6_979_321_856
    assert abs(c.weight_bytes() - 16.06e9) < 0.01e9
    assert get_config("b200/TinyLlama-1.1B").kv_bytes_per_token == 22 * 1024
    assert is_local_model("llama-3-70b") and not is_local_model("gpt-4o")


# ---- C ABI -------------------------------------------------------------------------------------
def _declared_functions():
    text = (ROOT / "include" / "kafka_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(kb_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    names = _declared_functions()
    assert len(names) >= 20 and "kb_decode" in names and "kb_attn_decode_static_scratch" in names
    lib = ctypes.CDLL(str(cabi.lib_path()))
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/kafka_b200.h but not exported"
    assert sorted(cabi.SIGNATURES) == names, "Python binding and header disagree on the entry points"
    loaded = cabi.load()
    assert loaded.kb_abi_version() == cabi.KB_ABI_VERSION
    assert loaded.kb_static_scratch_capacity() == 2 * 1024 * 1024
    assert (loaded.kb_last_error() or b"") == b""


def test_header_constants_match_binding():
    text = (ROOT / "include" / "kafka_b200.h").read_text()
    assert int(re.search(r"#define KB_BLOCK_TOKENS (\d+)", text).group(1)) == cabi.KB_BLOCK_TOKENS
    assert int(re.search(r"#define KB_MAX_SPLITS (\d+)", text).group(1)) == cabi.KB_MAX_SPLITS
    assert int(re.search(r"#define KB_PROF_CATS (\d+)", text).group(1)) == len(cabi.PROF_CATS)
    assert ctypes.sizeof(cabi.ModelCfg) == 40 and ctypes.sizeof(cabi.Sampling) == 80     # 64 + top_p (4, padded to 8) + seq_top_p


def test_no_gpu_means_loud_failure():
    """No CPU fallback: without a device the engine refuses to start and says why."""
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    lib = cabi.load()
    cfg = get_config("tiny-test")
    c = cabi.ModelCfg(cfg.hidden, cfg.n_layers, cfg.n_heads, cfg.n_kv_heads, cfg.head_dim, cfg.ffn, cfg.vocab,
                      cfg.max_pos, cfg.rope_theta, cfg.rms_eps)
    h = ctypes.c_void_p()
    assert lib.kb_engine_create(0, ctypes.byref(c), ctypes.byref(h)) != 0
    assert "no CPU fallback" in cabi.last_error()
    from kafka_b200.engine_core import EngineCore
    with pytest.raises(cabi.KafkaB200Error, match="no CPU fallback"):
        EngineCore(cfg, {}, torch.device("cpu"), 1, 1, 1)
    bad = cabi.ModelCfg(100, 1, 1, 1, 64, 64, 8, 16, 10000.0, 1e-5)
    assert lib.kb_workspace_bytes(ctypes.byref(bad), 4, 4) == 0                         # hidden % 64 != 0


def test_kv_layout_matches_kernel_swizzle():
    from kafka_b200.kv_layout import tile_permutation
    for hd in (64, 128):
        perm = tile_permutation(hd).view(64, hd)
        for t in (0, 1, 7, 8, 13, 63):
            for d in (0, 7, 8, 63, hd - 1):
                # kb_common.cuh kv_swizzled_elem: [dim plane][token][64 dims], chunks XOR (token & 7)
                want = (d >> 6) * 64 * 64 + t * 64 + ((((d & 63) >> 3) ^ (t & 7)) << 3) + (d & 7)
                assert int(perm[t, d]) == want
        assert sorted(perm.view(-1).tolist()) == list(range(64 * hd))                    # a permutation of the tile


def test_product_never_imports_the_oracle():
    pkg = ROOT / "kafka-llm-service_b200"
    for f in pkg.rglob("*.py"):
        src = f.read_text()
        assert "import oracle" not in src and "from oracle" not in src, f
    for f in (pkg / "csrc").glob("*.cu*"):
        assert "oracle" not in f.read_text().lower(), f


def test_batched_oracle_equals_the_one_conversation_provider():
    """oracle/cpu_provider.BatchedOracle (packed projections, K/V kept between turns) must produce exactly
    the text of CPUOracleProvider (stateless, one conversation at a time) over several turns."""
    import asyncio
    from kafka_b200.llm import Message
    from kafka_b200.model_config import get_config
    from kafka_b200.weights import generate_weights
    from oracle.cpu_provider import BatchedOracle, CPUOracleProvider
    cfg = get_config("tiny-test")
    w = generate_weights(cfg, 0, init="margin")
    one, many = CPUOracleProvider(cfg, w), BatchedOracle(cfg, w)
    convs = [[Message(role="user", content=f"hello {i} " * (i + 1))] for i in range(5)]
    for turn in range(3):
        want = [asyncio.run(one.completion(c, max_tokens=12)) for c in convs]
        got = many.complete([f"k{i}" for i in range(5)], convs, 12)
        assert [(x.content, x.finish_reason) for x in want] == got
        for c, (t, _) in zip(convs, got):
            c += [Message(role="assistant", content=t), Message(role="user", content="more " + t[:3])]
    assert many.tokens_prefilled < sum(len(one.template.render(c)) for c in convs) * 2   # follow-ups reuse K/V
