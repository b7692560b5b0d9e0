"""Per-kernel parity on the B200: every CUDA kernel, called through the C ABI, against a plain fp32
PyTorch statement of the same op on the same seeded inputs.

Tolerances (written where used): outputs that are rounded to bf16 are compared with rtol = atol = 2e-2
of the fp32 result (bf16 has 8 mantissa bits, 2^-8 = 3.9e-3 per rounding, a few roundings per op);
fp32 outputs of bf16-input GEMMs with rtol = atol = 2e-3.
"""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def kb():
    from kafka_b200 import cabi
    cabi.load()
    return cabi


def _bf16(t):
    return t.to(torch.bfloat16)


def _f16(t):
    """q / k / v and the KV pools are fp16 (include/kafka_b200.h)."""
    return t.to(torch.float16)


# ------------------------------------------------------------------------------------------------
def test_embed(kb, cuda):
    g = torch.Generator(device="cpu").manual_seed(0)
    table = _bf16(torch.randn(300, 256, generator=g)).to(cuda)
    toks = torch.randint(0, 300, (37,), generator=g, dtype=torch.int32).to(cuda)
    x = kb.embed(toks, table)
    torch.cuda.synchronize()
    assert torch.equal(x, table[toks.long()].float())  # exact: a gather + widening


@pytest.mark.parametrize("n,hidden", [(1, 256), (5, 2048), (33, 4096), (7, 8192)])
def test_rmsnorm(kb, cuda, n, hidden):
    g = torch.Generator(device="cpu").manual_seed(n * 1000 + hidden)
    x = (torch.randn(n, hidden, generator=g) * 3).to(cuda)
    w = _bf16(1 + 0.1 * torch.randn(hidden, generator=g)).to(cuda)
    y = kb.rmsnorm(x, w, 1e-5)
    ref = x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + 1e-5) * w.float()
    torch.testing.assert_close(y.float(), ref, rtol=2e-2, atol=2e-2)


# ------------------------------------------------------------------------------------------------
GEMM_SHAPES = [
    # M, N, K
    (128, 256, 64),
    (128, 128, 128),
    (1, 512, 256),
    (17, 136, 192),      # partial M and N tiles
    (300, 1000, 320),    # partial everything, N % 8 == 0
    (512, 4096, 4096),   # decode-step O projection
    (512, 6144, 4096),   # decode-step QKV
    (64, 512, 14336),    # long K
    (1000, 768, 512),    # many M tiles
]


# negative tile_n = CTA-pair kernel, 256 x |tile_n| tile; beyond 256 = the wide tile (one 512-column accumulator,
# two MMAs per k-step: N = 256 and N = |tile_n| - 256)
@pytest.mark.parametrize("tile_n", [0, 64, 128, 256, -128, -256, -272, -400, -512])
@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
def test_gemm_bf16(kb, cuda, M, N, K, tile_n):
    g = torch.Generator(device="cpu").manual_seed(M * 7 + N * 3 + K)
    a = _bf16(torch.randn(M, K, generator=g)).to(cuda)
    w = _bf16(torch.randn(N, K, generator=g) / math.sqrt(K)).to(cuda)
    out = kb.gemm(a, w, kb.EPI_BF16, tile_n=tile_n)
    ref = a.float() @ w.float().T
    torch.testing.assert_close(out.float(), ref, rtol=2e-2, atol=2e-2)


@pytest.mark.parametrize("tile_n", [64, 128, 256, -128, -256, -320, -512])
@pytest.mark.parametrize("M,N,K", [(17, 136, 192), (300, 1000, 320), (512, 4096, 1024)])
def test_gemm_f32_and_residual(kb, cuda, M, N, K, tile_n):
    g = torch.Generator(device="cpu").manual_seed(M + N + K)
    a = _bf16(torch.randn(M, K, generator=g)).to(cuda)
    w = _bf16(torch.randn(N, K, generator=g) / math.sqrt(K)).to(cuda)
    ref = a.float() @ w.float().T
    out = kb.gemm(a, w, kb.EPI_F32, tile_n=tile_n)
    torch.testing.assert_close(out, ref, rtol=2e-3, atol=2e-3)
    resid = torch.randn(M, N, generator=g).to(cuda)
    want = resid + ref
    out2 = kb.gemm(a, w, kb.EPI_RESID_F32, resid=resid, tile_n=tile_n)
    torch.testing.assert_close(out2, want, rtol=2e-3, atol=2e-3)
    # in place: out aliases resid (how the engine uses it)
    r2 = resid.clone()
    kb.gemm(a, w, kb.EPI_RESID_F32, resid=r2, out=r2, tile_n=tile_n)
    torch.testing.assert_close(r2, want, rtol=2e-3, atol=2e-3)


@pytest.mark.parametrize("tile_n", [64, 128, 256, -128, -256, -384, -512])
@pytest.mark.parametrize("M,F,K", [(17, 80, 192), (300, 512, 256), (512, 1408, 1024)])
def test_gemm_swiglu(kb, cuda, M, F, K, tile_n):
    g = torch.Generator(device="cpu").manual_seed(M + F + K)
    a = _bf16(torch.randn(M, K, generator=g)).to(cuda)
    wg = _bf16(torch.randn(F, K, generator=g) / math.sqrt(K)).to(cuda)
    wu = _bf16(torch.randn(F, K, generator=g) / math.sqrt(K)).to(cuda)
    w = torch.stack([wg, wu], dim=1).reshape(2 * F, K).contiguous()
    out = kb.gemm(a, w, kb.EPI_SWIGLU_BF16, tile_n=tile_n)
    ref = torch.nn.functional.silu(a.float() @ wg.float().T) * (a.float() @ wu.float().T)
    torch.testing.assert_close(out.float(), ref, rtol=2e-2, atol=2e-2)


@pytest.mark.parametrize("tile_n", [64, 128, 256, -128, -256, -288, -512])
def test_gemm_fp16_operands(kb, cuda, tile_n):
    """The hot path's operand format: A (activations) and W (bf16 values re-encoded as fp16) both fp16,
    fp32 accumulate. The fp32 epilogue must match an fp32 matmul of the same values to
    accumulation-order error only (rtol = atol = 2e-3 as for the bf16 case)."""
    g = torch.Generator(device="cpu").manual_seed(99 + tile_n)
    M, N, K = 300, 1000, 320
    a = torch.randn(M, K, generator=g).half().to(cuda)
    w = _bf16(torch.randn(N, K, generator=g) / math.sqrt(K)).half().to(cuda)
    assert torch.equal(w.float(), w.float().bfloat16().float())      # still exactly the bf16 values
    ref = a.float() @ w.float().T
    torch.testing.assert_close(kb.gemm(a, w, kb.EPI_F32, tile_n=tile_n), ref, rtol=2e-3, atol=2e-3)
    with pytest.raises(kb.KafkaB200Error, match="formats must match"):
        kb.gemm(a, w.bfloat16())
    torch.testing.assert_close(kb.gemm(a, w, kb.EPI_BF16, tile_n=tile_n).float(), ref, rtol=2e-2, atol=2e-2)
    F = 256
    wg = _bf16(torch.randn(F, K, generator=g) / math.sqrt(K)).half().to(cuda)
    wu = _bf16(torch.randn(F, K, generator=g) / math.sqrt(K)).half().to(cuda)
    wi = torch.stack([wg, wu], dim=1).reshape(2 * F, K).contiguous()
    h = kb.gemm(a, wi, kb.EPI_SWIGLU_F16, tile_n=tile_n)
    assert h.dtype == torch.float16
    want = torch.nn.functional.silu(a.float() @ wg.float().T) * (a.float() @ wu.float().T)
    # fp16 output: 2^-11 relative rounding -> 2e-3
    torch.testing.assert_close(h.float(), want, rtol=2e-3, atol=2e-3)


@pytest.mark.parametrize("splits", [0, 1, 2, 4])
@pytest.mark.parametrize("M,N,K", [(1, 512, 256), (5, 4096, 4096), (16, 136, 192), (17, 1000, 320), (33, 6144, 1024),
                                   (64, 4096, 14336), (100, 384, 512), (128, 4096, 4096)])
def test_gemm_small_m_kernel(kb, cuda, M, N, K, splits):
    """The small-M kernel (operands swapped, cluster split-K with a rank-ordered reduction through distributed
    shared memory): every epilogue, forced and automatic split counts, ragged M / N."""
    g = torch.Generator(device="cpu").manual_seed(M * 11 + N * 5 + K + splits)
    a = torch.randn(M, K, generator=g).half().to(cuda)
    w = _bf16(torch.randn(N, K, generator=g) / math.sqrt(K)).half().to(cuda)
    ref = a.float() @ w.float().T
    torch.testing.assert_close(kb.gemm(a, w, kb.EPI_F32, tile_n=splits), ref, rtol=2e-3, atol=2e-3)
    torch.testing.assert_close(kb.gemm(a, w, kb.EPI_BF16, tile_n=splits).float(), ref, rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(kb.gemm(a, w, kb.EPI_F16, tile_n=splits).float(), ref, rtol=2e-3, atol=2e-3)
    resid = torch.randn(M, N, generator=g).to(cuda)
    r2 = resid.clone()
    kb.gemm(a, w, kb.EPI_RESID_F32, resid=r2, out=r2, tile_n=splits)          # in place, as the engine runs it
    torch.testing.assert_close(r2, resid + ref, rtol=2e-3, atol=2e-3)
    out3 = kb.gemm(a, w, kb.EPI_RESID_F32, resid=resid, tile_n=splits)
    torch.testing.assert_close(out3, resid + ref, rtol=2e-3, atol=2e-3)
    if N % 16 == 0:
        F = N // 2
        wg, wu = w[0::2].float(), w[1::2].float()                             # interleaved (gate_j, up_j) rows
        want = torch.nn.functional.silu(a.float() @ wg.T) * (a.float() @ wu.T)
        h = kb.gemm(a, w, kb.EPI_SWIGLU_F16, tile_n=splits)
        assert h.shape == (M, F)
        torch.testing.assert_close(h.float(), want, rtol=2e-3, atol=2e-3)
        hb = kb.gemm(a, w, kb.EPI_SWIGLU_BF16, tile_n=splits)
        torch.testing.assert_close(hb.float(), want, rtol=2e-2, atol=2e-2)
    # the reduction order is fixed: two runs give the same bits
    assert torch.equal(kb.gemm(a, w, kb.EPI_F32, tile_n=splits), kb.gemm(a, w, kb.EPI_F32, tile_n=splits))


# ---- the decode step's projections at Llama-3-8B size (the launches bench.py times) -------------------
def _llama_operands(M, N, K, seed, cuda):
    """fp16 activations ~N(0,1) and bf16-valued fp16 weights ~N(0, 0.02), drawn on the device."""
    g = torch.Generator(device=cuda).manual_seed(seed)
    a = torch.randn(M, K, generator=g, device=cuda).half()
    w = (torch.randn(N, K, generator=g, device=cuda) * 0.02).bfloat16().half()
    return a, w


def _fp32_matmul(a, w):
    """a @ w.T in true fp32 on the device (TF32 off), in row blocks to bound memory."""
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        wf = w.float()
        return torch.cat([a[i:i + 128].float() @ wf.T for i in range(0, a.shape[0], 128)])
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev


@pytest.mark.parametrize("tile_n", [0, -208, -256, 256, -400, -448, -512])
def test_gemm_gate_up_at_llama3_8b_size(kb, cuda, tile_n):
    """512 x 28 672 x 4 096 with the SwiGLU fp16 epilogue: the heuristic's tile, the run-time width the
    decode step uses (bn = 208), the widest pair tile and the 1-CTA tile."""
    M, F, K = 512, 14336, 4096
    a, w = _llama_operands(M, 2 * F, K, 31, cuda)
    h = kb.gemm(a, w, kb.EPI_SWIGLU_F16, tile_n=tile_n)
    d = _fp32_matmul(a, w)
    want = torch.nn.functional.silu(d[:, 0::2]) * d[:, 1::2]
    assert h.shape == (M, F) and h.dtype == torch.float16
    torch.testing.assert_close(h.float(), want, rtol=2e-3, atol=2e-3)      # fp16 output: 2^-11 relative


SPLIT2, SPLIT4 = 2 * 1024, 4 * 1024      # kb_gemm_ex: tile_n = -(1024 * splits + bn) forces split-K on the CTA-pair kernel


@pytest.mark.parametrize("tile_n", [0, -112, -256, 128, -(SPLIT2 + 240), -(SPLIT4 + 256), -(SPLIT2 + 112), -512,
                                    -(SPLIT4 + 512)])
def test_gemm_down_in_place_residual_at_llama3_8b_size(kb, cuda, tile_n):
    """512 x 4 096 x 14 336, x += h W^T in place (red.global.add epilogue), K = 14 336; the heuristic (split-K in
    2 parts on 256 x 240 tiles), plain tiles and forced split-K."""
    M, N, K = 512, 4096, 14336
    a, w = _llama_operands(M, N, K, 32, cuda)
    g = torch.Generator(device=cuda).manual_seed(33)
    x0 = torch.randn(M, N, generator=g, device=cuda) * 4
    want = x0 + _fp32_matmul(a, w)
    x = x0.clone()
    kb.gemm(a, w, kb.EPI_RESID_F32, resid=x, out=x, tile_n=tile_n)
    torch.testing.assert_close(x, want, rtol=2e-3, atol=2e-3)
    # the partial sums of a tile are added in a fixed order: every run gives the same bits
    for _ in range(3):
        y = x0.clone()
        kb.gemm(a, w, kb.EPI_RESID_F32, resid=y, out=y, tile_n=tile_n)
        assert torch.equal(x, y)


@pytest.mark.parametrize("M,N,K,tile_n", [(300, 1000, 512, -(SPLIT2 + 128)), (512, 4096, 4096, -(SPLIT2 + 240)),
                                          (512, 4096, 4096, -(SPLIT4 + 256)), (1000, 768, 1024, -(SPLIT2 + 64)),
                                          (200, 4096, 14336, 0), (130, 4096, 4096, 0)])
def test_gemm_split_k_ragged_and_multi_round(kb, cuda, M, N, K, tile_n):
    """Split-K work units: partial M / N tiles, more units than CTA pairs (several rounds: a unit may wait for a
    unit of an earlier round), and the shapes where the heuristic picks it on its own."""
    a, w = _llama_operands(M, N, K, 41, cuda)
    g = torch.Generator(device=cuda).manual_seed(42)
    x0 = torch.randn(M, N, generator=g, device=cuda)
    want = x0 + _fp32_matmul(a, w)
    x = x0.clone()
    kb.gemm(a, w, kb.EPI_RESID_F32, resid=x, out=x, tile_n=tile_n)
    torch.testing.assert_close(x, want, rtol=2e-3, atol=2e-3)
    y = x0.clone()
    kb.gemm(a, w, kb.EPI_RESID_F32, resid=y, out=y, tile_n=tile_n)
    assert torch.equal(x, y)
    if tile_n < -1024:      # split-K is only defined for the in-place residual epilogue
        with pytest.raises(kb.KafkaB200Error, match="in-place residual"):
            kb.gemm(a, w, kb.EPI_F32, tile_n=tile_n)


@pytest.mark.parametrize("M,tile_n", [(512, 0), (512, -256), (200, 0), (8, 0), (8, 4), (1, 0)])
def test_gemm_lm_head_at_llama3_8b_size(kb, cuda, M, tile_n):
    """M x 128 256 x 4 096 with fp32 logits out: the CTA-pair kernel at M = 512 / 200 and the small-M kernel
    (operands swapped, cluster split-K) at M = 8 / 1."""
    N, K = 128256, 4096
    a, w = _llama_operands(M, N, K, 34, cuda)
    w = (w.float() * (1.0 / 0.02 / 64.0)).bfloat16().half()                 # lm_head scale: std 1 / sqrt(H)
    out = kb.gemm(a, w, kb.EPI_F32, tile_n=tile_n)
    torch.testing.assert_close(out, _fp32_matmul(a, w), rtol=2e-3, atol=2e-3)


@pytest.mark.parametrize("M", [512, 40])
def test_gemm_qkv_and_o_at_llama3_8b_size(kb, cuda, M):
    a, w = _llama_operands(M, 6144, 4096, 35, cuda)
    torch.testing.assert_close(kb.gemm(a, w, kb.EPI_BF16).float(), _fp32_matmul(a, w), rtol=2e-2, atol=2e-2)
    a, w = _llama_operands(M, 4096, 4096, 36, cuda)
    g = torch.Generator(device=cuda).manual_seed(37)
    x = torch.randn(M, 4096, generator=g, device=cuda)
    want = x + _fp32_matmul(a, w)
    kb.gemm(a, w, kb.EPI_RESID_F32, resid=x, out=x)
    torch.testing.assert_close(x, want, rtol=2e-3, atol=2e-3)


def test_fp16_activation_outputs_of_norm_and_attention(kb, cuda):
    from kafka_b200 import kv_layout
    g = torch.Generator(device="cpu").manual_seed(5)
    x = (torch.randn(9, 2048, generator=g) * 3).to(cuda)
    w = _bf16(1 + 0.1 * torch.randn(2048, generator=g)).to(cuda)
    y = kb.rmsnorm(x, w, 1e-5, out_f16=True)
    ref = x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + 1e-5) * w.float()
    assert y.dtype == torch.float16
    torch.testing.assert_close(y.float(), ref, rtol=2e-3, atol=2e-3)
    # attention: the fp16 output must equal the bf16 output up to the output rounding
    hd, n_heads, n_kv = 128, 8, 2
    kp, vp = _alloc_pool(8, n_kv, hd, cuda, 6)
    bt = torch.arange(8, dtype=torch.int32, device=cuda).view(2, 4)
    q = _f16(torch.randn(2, n_heads * hd, generator=g)).to(cuda)
    lens = torch.tensor([200, 77], dtype=torch.int32, device=cuda)
    a16 = kb.attn_decode(q, kp, vp, lens, bt, n_heads, n_kv, hd, out_f16=True)
    ab = kb.attn_decode(q, kp, vp, lens, bt, n_heads, n_kv, hd)
    assert a16.dtype == torch.float16
    torch.testing.assert_close(a16.float(), ab.float(), rtol=1e-2, atol=1e-2)
    cu = torch.tensor([0, 5, 12], dtype=torch.int32, device=cuda)
    ctx = torch.tensor([100, 3], dtype=torch.int32, device=cuda)
    qq = _f16(torch.randn(12, n_heads * hd, generator=g)).to(cuda)
    p16 = kb.attn_prefill(qq, kp, vp, cu, ctx, bt, 7, n_heads, n_kv, hd, out_f16=True)
    pb = kb.attn_prefill(qq, kp, vp, cu, ctx, bt, 7, n_heads, n_kv, hd)
    torch.testing.assert_close(p16.float(), pb.float(), rtol=1e-2, atol=1e-2)
    del kv_layout


def test_fp16_outputs_saturate_instead_of_overflowing(kb, cuda):
    """Every 16-bit activation is fp16, whose range ends at 65 504. Random-init weights never get near it, a
    real checkpoint's outlier channels can: the epilogues must clamp to +-65504 rather than emit inf (which
    would turn the rest of the forward pass, and every later token reading that K/V, into NaN)."""
    g = torch.Generator(device="cpu").manual_seed(77)
    M, K, F = 64, 256, 256
    a = torch.randn(M, K, generator=g).half().to(cuda)
    # gate / up outputs with std ~ 640: silu(g) * u reaches several 1e5
    wg = (torch.randn(F, K, generator=g) * 40.0 / math.sqrt(K) * 16).bfloat16().half().to(cuda)
    wu = (torch.randn(F, K, generator=g) * 40.0 / math.sqrt(K) * 16).bfloat16().half().to(cuda)
    wi = torch.stack([wg, wu], dim=1).reshape(2 * F, K).contiguous()
    want = torch.nn.functional.silu(a.float() @ wg.float().T) * (a.float() @ wu.float().T)
    assert want.abs().max() > 2e5                                    # the case really overflows fp16
    for tile_n in (0, 128, -128, 1):
        h = kb.gemm(a[:32] if tile_n == 1 else a, wi, kb.EPI_SWIGLU_F16, tile_n=tile_n)
        w_ = want[:32] if tile_n == 1 else want
        assert torch.isfinite(h).all(), f"tile_n={tile_n}: inf / nan in the fp16 SwiGLU output"
        # products of two O(1000) values: one fp16 rounding plus the fp32 accumulation-order error of both factors
        torch.testing.assert_close(h.float(), w_.clamp(-65504, 65504), rtol=4e-3, atol=4e-3)
        # the saturated activations go through the down projection without trouble (fp32 accumulate)
        wd = (torch.randn(K, F, generator=g) * 0.02).bfloat16().half().to(cuda)
        x = torch.zeros(h.shape[0], K, device=cuda)
        kb.gemm(h, wd, kb.EPI_RESID_F32, resid=x, out=x, tile_n=0)
        assert torch.isfinite(x).all()
    # plain fp16 output (QKV projection -> q / k / v) and the RMSNorm output
    big = kb.gemm(a, (wg.float() * 200).bfloat16().half(), kb.EPI_F16)
    ref = (a.float() @ (wg.float() * 200).bfloat16().float().T).clamp(-65504, 65504)
    assert ref.abs().max() == 65504 and torch.isfinite(big).all()
    torch.testing.assert_close(big.float(), ref, rtol=2e-3, atol=2e-3)
    x = (torch.randn(4, 512, generator=g) * 3).to(cuda)
    w = torch.full((512,), 6e4, dtype=torch.bfloat16, device=cuda)     # a norm weight that pushes 1.x * w past 65504
    y = kb.rmsnorm(x, w, 1e-5, out_f16=True)
    assert torch.isfinite(y).all() and y.abs().max() == 65504


def test_gemm_rejects_bad_shapes(kb, cuda):
    a = torch.zeros(4, 100, dtype=torch.bfloat16, device=cuda)
    w = torch.zeros(8, 100, dtype=torch.bfloat16, device=cuda)
    with pytest.raises(kb.KafkaB200Error, match="multiple of 64"):
        kb.gemm(a, w)


# ------------------------------------------------------------------------------------------------
def _rope_ref(x, pos, theta):
    """x [T, H, hd] fp32, HF rotate-half."""
    hd = x.shape[-1]
    inv = 1.0 / (theta ** (torch.arange(0, hd, 2, dtype=torch.float32, device=x.device) / hd))
    f = pos.float()[:, None] * inv[None]
    c, s = f.cos()[:, None, :], f.sin()[:, None, :]
    x1, x2 = x[..., : hd // 2], x[..., hd // 2:]
    return torch.cat([x1 * c - x2 * s, x2 * c + x1 * s], -1)


def _alloc_pool(n_blocks, n_kv, hd, device, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    k = _f16(torch.randn(n_blocks, n_kv, 64 * hd, generator=g)).to(device)
    v = _f16(torch.randn(n_blocks, n_kv, 64 * hd, generator=g)).to(device)
    return k, v


@pytest.mark.parametrize("hd,n_heads,n_kv,theta", [(64, 4, 2, 10000.0), (128, 8, 2, 500000.0)])
def test_rope_append(kb, cuda, hd, n_heads, n_kv, theta):
    from kafka_b200 import kv_layout
    g = torch.Generator(device="cpu").manual_seed(hd)
    new = [5, 70, 1]          # new tokens per sequence
    ctx = [0, 60, 130]        # cached tokens per sequence (second one crosses a block boundary)
    n_seqs, T = len(new), sum(new)
    max_blocks = 4
    # distinct, shuffled physical blocks
    perm = torch.randperm(n_seqs * max_blocks, generator=g).view(n_seqs, max_blocks).int()
    kp, vp = _alloc_pool(n_seqs * max_blocks, n_kv, hd, cuda, 1)
    kp0, vp0 = kp.clone(), vp.clone()
    qkv = _f16(torch.randn(T, (n_heads + 2 * n_kv) * hd, generator=g)).to(cuda)
    qkv0 = qkv.clone()
    cu = torch.tensor([0] + list(torch.tensor(new).cumsum(0)), dtype=torch.int32, device=cuda)
    ctx_t = torch.tensor(ctx, dtype=torch.int32, device=cuda)
    bt = perm.to(cuda)
    kb.rope_append(qkv, cu, ctx_t, bt, kp, vp, n_heads, n_kv, hd, theta)
    torch.cuda.synchronize()

    pos = torch.cat([torch.arange(c, c + n) for c, n in zip(ctx, new)]).to(cuda)
    q0 = qkv0[:, : n_heads * hd].float().view(T, n_heads, hd)
    k0 = qkv0[:, n_heads * hd: (n_heads + n_kv) * hd].float().view(T, n_kv, hd)
    v0 = qkv0[:, (n_heads + n_kv) * hd:].view(T, n_kv, hd)
    q_ref = _rope_ref(q0, pos, theta)
    k_ref = _rope_ref(k0, pos, theta)
    # tolerance: one fp16 rounding of an O(1) value (2^-11 relative) plus fp32 sin/cos of angles up to ~200 rad
    torch.testing.assert_close(qkv[:, : n_heads * hd].float().view(T, n_heads, hd), q_ref, rtol=2e-3, atol=2e-3)
    # v columns of qkv untouched
    assert torch.equal(qkv[:, (n_heads + n_kv) * hd:], qkv0[:, (n_heads + n_kv) * hd:])
    off = 0
    touched_k = torch.zeros_like(kp, dtype=torch.bool)
    for s in range(n_seqs):
        L = ctx[s] + new[s]
        k_all = kv_layout.read_tokens(kp, perm[s].tolist(), L, hd)
        v_all = kv_layout.read_tokens(vp, perm[s].tolist(), L, hd)
        torch.testing.assert_close(k_all[ctx[s]:].float(), k_ref[off: off + new[s]], rtol=2e-3, atol=2e-3)
        assert torch.equal(v_all[ctx[s]:], v0[off: off + new[s]])  # V is a pure copy
        # cached prefix untouched
        k_before = kv_layout.read_tokens(kp0, perm[s].tolist(), L, hd)
        assert torch.equal(k_all[: ctx[s]], k_before[: ctx[s]])
        off += new[s]
    # nothing outside the appended rows changed: count of changed elements == appended elements
    changed = (kp != kp0).sum().item()
    assert changed <= T * n_kv * hd
    del touched_k


# ------------------------------------------------------------------------------------------------
def _attn_ref(q, K, V, n_heads, n_kv, q_pos):
    """q [Tq, n_heads, hd], K/V [L, n_kv, hd] fp32; query i sees keys 0..q_pos[i]."""
    hd = q.shape[-1]
    grp = n_heads // n_kv
    Kh = K.repeat_interleave(grp, dim=1)
    Vh = V.repeat_interleave(grp, dim=1)
    s = torch.einsum("thd,lhd->htl", q, Kh) / math.sqrt(hd)
    L = K.shape[0]
    mask = torch.arange(L, device=q.device)[None, :] > q_pos[:, None]
    s = s.masked_fill(mask[None], float("-inf"))
    return torch.einsum("htl,lhd->thd", torch.softmax(s, -1), Vh)


DECODE_CASES = [
    # hd, n_heads, n_kv, kv_lens
    (128, 8, 2, [1, 31, 32, 33, 64, 65, 200]),
    (128, 4, 4, [7, 130]),            # group 1
    (128, 16, 2, [2112, 17, 1000]),   # group 8, long
    (64, 8, 2, [1, 63, 64, 65, 500]),
    (64, 32, 4, [300, 2049]),         # TinyLlama grouping (8)
    (64, 4, 2, [90]),                 # group 2
    (128, 32, 8, [8192]),             # Llama-3-8B heads at the context limit, alone: 16 KV splits
    (128, 32, 8, [8191, 8192, 4097, 1, 64]),   # ragged around the limit and the block boundaries
    (128, 8, 1, [3000, 5]),           # Llama-3-70B's per-GPU share at TP 8: one KV head, group 8
]


@pytest.mark.parametrize("use_scratch", [False, True])
@pytest.mark.parametrize("hd,n_heads,n_kv,kv_lens", DECODE_CASES)
def test_attn_decode(kb, cuda, hd, n_heads, n_kv, kv_lens, use_scratch):
    from kafka_b200 import kv_layout
    g = torch.Generator(device="cpu").manual_seed(sum(kv_lens) + hd)
    B = len(kv_lens)
    max_blocks = (max(kv_lens) + 63) // 64 + 1
    n_blocks = B * max_blocks
    perm = torch.randperm(n_blocks, generator=g).view(B, max_blocks).int()
    kp = torch.zeros(n_blocks, n_kv, 64 * hd, dtype=torch.float16, device=cuda)
    vp = torch.zeros_like(kp)
    Ks, Vs = [], []
    for s, L in enumerate(kv_lens):
        K = _f16(torch.randn(L, n_kv, hd, generator=g)).to(cuda)
        V = _f16(torch.randn(L, n_kv, hd, generator=g)).to(cuda)
        kv_layout.write_tokens(kp, perm[s].tolist(), 0, K)
        kv_layout.write_tokens(vp, perm[s].tolist(), 0, V)
        Ks.append(K.float())
        Vs.append(V.float())
    # q lives inside a wider row (as in the fused qkv buffer)
    qkv = _f16(torch.randn(B, (n_heads + 2 * n_kv) * hd, generator=g)).to(cuda)
    lens = torch.tensor(kv_lens, dtype=torch.int32, device=cuda)
    scratch = None
    if use_scratch:
        scratch = torch.empty(kb.attn_decode_scratch_bytes(B, n_heads, hd), dtype=torch.uint8, device=cuda)
    out = kb.attn_decode(qkv, kp, vp, lens, perm.to(cuda), n_heads, n_kv, hd, kv_len_bias=0,
                         max_kv_len=max(kv_lens), scratch=scratch)
    torch.cuda.synchronize()
    for s, L in enumerate(kv_lens):
        q = qkv[s, : n_heads * hd].float().view(1, n_heads, hd)
        ref = _attn_ref(q, Ks[s], Vs[s], n_heads, n_kv, torch.tensor([L - 1], device=cuda))
        # P and the output are rounded to bf16: 2e-2
        torch.testing.assert_close(out[s].float().view(1, n_heads, hd), ref, rtol=2e-2, atol=2e-2)


def test_attn_decode_bias_matches_explicit_length(kb, cuda):
    """kv_len_bias=1 over ctx_lens must equal kv_lens = ctx_lens + 1 (how the engine calls it)."""
    from kafka_b200 import kv_layout
    g = torch.Generator(device="cpu").manual_seed(3)
    hd, n_heads, n_kv = 128, 8, 2
    lens = [40, 129]
    perm = torch.arange(8).view(2, 4).int()
    kp, vp = _alloc_pool(8, n_kv, hd, cuda, 5)
    q = _f16(torch.randn(2, n_heads * hd, generator=g)).to(cuda)
    a = kb.attn_decode(q, kp, vp, torch.tensor(lens, dtype=torch.int32, device=cuda), perm.to(cuda), n_heads, n_kv, hd)
    b = kb.attn_decode(q, kp, vp, torch.tensor([l - 1 for l in lens], dtype=torch.int32, device=cuda), perm.to(cuda),
                       n_heads, n_kv, hd, kv_len_bias=1)
    assert torch.equal(a, b)
    del kv_layout


PREFILL_CASES = [
    # hd, n_heads, n_kv, [(ctx, new), ...]
    (128, 8, 2, [(0, 5), (100, 64), (60, 9)]),
    (128, 8, 2, [(2048, 64)]),
    (128, 4, 4, [(0, 130)]),            # group 1: 64 query tokens per CTA
    (64, 32, 4, [(0, 77), (500, 3)]),   # group 8
    (64, 4, 2, [(63, 2), (0, 1)]),
    (128, 16, 2, [(10, 200)]),
    # the two-query-tiles-per-CTA kernel (head_dim 128, more than 128 query rows in some sequence): Llama-3-8B's
    # grouping with the benchmark's turn, ragged pairs (second tile partly filled / empty / first tile only),
    # several pairs per sequence, a pair whose first tile is fully masked in the last K/V tile
    (128, 32, 8, [(2048, 64)]),
    (128, 32, 8, [(2048, 64), (0, 40), (700, 33), (5, 32), (63, 1), (129, 97)]),
    (128, 8, 2, [(0, 300), (4000, 65)]),
    (128, 8, 8, [(100, 129), (0, 256)]),   # group 1: 128 query tokens per tile
    (128, 16, 2, [(31, 16), (64, 17)]),    # group 8: 16 tokens per tile
    # 128-key tiles: odd and even numbers of KV blocks, a lone block, lengths at the block and tile boundaries
    (128, 8, 2, [(0, 64), (0, 65), (64, 64), (100, 28), (127, 1), (128, 1), (191, 2), (320, 64)]),
]


@pytest.mark.parametrize("variant", ["default", "independent_halves", "tiles_of_64", "two_tiles_per_cta", "two_ctas_per_sm"])
@pytest.mark.parametrize("hd,n_heads,n_kv,seqs", PREFILL_CASES)
def test_attn_prefill(kb, cuda, hd, n_heads, n_kv, seqs, variant, monkeypatch):
    """`variant`: the head_dim-128 tcgen05 kernel exists in four forms, selected per call through the environment
    (csrc/kb_attn_tc.cu): the default (128-key tiles, two MMA-issuing threads) and, behind KB_ATTN_TC64=1, the
    64-key-tile kernels: one tile per CTA, KB_ATTN_TC2 (two query tiles per CTA) and KB_ATTN_TCS (two CTAs per SM)."""
    from kafka_b200 import kv_layout
    for k in ("KB_ATTN_TC64", "KB_ATTN_TC2", "KB_ATTN_TCS", "KB_ATTN_T128H"):
        monkeypatch.delenv(k, raising=False)
    if variant != "default" and hd != 128:
        pytest.skip("head_dim 64 has one kernel (mma.sync)")
    if variant == "independent_halves":
        monkeypatch.setenv("KB_ATTN_T128H", "1")       # 128-key tiles, the two key halves never meet inside the loop
    elif variant != "default":
        monkeypatch.setenv("KB_ATTN_TC64", "1")
    if variant == "two_tiles_per_cta":
        if max(n for _, n in seqs) * (n_heads // n_kv) <= 128:
            pytest.skip("needs more than 128 query rows in a sequence")
        monkeypatch.setenv("KB_ATTN_TC2", "1")
    elif variant == "two_ctas_per_sm":
        monkeypatch.setenv("KB_ATTN_TCS", "1")
    g = torch.Generator(device="cpu").manual_seed(hd + len(seqs) + seqs[0][1])
    B = len(seqs)
    max_blocks = max((c + n + 63) // 64 for c, n in seqs) + 1
    n_blocks = B * max_blocks
    perm = torch.randperm(n_blocks, generator=g).view(B, max_blocks).int()
    kp = torch.zeros(n_blocks, n_kv, 64 * hd, dtype=torch.float16, device=cuda)
    vp = torch.zeros_like(kp)
    T = sum(n for _, n in seqs)
    qkv = _f16(torch.randn(T, (n_heads + 2 * n_kv) * hd, generator=g)).to(cuda)
    refs = []
    off = 0
    for s, (c, n) in enumerate(seqs):
        K = _f16(torch.randn(c + n, n_kv, hd, generator=g)).to(cuda)
        V = _f16(torch.randn(c + n, n_kv, hd, generator=g)).to(cuda)
        kv_layout.write_tokens(kp, perm[s].tolist(), 0, K)
        kv_layout.write_tokens(vp, perm[s].tolist(), 0, V)
        q = qkv[off: off + n, : n_heads * hd].float().view(n, n_heads, hd)
        refs.append(_attn_ref(q, K.float(), V.float(), n_heads, n_kv, torch.arange(c, c + n, device=cuda)))
        off += n
    cu = torch.tensor([0] + list(torch.tensor([n for _, n in seqs]).cumsum(0)), dtype=torch.int32, device=cuda)
    ctx = torch.tensor([c for c, _ in seqs], dtype=torch.int32, device=cuda)
    out = kb.attn_prefill(qkv, kp, vp, cu, ctx, perm.to(cuda), max(n for _, n in seqs), n_heads, n_kv, hd)
    torch.cuda.synchronize()
    ref = torch.cat(refs, 0)
    torch.testing.assert_close(out.float().view(T, n_heads, hd), ref, rtol=2e-2, atol=2e-2)


# ------------------------------------------------------------------------------------------------
def test_sample_greedy_and_masks(kb, cuda):
    g = torch.Generator(device="cpu").manual_seed(11)
    logits = torch.randn(9, 32000, generator=g).to(cuda)
    logits[3, 100] = logits[3, 7] = 50.0  # tie -> lowest id
    out = kb.sample(logits)
    ref = torch.argmax(logits, -1).int()
    ref[3] = 7
    assert torch.equal(out, ref)
    # banned token
    out_b = kb.sample(logits, banned_token=int(ref[0]))
    l0 = logits[0].clone()
    l0[int(ref[0])] = float("-inf")
    assert int(out_b[0]) == int(torch.argmax(l0))
    # allowed range + extras
    out_r = kb.sample(logits, allowed_lo=32, allowed_hi=127, extra_a=10, extra_b=31999)
    allowed = torch.tensor(list(range(32, 127)) + [10, 31999], device=cuda)
    want = allowed[torch.argmax(logits[:, allowed], -1)].int()
    assert torch.equal(out_r, want)


@pytest.mark.parametrize("vocab", [1000, 1003, 128256])
def test_sample_ragged_ranges_and_vector_path(kb, cuda, vocab):
    """The aligned middle of the candidate range is scanned with 16-byte loads, its ragged ends and the extras
    one by one: every (lo, hi) alignment must give the argmax of exactly the allowed ids, greedy and with Gumbel
    noise (which depends on the id only, so masking logits to -inf must equal restricting the range)."""
    g = torch.Generator(device="cpu").manual_seed(vocab)
    n = 7
    logits = torch.randn(n, vocab, generator=g).to(cuda)
    assert torch.equal(kb.sample(logits), torch.argmax(logits, -1).int())
    for lo, hi, ea, eb in [(0, vocab, -1, -1), (1, vocab - 1, -1, -1), (2, 7, 0, -1), (5, 6, -1, vocab - 1),
                           (33, 127, 10, vocab - 2), (3, 4, -1, -1), (vocab - 9, vocab - 2, 1, 2), (4, 8, -1, -1),
                           (6, 262, 300, 1)]:
        ids = list(range(lo, hi)) + [e for e in (ea, eb) if e >= 0]
        allowed = torch.tensor(ids, device=cuda)
        want = allowed[torch.argmax(logits[:, allowed], -1)].int()
        got = kb.sample(logits, allowed_lo=lo, allowed_hi=hi, extra_a=ea, extra_b=eb)
        assert torch.equal(got, want), (lo, hi, ea, eb)
        masked = torch.full_like(logits, float("-inf"))
        masked[:, allowed] = logits[:, allowed]
        a = kb.sample(logits, temperature=0.9, seed=3, step=2, allowed_lo=lo, allowed_hi=hi, extra_a=ea, extra_b=eb)
        b = kb.sample(masked, temperature=0.9, seed=3, step=2)
        assert torch.equal(a, b), (lo, hi, ea, eb)
    banned = int(torch.argmax(logits[0]))
    l0 = logits[0].clone()
    l0[banned] = float("-inf")
    assert int(kb.sample(logits, banned_token=banned)[0]) == int(torch.argmax(l0))


def test_sample_temperature_distribution(kb, cuda):
    """Gumbel-max sampling: deterministic per (seed, step), and frequencies follow softmax(logits / T)."""
    probs = torch.tensor([0.5, 0.25, 0.125, 0.125])
    T = 0.7
    logits = (probs.log() * T).repeat(4096, 1).to(cuda).contiguous()  # softmax(logits / T) == probs
    # vocab must be a multiple of 8 for the engine, not for kb_sample itself
    a = kb.sample(logits, temperature=T, seed=1, step=5)
    b = kb.sample(logits, temperature=T, seed=1, step=5)
    c = kb.sample(logits, temperature=T, seed=1, step=6)
    assert torch.equal(a, b)
    assert not torch.equal(a, c)
    freq = torch.bincount(a.long(), minlength=4).float().cpu() / 4096
    # 4096 draws: std of a frequency <= sqrt(.25/4096) = 0.0078; allow 5 sigma
    assert torch.all((freq - probs).abs() < 0.04), freq


@pytest.mark.parametrize("n_slices", [2, 8])
def test_sample_by_vocabulary_slices_picks_the_unsharded_token(kb, cuda, n_slices):
    """kb_sample_slice (one rank's part of vocabulary-sharded sampling under tensor parallelism): the largest of
    the slices' (value, id) pairs — ties to the lowest id, the rule tp_argmax_kernel applies across the ranks —
    is exactly kb_sample's token on the full row: greedy, with Gumbel noise, with a candidate range + extras,
    a banned id, and rows full of exact ties. The cross-GPU exchange itself is tests/test_gpu_tp.py."""
    S, V = 257, 4096
    nv = V // n_slices
    g = torch.Generator(device="cpu").manual_seed(17 + n_slices)
    cases = [dict(), dict(temperature=0.8, seed=3, step=9), dict(allowed_lo=100, allowed_hi=360, extra_a=4000, extra_b=7),
             dict(temperature=1.3, seed=1, step=2, allowed_lo=nv - 3, allowed_hi=nv + 5, banned_token=nv),
             dict(allowed_lo=10, allowed_hi=20), dict(banned_token=0)]
    for it, kw in enumerate(cases * 2):
        logits = torch.randn(S, V, generator=g)
        if it % 2 == 1:
            logits = (logits * 2).round() / 2
        full = logits.to(cuda)
        want = kb.sample(full, **kw).cpu()
        best_v = torch.full((S,), float("-inf"))
        best_i = torch.full((S,), 0x7fffffff, dtype=torch.int64)
        for r in range(n_slices):
            pairs = kb.sample_slice(full[:, r * nv:(r + 1) * nv].contiguous(), V, r * nv, **kw).cpu()
            val = pairs[:, 0].contiguous().view(torch.float32)
            idx = pairs[:, 1].long()
            assert bool(((idx == 0x7fffffff) | ((idx >= r * nv) & (idx < (r + 1) * nv))).all())
            better = (val > best_v) | ((val == best_v) & (idx < best_i))
            best_v = torch.where(better, val, best_v)
            best_i = torch.where(better, idx, best_i)
        got = torch.where(best_i == 0x7fffffff, torch.zeros_like(best_i), best_i).int()
        assert torch.equal(got, want), (kw, it)


def _nucleus(logits_row, top_p, T=1.0):
    """The ids nucleus sampling may return: the most probable ones whose probabilities reach top_p (torch restatement)."""
    p = torch.softmax(logits_row.double() / T, -1)
    order = torch.argsort(p, descending=True)
    cum = torch.cumsum(p[order], 0)
    n = int((cum < top_p).sum()) + 1
    return set(order[:n].tolist()), p, float(p[order[n - 1]])


def test_sample_top_p_known_distribution(kb, cuda):
    """Nucleus sampling (the `top_p` of the reference's provider API): only the smallest set of most probable tokens
    whose mass reaches top_p is sampled, with the probabilities renormalised; top_p = 0 / 1 leave the sampler as it
    was (bit-equal), a tiny top_p is the arg-max whatever the seed, and a seed reproduces."""
    probs = torch.tensor([0.4, 0.3, 0.15, 0.1, 0.03, 0.01, 0.005, 0.005])
    logits = probs.log().repeat(8192, 1).to(cuda).contiguous()
    a = kb.sample(logits, temperature=1.0, seed=4, step=1, top_p=0.8)
    assert torch.equal(a, kb.sample(logits, temperature=1.0, seed=4, step=1, top_p=0.8))
    assert int(a.max()) <= 2                                    # {0.4, 0.3, 0.15}: 0.85 >= 0.8, two of them only 0.7
    freq = torch.bincount(a.long(), minlength=8).float().cpu() / 8192
    want = torch.tensor([0.4, 0.3, 0.15, 0, 0, 0, 0, 0]) / 0.85
    assert torch.all((freq - want).abs() < 0.03), freq
    plain = kb.sample(logits, temperature=1.0, seed=4, step=1)
    assert torch.equal(plain, kb.sample(logits, temperature=1.0, seed=4, step=1, top_p=1.0))
    assert torch.equal(plain, kb.sample(logits, temperature=1.0, seed=4, step=1, top_p=0.0))
    assert int(plain.max()) > 2                                 # without the cut the tail is sampled
    for seed in (1, 2, 3):
        assert bool((kb.sample(logits, temperature=1.0, seed=seed, step=7, top_p=0.05) == 0).all())
    # temperature reshapes the distribution before the cut: at T = 0.5 p ~ probs^2 -> {0.4, 0.3} hold 0.87 of the mass
    hot = kb.sample(logits * 1.0, temperature=0.5, seed=9, step=3, top_p=0.8)
    assert int(hot.max()) <= 1
    # greedy ignores top_p
    assert bool((kb.sample(logits, temperature=0.0, top_p=0.5) == 0).all())
    # near-ties at the top (1 % apart: the same coarse histogram bin, resolved by the second level): the cut still
    # falls between them
    close = torch.tensor([0.300, 0.297, 0.2, 0.1, 0.05, 0.03, 0.013, 0.01]).log().repeat(2048, 1).to(cuda).contiguous()
    for seed in (1, 2):
        assert bool((kb.sample(close, temperature=1.0, seed=seed, step=1, top_p=0.2) == 0).all())
        assert int(kb.sample(close, temperature=1.0, seed=seed, step=1, top_p=0.5).max()) == 1
    # per-sequence values: row 0 off, row 1 cut to the arg-max
    two = logits[:2].contiguous()
    seq_tp = torch.tensor([0.0, 0.05], device=cuda)
    outs = torch.stack([kb.sample(two, temperature=1.0, seed=s, step=0, seq_top_p=seq_tp) for s in range(200)])
    assert int(outs[:, 1].max()) == 0 and int(outs[:, 0].max()) > 0


@pytest.mark.parametrize("vocab,lo,hi,banned", [(128256, 0, 0, -1), (128256, 1000, 90000, -2), (4096, 0, 0, -1), (520, 3, 517, -2)])
def test_sample_top_p_on_wide_rows_matches_the_sorted_definition(kb, cuda, vocab, lo, hi, banned):
    """Full-vocabulary rows with a peaked distribution: every sampled id lies in the nucleus computed by sorting (or
    within 1e-4 in log-probability of its cut — the kernel resolves the cut to ~3e-5), and ids outside are never
    produced over many draws; with a candidate range and a banned id the nucleus is that of the remaining candidates."""
    g = torch.Generator(device="cpu").manual_seed(vocab + lo)
    S, T, top_p = 16, 0.9, 0.9
    logits = (torch.randn(S, vocab, generator=g) * 3.0)
    dev_logits = logits.to(cuda)
    kw = {}
    masked = logits.clone()
    if hi > lo:
        kw.update(allowed_lo=lo, allowed_hi=hi)
        masked[:, :lo] = float("-inf")
        masked[:, hi:] = float("-inf")
    if banned == -2:                                            # ban each row's favourite... of row 0 (one id for the call)
        banned = int(masked[0].argmax())
    if banned >= 0:
        kw.update(banned_token=banned)
        masked[:, banned] = float("-inf")
    allowed, slack = [], []
    for r in range(S):
        keep, p, p_cut = _nucleus(masked[r], top_p, T)
        allowed.append(keep)
        near = ((p.log() - torch.log(torch.tensor(p_cut, dtype=torch.float64))).abs() < 1e-4).nonzero().flatten().tolist()
        slack.append(set(near))
    seen = [set() for _ in range(S)]
    for step in range(300):
        out = kb.sample(dev_logits, temperature=T, seed=11, step=step, top_p=top_p, **kw).cpu().tolist()
        for r, t in enumerate(out):
            assert t in allowed[r] or t in slack[r], (r, t, step, len(allowed[r]))
            seen[r].add(t)
    # the draws really spread over the nucleus (not just its head)
    assert all(len(s) >= min(len(a), 20) // 2 for s, a in zip(seen, allowed))
