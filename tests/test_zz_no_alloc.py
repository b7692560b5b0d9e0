"""Decode attention without caller-provided scratch (kb_attn_decode_static_scratch).

Who needs it: a host of this library that receives its input and output buffers from somewhere else
and has no device allocator of its own (a foreign-language binding on IPC / pool memory, a process
with pre-partitioned HBM, a loop that cannot tolerate cudaMalloc's device synchronisation). Such a
caller cannot hand kb_attn_decode the split-KV scratch it wants, so the library keeps the partials in
a fixed 2 MiB array of its own.

What is checked, through the C ABI on the GPU: the new entry point gives the same attention output as
kb_attn_decode with caller scratch AND as a plain fp32 statement of the attention (softmax(q K^T / sqrt(d)) V
per query head, written out in torch below), on two calls in a row whose inputs differ (the second call
finds the first call's partials still in the shared array and must not be affected by them), and a shape
whose partials do not fit is refused with a clear message rather than computed differently.

Collected last (file name) so that, under `pytest -x`, every other test has already run.
Tolerance: rtol = atol = 2e-2 on bf16-rounded outputs, the same as tests/test_gpu_kernels.py.
"""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _fp32_attention(q, Ks, Vs, n_heads, n_kv, hd):
    """One query token per sequence over its whole K/V, in fp32: out[s, h] = softmax(q_h K_h^T / sqrt(hd)) V_h."""
    out = []
    for s, (K, V) in enumerate(zip(Ks, Vs)):
        qs = q[s].float().view(n_heads, hd)
        Kh = K.float().repeat_interleave(n_heads // n_kv, dim=1)      # [L, n_heads, hd]
        Vh = V.float().repeat_interleave(n_heads // n_kv, dim=1)
        p = torch.softmax(torch.einsum("hd,lhd->hl", qs, Kh) / math.sqrt(hd), dim=-1)
        out.append(torch.einsum("hl,lhd->hd", p, Vh).reshape(n_heads * hd))
    return torch.stack(out)


def _case(cuda, seed, kv_lens, hd=128, n_heads=32, n_kv=8):
    from kafka_b200 import kv_layout
    g = torch.Generator(device="cpu").manual_seed(seed)
    B = len(kv_lens)
    max_blocks = (max(kv_lens) + 63) // 64 + 1
    perm = torch.randperm(B * max_blocks, generator=g).view(B, max_blocks).int()
    kp = torch.zeros(B * max_blocks, n_kv, 64 * hd, dtype=torch.float16, device=cuda)
    vp = torch.zeros_like(kp)
    Ks, Vs = [], []
    for s, L in enumerate(kv_lens):
        Ks.append(torch.randn(L, n_kv, hd, generator=g).half().to(cuda))
        Vs.append(torch.randn(L, n_kv, hd, generator=g).half().to(cuda))
        kv_layout.write_tokens(kp, perm[s].tolist(), 0, Ks[-1])
        kv_layout.write_tokens(vp, perm[s].tolist(), 0, Vs[-1])
    q = torch.randn(B, n_heads * hd, generator=g).half().to(cuda)
    lens = torch.tensor(kv_lens, dtype=torch.int32, device=cuda)
    ref = _fp32_attention(q, Ks, Vs, n_heads, n_kv, hd)
    return q, kp, vp, lens, perm.to(cuda), ref


def test_static_scratch_matches_caller_scratch_twice_in_a_row(cuda):
    from kafka_b200 import cabi
    hd, n_heads, n_kv = 128, 32, 8
    assert cabi.load().kb_static_scratch_capacity() == 2 * 1024 * 1024
    # small batches over long histories: the shapes for which the kernel splits the KV range
    for seed, kv_lens in ((1, [2112, 700, 1500, 64]), (2, [900, 3000])):
        q, kp, vp, lens, bt, ref = _case(cuda, seed, kv_lens, hd, n_heads, n_kv)
        scratch = torch.empty(cabi.attn_decode_scratch_bytes(len(kv_lens), n_heads, hd), dtype=torch.uint8,
                              device=cuda)
        want = cabi.attn_decode(q, kp, vp, lens, bt, n_heads, n_kv, hd, max_kv_len=max(kv_lens), scratch=scratch)
        out = torch.empty_like(want)  # the caller's output buffer; the call itself allocates nothing
        cabi.attn_decode_static_scratch(q, kp, vp, lens, bt, n_heads, n_kv, hd, max_kv_len=max(kv_lens), out=out)
        torch.cuda.synchronize()
        torch.testing.assert_close(out.float(), want.float(), rtol=2e-2, atol=2e-2)
        # ... and both against the fp32 statement of the op (P and the output are rounded to bf16: 2e-2)
        torch.testing.assert_close(out.float(), ref, rtol=2e-2, atol=2e-2)
        torch.testing.assert_close(want.float(), ref, rtol=2e-2, atol=2e-2)


def test_static_scratch_large_batch_needs_no_partials(cuda):
    """A batch that fills the GPU on its own is not split, so it fits whatever its size."""
    from kafka_b200 import cabi
    hd, n_heads, n_kv = 128, 32, 8
    kv_lens = [130 + 3 * i for i in range(48)]  # 48 x 8 KV heads >= 2 x 148 SMs
    q, kp, vp, lens, bt, ref = _case(cuda, 3, kv_lens, hd, n_heads, n_kv)
    want = cabi.attn_decode(q, kp, vp, lens, bt, n_heads, n_kv, hd, max_kv_len=max(kv_lens))
    out = cabi.attn_decode_static_scratch(q, kp, vp, lens, bt, n_heads, n_kv, hd, max_kv_len=max(kv_lens))
    torch.cuda.synchronize()
    torch.testing.assert_close(out.float(), want.float(), rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(out.float(), ref, rtol=2e-2, atol=2e-2)


def test_static_scratch_rejects_oversize(cuda):
    from kafka_b200 import cabi
    hd, n_heads, n_kv = 128, 64, 8           # 16 sequences x 64 heads x 5 splits x 520 B > 2 MiB
    kv_lens = [2000] * 16
    q, kp, vp, lens, bt, _ = _case(cuda, 4, kv_lens, hd, n_heads, n_kv)
    with pytest.raises(cabi.KafkaB200Error, match="more than the built-in"):
        cabi.attn_decode_static_scratch(q, kp, vp, lens, bt, n_heads, n_kv, hd, max_kv_len=2000)
