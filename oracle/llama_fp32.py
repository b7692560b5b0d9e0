"""ORACLE — fp32 CPU eager forward of the Llama architecture.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import this module; the product (``kafka-llm-service_b200/``) never does, and fails loudly
when its CUDA library is missing instead of falling back to anything here.

PARITY UNPINNED.  The reference (egrigokhan/kafka-llm-service) holds no arithmetic for this path: its
only provider forwards the conversation to a remote gateway
(/root/reference/src/llm/portkey.py:379 and :616) and it ships no tests, golden vectors or fixtures
for completions.  BASELINE.json's north_star therefore *defines* the oracle as an
fp32 PyTorch CPU eager forward of the same random-init architecture; this file is that restatement of
the published Llama algorithm (RMSNorm -> GQA attention with rotate-half RoPE -> SwiGLU MLP).  What
pins it instead: ``tests/test_oracle.py`` checks it against ``transformers.LlamaForCausalLM`` (an
independent implementation present in the image) on random-init weights, and the generating script
``tests/golden/make_golden.py`` freezes its outputs as fixtures.

Shapes follow kafka_b200.model_config.LlamaConfig; weights follow kafka_b200.weights.generate_weights
(canonical, unfused names).
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch


def rmsnorm(x: torch.Tensor, w: torch.Tensor, eps: float) -> torch.Tensor:
    var = x.pow(2).mean(-1, keepdim=True)
    return x * torch.rsqrt(var + eps) * w


def rope_cos_sin(positions: torch.Tensor, head_dim: int, theta: float, inv_freq=None):
    """`inv_freq`: the head_dim / 2 frequencies when they are not the plain theta^(-2i/d) (scaled RoPE)."""
    if inv_freq is None:
        inv_freq = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.float32) / head_dim))
    freqs = positions.to(torch.float32)[:, None] * inv_freq[None, :]
    return freqs.cos(), freqs.sin()  # [T, hd/2]


def apply_rope(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
    """x: [T, n_heads, hd]; HF rotate-half convention (pairs i, i + hd/2)."""
    half = x.shape[-1] // 2
    x1, x2 = x[..., :half], x[..., half:]
    c, s = cos[:, None, :], sin[:, None, :]
    return torch.cat([x1 * c - x2 * s, x2 * c + x1 * s], dim=-1)


class KVCache:
    """Contiguous per-sequence cache, fp32: per layer a [capacity, n_kv_heads, head_dim] buffer of
    which the first ``length`` rows are valid (amortised-doubling append, so a decode step costs
    O(1) copies like a real server's cache would)."""

    def __init__(self, n_layers: int):
        self._k: List[Optional[torch.Tensor]] = [None] * n_layers
        self._v: List[Optional[torch.Tensor]] = [None] * n_layers
        self._len = [0] * n_layers

    @property
    def length(self) -> int:
        return self._len[0]

    def k(self, i: int) -> torch.Tensor:
        return self._k[i][: self._len[i]]

    def v(self, i: int) -> torch.Tensor:
        return self._v[i][: self._len[i]]

    def append(self, i: int, k: torch.Tensor, v: torch.Tensor) -> None:
        n, need = k.shape[0], self._len[i] + k.shape[0]
        if self._k[i] is None or need > self._k[i].shape[0]:
            cap = max(need, 2 * (0 if self._k[i] is None else self._k[i].shape[0]), 64)
            nk = torch.empty((cap,) + tuple(k.shape[1:]), dtype=k.dtype)
            nv = torch.empty_like(nk)
            if self._len[i]:
                nk[: self._len[i]] = self._k[i][: self._len[i]]
                nv[: self._len[i]] = self._v[i][: self._len[i]]
            self._k[i], self._v[i] = nk, nv
        self._k[i][self._len[i]: need] = k
        self._v[i][self._len[i]: need] = v
        self._len[i] = need

    def truncate(self, n: int) -> None:
        self._len = [min(n, l) for l in self._len]


class OracleLlama:
    """fp32 forward. ``weights`` are the canonical bf16-representable tensors, widened to fp32."""

    def __init__(self, cfg, weights: Dict[str, torch.Tensor], allreduce=None, tp_rank: int = 0):
        """`allreduce` / `tp_rank`: tensor-parallel restatement. With a shard's local cfg and weights
        (kafka_b200.tp.local_config / shard_weights) the two row-parallel projections yield partial sums;
        `allreduce(t)` must return their sum over the ranks. Rank 0 folds the residual in, exactly like
        the engine (kb_set_tensor_parallel)."""
        self.cfg = cfg
        self.w = {k: v.detach().to("cpu", torch.float32) for k, v in weights.items()}
        self.allreduce = allreduce
        self.tp_rank = tp_rank
        # scaled RoPE (Llama-3.1's rule): the frequencies come from the config; plain RoPE computes them in place
        self.inv_freq = None
        if getattr(cfg, "rope_llama3", None) is not None:
            from kafka_b200.model_config import inv_freq
            self.inv_freq = torch.tensor(inv_freq(cfg), dtype=torch.float32)

    def _residual(self, x: torch.Tensor, delta: torch.Tensor) -> torch.Tensor:
        if self.allreduce is None:
            return x + delta
        return self.allreduce(delta + x if self.tp_rank == 0 else delta)

    def new_cache(self) -> KVCache:
        return KVCache(self.cfg.n_layers)

    @torch.no_grad()
    def forward(self, tokens: torch.Tensor, cache: KVCache, all_logits: bool = False,
                taps: Optional[Dict[str, torch.Tensor]] = None) -> torch.Tensor:
        """Append ``tokens`` [T] to ``cache`` and return logits of the last position [V]
        (or of every new position [T, V] when all_logits)."""
        cfg, w = self.cfg, self.w
        T = tokens.numel()
        past = cache.length
        pos = torch.arange(past, past + T)
        cos, sin = rope_cos_sin(pos, cfg.head_dim, cfg.rope_theta, self.inv_freq)
        x = w["embed"][tokens.long()]
        scale = cfg.head_dim ** -0.5
        for i in range(cfg.n_layers):
            p = f"l{i}."
            h = rmsnorm(x, w[p + "attn_norm"], cfg.rms_eps)
            q = (h @ w[p + "wq"].T).view(T, cfg.n_heads, cfg.head_dim)
            k = (h @ w[p + "wk"].T).view(T, cfg.n_kv_heads, cfg.head_dim)
            v = (h @ w[p + "wv"].T).view(T, cfg.n_kv_heads, cfg.head_dim)
            q = apply_rope(q, cos, sin)
            k = apply_rope(k, cos, sin)
            cache.append(i, k, v)
            K = cache.k(i).repeat_interleave(cfg.group, dim=1)  # [L, n_heads, hd]
            V = cache.v(i).repeat_interleave(cfg.group, dim=1)
            s = torch.einsum("thd,lhd->htl", q, K) * scale      # [n_heads, T, L]
            L = K.shape[0]
            causal = torch.arange(L)[None, :] > (past + torch.arange(T))[:, None]
            s = s.masked_fill(causal[None], float("-inf"))
            a = torch.softmax(s, dim=-1)
            o = torch.einsum("htl,lhd->thd", a, V).reshape(T, cfg.q_dim)
            if taps is not None and i == 0:
                taps["q0"], taps["attn0"] = q.clone(), o.clone()
            x = self._residual(x, o @ w[p + "wo"].T)
            h = rmsnorm(x, w[p + "mlp_norm"], cfg.rms_eps)
            g = h @ w[p + "w_gate"].T
            u = h @ w[p + "w_up"].T
            x = self._residual(x, (torch.nn.functional.silu(g) * u) @ w[p + "w_down"].T)
        xl = x if all_logits else x[-1:]
        logits = rmsnorm(xl, w["final_norm"], cfg.rms_eps) @ w["lm_head"].T
        return logits if all_logits else logits[0]

    @torch.no_grad()
    def forward_packed(self, seqs: List[torch.Tensor], caches: List[KVCache]) -> torch.Tensor:
        """``forward`` for several sequences at once: the new tokens of all sequences are packed for the
        projections (one [sum T, H] x [H, N] GEMM each, so the host cores run a config-2-sized turn in
        minutes instead of tens of minutes), attention runs per sequence against its own cache. Same
        arithmetic per row as ``forward``; returns the last position's logits of every sequence [n, V]."""
        cfg, w = self.cfg, self.w
        lens = [int(s.numel()) for s in seqs]
        offs = [0]
        for n in lens:
            offs.append(offs[-1] + n)
        pasts = [c.length for c in caches]
        pos = torch.cat([torch.arange(p, p + n) for p, n in zip(pasts, lens)])
        cos, sin = rope_cos_sin(pos, cfg.head_dim, cfg.rope_theta, self.inv_freq)
        x = w["embed"][torch.cat([s.long() for s in seqs])]
        T = x.shape[0]
        scale = cfg.head_dim ** -0.5
        for i in range(cfg.n_layers):
            p = f"l{i}."
            h = rmsnorm(x, w[p + "attn_norm"], cfg.rms_eps)
            q = apply_rope((h @ w[p + "wq"].T).view(T, cfg.n_heads, cfg.head_dim), cos, sin)
            k = apply_rope((h @ w[p + "wk"].T).view(T, cfg.n_kv_heads, cfg.head_dim), cos, sin)
            v = (h @ w[p + "wv"].T).view(T, cfg.n_kv_heads, cfg.head_dim)
            o = torch.empty(T, cfg.q_dim)
            for b, c in enumerate(caches):
                lo, hi = offs[b], offs[b + 1]
                n = hi - lo
                c.append(i, k[lo:hi], v[lo:hi])
                K, V = c.k(i), c.v(i)                                      # [L, kvh, hd]
                sc = torch.einsum("tkgd,lkd->kgtl", q[lo:hi].view(n, cfg.n_kv_heads, cfg.group, cfg.head_dim),
                                  K) * scale
                causal = torch.arange(K.shape[0])[None, :] > (pasts[b] + torch.arange(n))[:, None]
                a = torch.softmax(sc.masked_fill(causal[None, None], float("-inf")), dim=-1)
                o[lo:hi] = torch.einsum("kgtl,lkd->tkgd", a, V).reshape(n, cfg.q_dim)
            x = self._residual(x, o @ w[p + "wo"].T)
            h = rmsnorm(x, w[p + "mlp_norm"], cfg.rms_eps)
            x = self._residual(x, (torch.nn.functional.silu(h @ w[p + "w_gate"].T) * (h @ w[p + "w_up"].T))
                               @ w[p + "w_down"].T)
        last = x[torch.tensor([o_ - 1 for o_ in offs[1:]])]
        return rmsnorm(last, w["final_norm"], cfg.rms_eps) @ w["lm_head"].T

    @torch.no_grad()
    def decode_batch(self, tokens: torch.Tensor, caches: List[KVCache],
                     taps: Optional[Dict[str, torch.Tensor]] = None) -> torch.Tensor:
        """One decode step for B sequences at once (projections batched over B, attention per
        sequence). Same arithmetic as ``forward`` with T=1 per sequence; returns logits [B, V].
        Used by bench.py's CPU baseline so the host cores see one [B,H]x[H,N] GEMM per projection."""
        cfg, w = self.cfg, self.w
        B = tokens.numel()
        pos = torch.tensor([c.length for c in caches])
        cos, sin = rope_cos_sin(pos, cfg.head_dim, cfg.rope_theta, self.inv_freq)
        x = w["embed"][tokens.long()]
        scale = cfg.head_dim ** -0.5
        padded = B > 1 and int(pos.max()) < 1024          # short contexts: one padded batch (see below)
        pad = torch.arange(int(pos.max()) + 1)[None, :] > pos[:, None]
        for i in range(cfg.n_layers):
            p = f"l{i}."
            h = rmsnorm(x, w[p + "attn_norm"], cfg.rms_eps)
            q = apply_rope((h @ w[p + "wq"].T).view(B, cfg.n_heads, cfg.head_dim), cos, sin)
            k = apply_rope((h @ w[p + "wk"].T).view(B, cfg.n_kv_heads, cfg.head_dim), cos, sin)
            v = (h @ w[p + "wv"].T).view(B, cfg.n_kv_heads, cfg.head_dim)
            for b, c in enumerate(caches):
                c.append(i, k[b:b + 1], v[b:b + 1])
            if padded:
                # attention of all B sequences as one padded batch (B tiny per-sequence einsums cost more
                # in dispatch than in arithmetic): padded key positions get -inf and contribute an exact 0
                Kp = torch.nn.utils.rnn.pad_sequence([c.k(i) for c in caches], batch_first=True)   # [B, Lmax, kvh, hd]
                Vp = torch.nn.utils.rnn.pad_sequence([c.v(i) for c in caches], batch_first=True)
                s = torch.einsum("bkgd,blkd->bkgl", q.view(B, cfg.n_kv_heads, cfg.group, cfg.head_dim), Kp) * scale
                s = s.masked_fill(pad[:, None, None, :], float("-inf"))
                o = torch.einsum("bkgl,blkd->bkgd", torch.softmax(s, -1), Vp).reshape(B, cfg.n_heads, cfg.head_dim)
            else:
                # long contexts: per sequence, straight from each cache (no padded copy of the K/V)
                o = torch.empty(B, cfg.n_heads, cfg.head_dim)
                for b, c in enumerate(caches):
                    qb = q[b].view(cfg.n_kv_heads, cfg.group, cfg.head_dim)
                    s = torch.einsum("kgd,lkd->kgl", qb, c.k(i)) * scale
                    o[b] = torch.einsum("kgl,lkd->kgd", torch.softmax(s, -1), c.v(i)).reshape(cfg.n_heads, cfg.head_dim)
            x = self._residual(x, o.view(B, cfg.q_dim) @ w[p + "wo"].T)
            h = rmsnorm(x, w[p + "mlp_norm"], cfg.rms_eps)
            inner = torch.nn.functional.silu(h @ w[p + "w_gate"].T) * (h @ w[p + "w_up"].T)
            if taps is not None and i == cfg.n_layers - 1:     # the LAST layer's intermediates (tools/stage_errors.py)
                taps.update(q=q.reshape(B, -1).clone(), k=k.reshape(B, -1).clone(), v=v.reshape(B, -1).clone(),
                            attn=o.reshape(B, -1).clone(), x_mid=x.clone(), xn_mlp=h.clone(), inner=inner.clone())
            x = self._residual(x, inner @ w[p + "w_down"].T)
        xl = rmsnorm(x, w["final_norm"], cfg.rms_eps)
        if taps is not None:
            taps.update(x_final=x.clone(), xl=xl.clone())
        return xl @ w["lm_head"].T

    @torch.no_grad()
    def greedy(self, prompt: torch.Tensor, n_new: int, cache: Optional[KVCache] = None,
               allowed: Optional[torch.Tensor] = None, eos: Optional[int] = None):
        """Greedy decode; returns (tokens, logits per step [n, V]). Ties -> lowest id (like kb_sample)."""
        cache = cache or self.new_cache()
        toks: List[int] = []
        lgs: List[torch.Tensor] = []
        cur = prompt
        for _ in range(n_new):
            lg = self.forward(cur, cache)
            lgs.append(lg)
            t = masked_argmax(lg, allowed)
            toks.append(t)
            if eos is not None and t == eos:
                break
            cur = torch.tensor([t])
        return toks, torch.stack(lgs)


def masked_argmax(logits: torch.Tensor, allowed: Optional[torch.Tensor]) -> int:
    if allowed is None:
        return int(torch.argmax(logits))
    sub = logits[allowed]
    best = sub.max()
    return int(allowed[sub == best].min())
