"""Tensor parallelism for models that do not fit one GPU (BASELINE config 5: Llama-3-70B over 8 B200).

Megatron-style sharding, one process per GPU:

    wq / wk / wv, w_gate / w_up   column-parallel: each rank owns n_heads/tp query heads, n_kv_heads/tp
                                  KV heads and ffn/tp MLP columns (rows of the [out, in] weight)
    wo, w_down                    row-parallel: each rank owns the matching K-slice; the partial sums are
                                  added over the ranks — 2 collectives over [T, hidden] per layer
    embed, norms                  replicated
    lm_head                       sharded by vocabulary rows when the peer region is in use: every rank computes
                                  its slice of the logits and its local winner, the ranks' (value, id) pairs are
                                  compared through the region (kb_set_lm_head_shard) — every rank ends with the
                                  token the unsharded sampler picks; replicated with the NCCL callback
    KV cache                      sharded by KV head: each rank stores only its heads' blocks

The engine is created with the rank's LOCAL shapes. The sums over ranks are done by our own kernels
over peer memory (csrc/kb_collective.cu; include/kafka_b200.h, kb_set_tensor_parallel_peers): this
module only provides the symmetric region they work on (`PeerRegion`: torch's symmetric-memory
rendezvous for allocation, peer mappings and the NVSwitch multicast mapping). Default: fp16 partial
sums and ONE kernel per projection that reduces, adds to the residual, RMS-norms and gathers. The
NCCL all-reduce through torch.distributed (`TPGroup`, kb_set_tensor_parallel) is kept as the
alternative. All ranks issue identical kb_prefill / kb_decode calls (SPMD).

The reference has no counterpart: its models run behind a gateway (/root/reference/src/llm/portkey.py:379).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import replace
from typing import Dict, Optional

import torch

from . import cabi
from .model_config import LlamaConfig
from .weights import GEMM_WEIGHT_DTYPE, _randn, fuse_for_engine


def local_config(cfg: LlamaConfig, tp: int) -> LlamaConfig:
    """Shapes one rank's engine is built with."""
    if tp == 1:
        return cfg
    if cfg.n_heads % tp or cfg.n_kv_heads % tp or cfg.ffn % (64 * tp):
        raise ValueError(f"{cfg.name}: heads {cfg.n_heads}/{cfg.n_kv_heads} and ffn {cfg.ffn} must divide by tp={tp} "
                         "(ffn/tp a multiple of 64)")
    return replace(cfg, name=f"{cfg.name}/tp{tp}", n_heads=cfg.n_heads // tp, n_kv_heads=cfg.n_kv_heads // tp,
                   ffn=cfg.ffn // tp)


def lm_head_slice(cfg: LlamaConfig, rank: int, tp: int):
    """(first row, rows) of the vocabulary slice rank `rank` holds, or None when the vocabulary does not cut into
    `tp` slices of a multiple of 8 rows."""
    if tp <= 1 or cfg.vocab % (8 * tp):
        return None
    n = cfg.vocab // tp
    return rank * n, n


def shard_weights(cfg: LlamaConfig, w: Dict[str, torch.Tensor], rank: int, tp: int,
                  shard_lm_head: bool = False) -> Dict[str, torch.Tensor]:
    """Canonical (unfused) weights of the full model -> the canonical weights of rank `rank`'s shard."""
    if tp == 1:
        return w
    loc = local_config(cfg, tp)
    q0, q1 = rank * loc.q_dim, (rank + 1) * loc.q_dim
    k0, k1 = rank * loc.kv_dim, (rank + 1) * loc.kv_dim
    f0, f1 = rank * loc.ffn, (rank + 1) * loc.ffn
    out = {k: w[k] for k in ("embed", "final_norm", "lm_head")}
    sl = lm_head_slice(cfg, rank, tp) if shard_lm_head else None
    if sl is not None:
        out["lm_head"] = w["lm_head"][sl[0]:sl[0] + sl[1]]
    for i in range(cfg.n_layers):
        p = f"l{i}."
        out[p + "attn_norm"], out[p + "mlp_norm"] = w[p + "attn_norm"], w[p + "mlp_norm"]
        out[p + "wq"] = w[p + "wq"][q0:q1]
        out[p + "wk"] = w[p + "wk"][k0:k1]
        out[p + "wv"] = w[p + "wv"][k0:k1]
        out[p + "wo"] = w[p + "wo"][:, q0:q1].contiguous()
        out[p + "w_gate"] = w[p + "w_gate"][f0:f1]
        out[p + "w_up"] = w[p + "w_up"][f0:f1]
        out[p + "w_down"] = w[p + "w_down"][:, f0:f1].contiguous()
    return out


def generate_sharded_on_device(cfg: LlamaConfig, seed: int, rank: int, tp: int, device,
                               shard_lm_head: bool = False) -> Dict[str, torch.Tensor]:
    """Benchmark path for shapes too large for the host: every rank draws the FULL tensor of a layer on
    its GPU from the same per-tensor seed, keeps its slice and frees the rest — ranks agree on the model
    without any transfer."""
    dev = str(device)
    loc = local_config(cfg, tp)
    H, F, f16 = cfg.hidden, cfg.ffn, GEMM_WEIGHT_DTYPE
    out: Dict[str, torch.Tensor] = {
        "embed": _randn((cfg.vocab, H), seed, "embed", 1.0, dev),
        "final_norm": _randn((H,), seed, "final_norm", 0.1, dev, mean=1.0),
        "lm_head": _randn((cfg.vocab, H), seed, "lm_head", 1.0 / (H ** 0.5), dev).to(f16),
    }
    sl = lm_head_slice(cfg, rank, tp) if shard_lm_head else None
    if sl is not None:
        out["lm_head"] = out["lm_head"][sl[0]:sl[0] + sl[1]].clone()
    q0, k0, f0 = rank * loc.q_dim, rank * loc.kv_dim, rank * loc.ffn
    for i in range(cfg.n_layers):
        p = f"l{i}."
        out[p + "attn_norm"] = _randn((H,), seed, p + "attn_norm", 0.1, dev, mean=1.0)
        out[p + "mlp_norm"] = _randn((H,), seed, p + "mlp_norm", 0.1, dev, mean=1.0)
        wq = _randn((cfg.q_dim, H), seed, p + "wq", 0.02, dev)[q0:q0 + loc.q_dim]
        wk = _randn((cfg.kv_dim, H), seed, p + "wk", 0.02, dev)[k0:k0 + loc.kv_dim]
        wv = _randn((cfg.kv_dim, H), seed, p + "wv", 0.02, dev)[k0:k0 + loc.kv_dim]
        out[p + "wqkv"] = torch.cat([wq, wk, wv], 0).to(f16).contiguous()
        out[p + "wo"] = _randn((H, cfg.q_dim), seed, p + "wo", 0.02, dev)[:, q0:q0 + loc.q_dim].to(f16).contiguous()
        g = _randn((F, H), seed, p + "w_gate", 0.02, dev)[f0:f0 + loc.ffn]
        u = _randn((F, H), seed, p + "w_up", 0.02, dev)[f0:f0 + loc.ffn]
        out[p + "w_gate_up"] = torch.stack([g, u], 1).reshape(2 * loc.ffn, H).to(f16).contiguous()
        out[p + "w_down"] = _randn((H, F), seed, p + "w_down", 0.02, dev)[:, f0:f0 + loc.ffn].to(f16).contiguous()
        del wq, wk, wv, g, u
    return out


class TPGroup:
    """The all-reduce the engine calls back into: NCCL (torch.distributed) on the current stream."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.dist = dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.size = dist.get_world_size(group)
        self.calls = 0
        self.bytes = 0
        self._views: Dict[int, torch.Tensor] = {}
        self._bases = []
        self._cb = cabi.ALLREDUCE_FN(self._callback)   # keep the ctypes thunk alive

    def attach(self, core, extra_base: Optional[torch.Tensor] = None) -> None:
        """Bind to an EngineCore built with local_config(): registers the callback in the library.
        `extra_base`: another uint8 tensor the engine's buffers may live in (a PeerRegion's)."""
        self._bases = [core.workspace] + ([extra_base] if extra_base is not None else [])
        self.mode = "nccl"
        cabi.check(core.lib.kb_set_tensor_parallel(core._h, self.rank, self.size, self._cb, None),
                   "kb_set_tensor_parallel")

    def _callback(self, ctx, buf, n, stream) -> int:
        try:
            key = (int(buf), int(n))
            t = self._views.get(key)
            if t is None:
                for ws in self._bases:
                    off = int(buf) - ws.data_ptr()
                    if off >= 0 and off + 4 * n <= ws.numel():
                        t = self._views[key] = ws[off:off + 4 * n].view(torch.float32)
                        break
                else:
                    return 1
            if torch.cuda.current_stream().cuda_stream != int(stream or 0):
                return 2        # the collective must be ordered on the stream the kernels run on
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
            self.calls += 1
            self.bytes += 4 * int(n)
            return 0
        except Exception:       # never let an exception cross the C boundary
            return 3


class PeerRegion:
    """A symmetric allocation (same size on every rank) mapped into every rank's address space, for the
    in-kernel all-reduce of csrc/kb_collective.cu. torch's symmetric-memory rendezvous is used for the
    MAPPING only (allocation, handle exchange, multicast binding); the collective itself is our kernel.

    Layout: [payload bytes | barrier flags (cabi.TP_FLAG_BYTES)]."""

    def __init__(self, payload_bytes: int, device, group=None, use_multicast: bool = True):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        group = group if group is not None else dist.group.WORLD
        self.rank, self.size = dist.get_rank(group), dist.get_world_size(group)
        if self.size > cabi.TP_MAX_RANKS:
            raise ValueError(f"at most {cabi.TP_MAX_RANKS} ranks")
        self.flags_offset = (payload_bytes + 255) & ~255
        self.bytes = self.flags_offset + cabi.TP_FLAG_BYTES
        self.buf = symm_mem.empty(self.bytes, dtype=torch.uint8, device=device)
        self.handle = symm_mem.rendezvous(self.buf, group)
        ptrs = [int(p) for p in self.handle.buffer_ptrs]
        off = self.buf.data_ptr() - ptrs[self.rank]      # the tensor may sit inside a larger mapped block
        mc = int(getattr(self.handle, "multicast_ptr", 0) or 0)
        self.multicast = bool(mc) and use_multicast
        self.peers = cabi.TpPeers()
        for r in range(self.size):
            self.peers.peer_base[r] = ptrs[r] + off
        self.peers.mc_base = (mc + off) if self.multicast else None
        self.peers.bytes = self.bytes
        self.peers.flags_offset = self.flags_offset
        # ONE epoch sequence per region, shared by allreduce_() and by every engine attached to it (the library
        # reads and advances this host word, kb_tp_peers.epoch_counter): a second engine on the same region, or
        # allreduce_() after attach(), continues where the last collective stopped instead of reusing an epoch
        # the flags have already passed
        self._epoch = C.c_uint32(1)
        self.peers.epoch_counter = C.pointer(self._epoch)
        self.peers.timeout_cycles = 0                    # library default: $KB_TP_TIMEOUT_S seconds (30)
        self.buf.zero_()
        torch.cuda.synchronize(device)
        dist.barrier(group)                              # every rank's flags are zero before anyone signals

    @property
    def epoch(self) -> int:
        """Epoch the next collective on this region must use (callers of the bare kernels advance it by 2)."""
        return int(self._epoch.value)

    @epoch.setter
    def epoch(self, v: int) -> None:
        self._epoch.value = v

    def payload(self, dtype=torch.float32) -> torch.Tensor:
        return self.buf[:self.flags_offset].view(dtype)

    def allreduce_(self, offset_bytes: int, n: int, blocks: int = 0, f16: bool = False) -> None:
        """In place on n fp32 (or fp16, added in fp32) at offset_bytes of the payload, on the current stream
        (all ranks, same order)."""
        cabi.check(cabi.load().kb_allreduce_peers(C.byref(self.peers), self.rank, self.size, offset_bytes, n,
                                                  int(f16), self._epoch.value, blocks,
                                                  torch.cuda.current_stream().cuda_stream),
                   "kb_allreduce_peers")
        self._epoch.value += 2

    WIRE = {"f32": 0, "f16": 1, "fused": 2}

    def attach(self, core, blocks: int = 0, wire: str = "f32") -> None:
        """Move the engine's residual / partial-sum buffers into this region and switch it to the in-kernel
        collective (the region must hold residual_region_bytes()). May be called again to change the grid
        or the wire format:
          "f32"    fp32 partial sums all-reduced in place (rank 0 folds the residual in)
          "f16"    fp16 partial sums (half the NVLink bytes, added in fp32), residual add in the next RMSNorm
          "fused"  fp16 partial sums; reduce-scatter + residual add + RMSNorm + all-gather of the normed rows
                   in one kernel per projection (each rank norms only its rows); hidden <= 8192"""
        cabi.check(core.lib.kb_set_tensor_parallel_peers(core._h, self.rank, self.size, C.byref(self.peers), blocks,
                                                         self.WIRE[wire]),
                   "kb_set_tensor_parallel_peers")
        self.wire = wire
        self.mode = ("peers-multicast" if self.multicast else "peers-p2p") + ("" if wire == "f32" else "-" + wire)


def residual_region_bytes(cfg: LlamaConfig, max_tokens: int, max_seqs: int = 0) -> int:
    """Payload of the engine's region: two [max_tokens, hidden] fp32 buffers, then (max_seqs > 0) the two
    (value, id) pair buffers of the vocabulary-sharded sampler, then 512 spare bytes."""
    each = (max_tokens * cfg.hidden * 4 + 255) & ~255
    return 2 * each + lm_pairs_bytes(max_seqs) + 512


def lm_pairs_bytes(max_seqs: int) -> int:
    return (2 * max_seqs * 8 + 255) & ~255


def build_tp_core(cfg: LlamaConfig, rank: int, tp: int, device, n_blocks: int, max_tokens: int, max_seqs: int,
                  weights: Optional[Dict[str, torch.Tensor]] = None, seed: int = 0, group=None,
                  allreduce: Optional[str] = None):
    """EngineCore of one tensor-parallel rank (+ the object that owns its all-reduce). `weights`: canonical
    FULL weights (tests); None draws the shard on the device (benchmarks).

    allreduce: "peers" (our kernel over the NVSwitch multicast mapping, fp32 partial sums), "peers-f16" (same,
    fp16 partial sums: half the NVLink bytes), "peers-fused" (fp16 partial sums; collective, residual add and
    RMSNorm in one kernel), "peers-p2p" (plain peer loads / stores, no multicast, fp32), "nccl"
    (torch.distributed callback), or None = $KAFKA_B200_TP_ALLREDUCE, default "auto" = "peers-fused" with a
    logged fall-back to NCCL when the symmetric mapping cannot be set up."""
    import os
    import sys
    from .engine_core import EngineCore
    loc = local_config(cfg, tp)
    mode = (allreduce or os.environ.get("KAFKA_B200_TP_ALLREDUCE", "auto")) if tp > 1 else "none"
    # lm_head by vocabulary slice: only with the peer region (the winners are compared through it), decided before
    # the weights are cut. "auto" may still fall back to NCCL below: then the slice cannot be used and the ranks
    # fail together, loudly (set KAFKA_B200_TP_LM_HEAD=replicated on such a machine).
    shard_lm = (tp > 1 and mode != "nccl" and lm_head_slice(cfg, rank, tp) is not None and
                os.environ.get("KAFKA_B200_TP_LM_HEAD", "sharded") != "replicated")
    from . import checkpoint
    if weights is not None:
        fused = fuse_for_engine(loc, shard_weights(cfg, weights, rank, tp, shard_lm_head=shard_lm), device)
    elif checkpoint.checkpoint_dir(cfg):        # real weights ($KAFKA_B200_CHECKPOINTS), cut while loading
        fused = checkpoint.load_fused_shard(checkpoint.checkpoint_dir(cfg), cfg, rank, tp, device, shard_lm_head=shard_lm)
    else:
        fused = generate_sharded_on_device(cfg, seed, rank, tp, device, shard_lm_head=shard_lm)
    core = EngineCore(loc, fused, device, n_blocks=n_blocks, max_tokens=max_tokens, max_seqs=max_seqs)
    core.lm_head_slice = None
    grp = None
    if tp > 1:
        if mode in ("auto", "peers", "peers-f16", "peers-fused", "peers-p2p"):
            err = None
            try:
                grp = PeerRegion(residual_region_bytes(loc, max_tokens, max_seqs), device, group,
                                 use_multicast=(mode != "peers-p2p"))
            except Exception as e:
                err, grp = e, None
            # the ranks must AGREE on the mode: if the region came up on some ranks only, those would spin in
            # the peer barrier while the others call NCCL. One MIN over a success flag settles it.
            import torch.distributed as dist
            ok = torch.tensor([0 if grp is None else 1], dtype=torch.int32, device=device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
            if int(ok.item()) == 0:
                if mode != "auto":
                    raise err if err is not None else RuntimeError(
                        "another rank could not set up the symmetric region for the peer-memory all-reduce")
                if shard_lm:
                    raise RuntimeError("symmetric memory unavailable on at least one rank, and the lm_head was cut by "
                                       "vocabulary for the peer region: set KAFKA_B200_TP_LM_HEAD=replicated (or "
                                       f"KAFKA_B200_TP_ALLREDUCE=nccl) on this machine ({err!r} here)")
                print(f"[kafka_b200.tp] symmetric memory unavailable on at least one rank ({err!r} here); every "
                      "rank falls back to NCCL", file=sys.stderr)
                grp = None
            else:
                grp.attach(core, int(os.environ.get("KAFKA_B200_TP_BLOCKS", "0")),
                           wire={"peers-f16": "f16", "peers-fused": "fused",
                                 "auto": "fused" if cfg.hidden <= 8192 else "f16"}.get(mode, "f32"))
                if shard_lm:
                    v0, n = lm_head_slice(cfg, rank, tp)
                    each = (max_tokens * loc.hidden * 4 + 255) & ~255
                    cabi.check(core.lib.kb_set_lm_head_shard(core._h, v0, n, 2 * each), "kb_set_lm_head_shard")
                    core.lm_head_slice = (v0, n)      # logits_out of prefill / decode is [n_seqs, n] on this rank
        if grp is None:
            grp = TPGroup(group)
            grp.attach(core)
    return core, grp
