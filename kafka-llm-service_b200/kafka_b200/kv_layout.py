"""Host-side description of the paged KV layout in HBM (DESIGN.md "Data layout in HBM").

    pool[layer][block][kv_head][token_in_block][head_dim]      fp16 (KV_DTYPE), K and V in separate pools

One (block, kv_head) tile is ``KB_BLOCK_TOKENS * head_dim`` contiguous elements, stored as
``[dim plane = dim // 64][token][64 dims]``: every 64-dim plane is a 64-row x 128-byte matrix whose
16-byte chunks (8 elements) are XOR-swizzled with ``token & 7`` — the canonical SWIZZLE_128B K-major form, so
a plain 1-D TMA bulk copy of a tile lands in shared memory ready for ``ldmatrix`` (no bank conflicts)
and for tcgen05 shared-memory descriptors alike.

These helpers exist for tests, smoke and debugging tools; the product path never moves KV through
the host (the append kernel writes the layout directly).
"""
from __future__ import annotations

import torch

from .cabi import KB_BLOCK_TOKENS

KV_DTYPE = torch.float16   # q / k / v, the pools and the attention probabilities (include/kafka_b200.h, KB_GEMM_F16)


def tile_permutation(head_dim: int) -> torch.Tensor:
    """perm[t * head_dim + d] = physical element offset of logical (token t, dim d) inside a tile."""
    t = torch.arange(KB_BLOCK_TOKENS).view(-1, 1)
    d = torch.arange(head_dim).view(1, -1)
    plane = d // 64
    chunk = ((d % 64) // 8) ^ (t & 7)
    return (plane * (KB_BLOCK_TOKENS * 64) + t * 64 + chunk * 8 + (d % 8)).reshape(-1)


def pool_shape(n_layers: int, n_blocks: int, n_kv_heads: int, head_dim: int):
    return (n_layers, n_blocks, n_kv_heads, KB_BLOCK_TOKENS * head_dim)


def write_tokens(pool_layer: torch.Tensor, block_ids, start_pos: int, x: torch.Tensor) -> None:
    """Write x[L, n_kv_heads, head_dim] at positions start_pos.. of a sequence owning ``block_ids``.

    pool_layer: [n_blocks, n_kv_heads, KB_BLOCK_TOKENS*head_dim]
    """
    L, n_kv, hd = x.shape
    perm = tile_permutation(hd).to(pool_layer.device)
    pos = torch.arange(start_pos, start_pos + L, device=pool_layer.device)
    blk = torch.as_tensor(block_ids, device=pool_layer.device, dtype=torch.long)[pos // KB_BLOCK_TOKENS]
    tib = pos % KB_BLOCK_TOKENS
    # physical offsets [L, hd]
    off = perm.view(KB_BLOCK_TOKENS, hd)[tib]
    for h in range(n_kv):
        pool_layer[blk.view(-1, 1).expand(L, hd), h, off] = x[:, h, :].to(pool_layer.dtype)


def read_tokens(pool_layer: torch.Tensor, block_ids, length: int, head_dim: int) -> torch.Tensor:
    """Inverse of write_tokens: returns [length, n_kv_heads, head_dim]."""
    n_kv = pool_layer.shape[1]
    perm = tile_permutation(head_dim).to(pool_layer.device)
    pos = torch.arange(0, length, device=pool_layer.device)
    blk = torch.as_tensor(block_ids, device=pool_layer.device, dtype=torch.long)[pos // KB_BLOCK_TOKENS]
    tib = pos % KB_BLOCK_TOKENS
    off = perm.view(KB_BLOCK_TOKENS, head_dim)[tib]
    out = torch.empty((length, n_kv, head_dim), dtype=pool_layer.dtype, device=pool_layer.device)
    for h in range(n_kv):
        out[:, h, :] = pool_layer[blk.view(-1, 1).expand(length, head_dim), h, off]
    return out
