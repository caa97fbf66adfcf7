"""Paged KV cache: fixed-size blocks + per-sequence block tables (vLLM's
PagedAttention memory model, restated). Tensors: k_cache, v_cache = [L][num_blocks, block_size,
KV_heads_local, head_dim] in the compute dtype; the block manager is pure host-side bookkeeping."""
from __future__ import annotations

import torch


class BlockManager:
    def __init__(self, num_blocks: int, block_size: int = 16):
        self.num_blocks, self.block_size = num_blocks, block_size
        self.free: list[int] = list(range(num_blocks - 1, -1, -1))

    def blocks_needed(self, num_tokens: int) -> int:
        return -(-num_tokens // self.block_size)

    def can_allocate(self, num_tokens: int, have: int = 0) -> bool:
        return self.blocks_needed(num_tokens) - have <= len(self.free)

    def ensure(self, blocks: list[int], num_tokens: int) -> bool:
        """Grow `blocks` so it covers num_tokens. False (nothing changed) if out of blocks."""
        need = self.blocks_needed(num_tokens) - len(blocks)
        if need > len(self.free):
            return False
        for _ in range(max(0, need)):
            blocks.append(self.free.pop())
        return True

    def release(self, blocks: list[int]) -> None:
        self.free.extend(reversed(blocks))
        blocks.clear()

    @property
    def num_free(self) -> int:
        return len(self.free)

    def usage(self) -> float:
        return 1.0 - len(self.free) / max(1, self.num_blocks)


class KVCache:
    def __init__(self, num_layers: int, num_blocks: int, block_size: int, kv_heads: int, head_dim: int,
                 dtype: torch.dtype, device):
        shape = (num_blocks, block_size, kv_heads, head_dim)
        self.k = [torch.zeros(shape, dtype=dtype, device=device) for _ in range(num_layers)]
        self.v = [torch.zeros(shape, dtype=dtype, device=device) for _ in range(num_layers)]
        self.num_blocks, self.block_size = num_blocks, block_size

    @staticmethod
    def bytes_per_block(num_layers, block_size, kv_heads, head_dim, dtype) -> int:
        return 2 * num_layers * block_size * kv_heads * head_dim * torch.empty(0, dtype=dtype).element_size()
