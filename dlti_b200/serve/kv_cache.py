"""Paged KV cache: fixed-size blocks + per-sequence block tables (vLLM's
PagedAttention memory model, restated). Tensors: k_cache, v_cache = [L][num_blocks, block_size,
KV_heads_local, head_dim] in the compute dtype; the block manager is pure host-side bookkeeping."""
from __future__ import annotations

from collections import deque

import torch


class BlockManager:
    """Free list + (optional) prefix cache over fixed-size KV blocks.

    Without prefix caching a block belongs to exactly one sequence. With `enable_prefix_caching` a FULL block
    whose content is determined by the token prefix it ends (hash chain over block_size-token groups, as in vLLM's
    automatic prefix caching, SP:vllm/v1/core/kv_cache_utils.py) is registered under that hash when its sequence
    has written it; a later prompt that starts with the same tokens takes a reference on the block instead of
    recomputing it. Reference counts keep shared blocks alive; a block whose count drops to zero goes to the
    free list but keeps its hash until it is handed out again (oldest freed first), so a hot prefix survives as
    long as memory allows."""

    def __init__(self, num_blocks: int, block_size: int = 16, enable_prefix_caching: bool = False):
        self.num_blocks, self.block_size = num_blocks, block_size
        self.free: deque[int] = deque(range(num_blocks - 1, -1, -1))      # handed out from the right end
        self.prefix_caching = enable_prefix_caching
        self.ref = [0] * num_blocks
        self.hash_of: dict[int, int] = {}          # block -> prefix hash
        self.cached: dict[int, int] = {}           # prefix hash -> block
        self.hits = self.queries = 0               # in tokens

    def blocks_needed(self, num_tokens: int) -> int:
        return -(-num_tokens // self.block_size)

    def can_allocate(self, num_tokens: int, have: int = 0) -> bool:
        return self.blocks_needed(num_tokens) - have <= len(self.free)

    def _take(self) -> int:
        b = self.free.pop()
        h = self.hash_of.pop(b, None)              # the block is about to be overwritten: forget its old content
        if h is not None and self.cached.get(h) == b:
            del self.cached[h]
        self.ref[b] = 1
        return b

    def ensure(self, blocks: list[int], num_tokens: int) -> bool:
        """Grow `blocks` so it covers num_tokens. False (nothing changed) if out of blocks."""
        need = self.blocks_needed(num_tokens) - len(blocks)
        if need > len(self.free):
            return False
        for _ in range(max(0, need)):
            blocks.append(self._take())
        return True

    def release(self, blocks: list[int]) -> None:
        if not self.prefix_caching:
            for b in blocks:
                self.ref[b] = 0
            self.free.extend(reversed(blocks))
            blocks.clear()
            return
        for b in blocks:                           # oldest-freed-first reuse: new entries go to the FRONT
            self.ref[b] -= 1
            if self.ref[b] == 0:
                self.free.appendleft(b)
        blocks.clear()

    # ---- prefix cache -----------------------------------------------------------------------------
    def _chain(self, tokens: list[int], n_blocks: int):
        h = 0
        bs = self.block_size
        for i in range(n_blocks):
            h = hash((h, tuple(tokens[i * bs:(i + 1) * bs])))
            yield h

    def match_prefix(self, tokens: list[int]) -> list[int]:
        """Blocks holding the longest cached prefix of `tokens` (whole blocks only, and never the whole
        sequence: at least one token must be computed to produce logits). Takes a reference on each."""
        if not self.prefix_caching:
            return []
        self.queries += len(tokens)
        out = []
        for h in self._chain(tokens, (len(tokens) - 1) // self.block_size):
            b = self.cached.get(h)
            if b is None:
                break
            out.append(b)
        for b in out:
            if self.ref[b] == 0:
                self.free.remove(b)                # resurrect a cached block that was waiting in the free list
            self.ref[b] += 1
        self.hits += len(out) * self.block_size
        return out

    def register(self, blocks: list[int], tokens: list[int], num_written: int) -> None:
        """The first num_written tokens of the sequence are in `blocks`: publish its full blocks."""
        if not self.prefix_caching:
            return
        for i, h in enumerate(self._chain(tokens, min(num_written // self.block_size, len(blocks)))):
            b = blocks[i]
            if b in self.hash_of:
                continue
            if h in self.cached:                   # another sequence published the same prefix first
                continue
            self.hash_of[b] = h
            self.cached[h] = b

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
