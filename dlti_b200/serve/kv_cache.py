"""Paged KV cache: fixed-size blocks + per-sequence block tables (vLLM's
PagedAttention memory model, restated). Tensors: k_cache, v_cache = [L][num_blocks, block_size,
KV_heads_local, head_dim] in the compute dtype; the block manager is pure host-side bookkeeping."""
from __future__ import annotations

from collections import OrderedDict

import torch


class BlockManager:
    """Free list + (optional) prefix cache over fixed-size KV blocks.

    Without prefix caching a block belongs to exactly one sequence. With `enable_prefix_caching` a FULL block
    whose content is determined by the token prefix it ends (hash chain over block_size-token groups, as in vLLM's
    automatic prefix caching, SP:vllm/v1/core/kv_cache_utils.py) is registered under that hash when its sequence
    has written it; a later prompt that starts with the same tokens takes a reference on the block instead of
    recomputing it. Reference counts keep shared blocks alive; a block whose count drops to zero goes to the
    free list but keeps its hash until it is handed out again (oldest freed first, and within one freed sequence
    the TAIL blocks first: losing block 0 would break the hash chain and strand every later block), so a hot
    prefix survives as long as memory allows. A hit is confirmed by comparing the block's tokens and its parent
    block, so a 64-bit hash collision can never hand one prompt another prompt's KV. The free list is an ordered
    dict: taking, freeing and resurrecting a cached block are all O(1)."""

    def __init__(self, num_blocks: int, block_size: int = 16, enable_prefix_caching: bool = False):
        self.num_blocks, self.block_size = num_blocks, block_size
        self.free: OrderedDict[int, None] = OrderedDict((b, None) for b in range(num_blocks))   # handed out from the front
        self.prefix_caching = enable_prefix_caching
        self.ref = [0] * num_blocks
        self.hash_of: dict[int, int] = {}          # block -> prefix hash
        self.cached: dict[int, tuple] = {}         # prefix hash -> (block, parent block or -1, tokens of the block)
        self.hits = self.queries = 0               # in tokens

    def blocks_needed(self, num_tokens: int) -> int:
        return -(-num_tokens // self.block_size)

    def can_allocate(self, num_tokens: int, have: int = 0) -> bool:
        return self.blocks_needed(num_tokens) - have <= len(self.free)

    def _take(self) -> int:
        b, _ = self.free.popitem(last=False)
        h = self.hash_of.pop(b, None)              # the block is about to be overwritten: forget its old content
        if h is not None and self.cached.get(h, (None,))[0] == b:
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
            for b in reversed(blocks):             # most recently freed first (warm in L2), blocks[0] at the front
                self.ref[b] = 0
                self.free[b] = None
                self.free.move_to_end(b, last=False)
            blocks.clear()
            return
        for b in reversed(blocks):                 # oldest-freed-first reuse: new entries go to the BACK, tail first
            self.ref[b] -= 1
            if self.ref[b] == 0:
                self.free[b] = None
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
        bs = self.block_size
        for i, h in enumerate(self._chain(tokens, (len(tokens) - 1) // bs)):
            e = self.cached.get(h)
            if e is None or e[1] != (out[-1] if out else -1) or e[2] != tuple(tokens[i * bs:(i + 1) * bs]):
                break                              # miss, or a hash collision (different tokens / different parent)
            out.append(e[0])
        for b in out:
            if self.ref[b] == 0:
                del self.free[b]                   # resurrect a cached block that was waiting in the free list
            self.ref[b] += 1
        self.hits += len(out) * self.block_size
        return out

    def register(self, blocks: list[int], tokens: list[int], num_written: int) -> None:
        """The first num_written tokens of the sequence are in `blocks`: publish its full blocks."""
        if not self.prefix_caching:
            return
        bs = self.block_size
        for i, h in enumerate(self._chain(tokens, min(num_written // bs, len(blocks)))):
            b = blocks[i]
            if b in self.hash_of:
                continue
            if h in self.cached:                   # another sequence published the same prefix first
                continue
            self.hash_of[b] = h
            self.cached[h] = (b, blocks[i - 1] if i else -1, tuple(tokens[i * bs:(i + 1) * bs]))

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
