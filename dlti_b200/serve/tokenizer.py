"""Tokenizer front-end for the server: the model directory's HF tokenizer when there is one, else a
byte-level fallback (ids 3..258 = bytes) so random-init / synthetic models can be served and load-
tested without any tokenizer files. Includes incremental detokenisation for streaming."""
from __future__ import annotations

import os


class ByteTokenizer:
    bos_token_id, eos_token_id, pad_token_id = 1, 2, 2

    def __init__(self, vocab_size: int = 32000):
        self.vocab_size = vocab_size

    def encode(self, text: str, add_bos: bool = True) -> list[int]:
        ids = [b + 3 for b in text.encode("utf-8")]
        return ([self.bos_token_id] if add_bos else []) + ids

    def decode(self, ids: list[int], skip_special_tokens: bool = True) -> str:
        bs = bytes((i - 3) % 256 for i in ids if i >= 3 or not skip_special_tokens)
        return bs.decode("utf-8", errors="replace")


class HFTokenizer:
    def __init__(self, path: str):
        from transformers import AutoTokenizer

        self.tok = AutoTokenizer.from_pretrained(path)
        self.bos_token_id = self.tok.bos_token_id
        self.eos_token_id = self.tok.eos_token_id
        self.pad_token_id = self.tok.pad_token_id if self.tok.pad_token_id is not None else self.eos_token_id
        self.vocab_size = len(self.tok)

    def encode(self, text: str, add_bos: bool = True) -> list[int]:
        return self.tok.encode(text, add_special_tokens=add_bos)

    def decode(self, ids: list[int], skip_special_tokens: bool = True) -> str:
        return self.tok.decode(ids, skip_special_tokens=skip_special_tokens)


def load_tokenizer(model_path: str | None, vocab_size: int):
    if model_path and any(os.path.exists(os.path.join(model_path, f))
                          for f in ("tokenizer.json", "tokenizer.model", "tokenizer_config.json")):
        return HFTokenizer(model_path)
    return ByteTokenizer(vocab_size)


class IncrementalDetokenizer:
    """Emit only text that can no longer change, in O(new tokens) per step: decode a short window of tokens —
    [prefix, read) was already emitted, [read, end) is new — and emit decode(prefix..end) minus decode(prefix..read)
    (the window gives the tokenizer the left context that decides spaces / merges). A trailing replacement
    character (an incomplete UTF-8 / multi-token glyph) is held back until the next token completes it."""
    WINDOW = 6

    def __init__(self, tokenizer, skip_special_tokens: bool = True):
        self.tok, self.skip = tokenizer, skip_special_tokens
        self.prefix = 0          # start of the context window
        self.read = 0            # tokens [0, read) have been turned into emitted text

    def step(self, output_ids: list[int], final: bool = False) -> str:
        n = len(output_ids)
        if n <= self.read and not final:
            return ""
        before = self.tok.decode(output_ids[self.prefix:self.read], skip_special_tokens=self.skip) if self.read > self.prefix else ""
        text = self.tok.decode(output_ids[self.prefix:n], skip_special_tokens=self.skip)
        if not final and text.endswith("\ufffd"):
            return ""            # wait for the rest of the character; nothing is consumed
        new = text[len(before):] if text.startswith(before) else text[len(before.rstrip("\ufffd")):]
        self.read = n
        self.prefix = max(self.prefix, n - self.WINDOW)
        return new
