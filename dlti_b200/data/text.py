"""Text-side data utilities: the Llama-2 chat template used by the reference's dataset prep,
tokenisation with truncation, tokenizer loading with pad = eos, and `load_from_disk` datasets
(reference scripts/prepare_dataset.py:12-25, training/train_baseline.py:115-117,149-165;
SURVEY.md D3, D26). `datasets` / `transformers` are optional dependencies, not hot paths.
"""
from __future__ import annotations


def format_conversation_for_llama2(example: dict) -> dict:
    """{question, answer} -> {"text": "<s>[INST] q [/INST] a</s>"}"""
    q = str(example.get("question", "")).strip()
    a = str(example.get("answer", "")).strip()
    return {"text": f"<s>[INST] {q} [/INST] {a}</s>"}


def format_conversations_for_llama2(batch: dict) -> dict:
    """Batched form of format_conversation_for_llama2 for `Dataset.map(..., batched=True)`: same text per row,
    one Python call per few thousand rows instead of one per row (136,109 synthetic rows: 600 k examples/s vs
    94 k row-wise on the build box; the reference's row-wise map logged 23,943 examples/s, train.ipynb:56)."""
    return {"text": [f"<s>[INST] {str(q).strip()} [/INST] {str(a).strip()}</s>"
                     for q, a in zip(batch["question"], batch["answer"])]}


def load_tokenizer(name_or_path: str):
    from transformers import AutoTokenizer

    tok = AutoTokenizer.from_pretrained(name_or_path)
    if tok.pad_token is None:
        tok.pad_token = tok.eos_token
    return tok


def tokenize_dataset(dataset, tokenizer, max_length: int = 512, desc: str | None = None):
    def tok(examples):
        return tokenizer(examples["text"], truncation=True, max_length=max_length, padding=False)

    return dataset.map(tok, batched=True, remove_columns=dataset.column_names, desc=desc)


def load_text_dataset(path: str):
    from datasets import load_from_disk

    return load_from_disk(path)
