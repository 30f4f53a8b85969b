"""Test-facing adapters with the reference's names (reference: tests/adapters.py:545-595).  Only the two
adapters on the tokenizer hot path exist; the model / optimizer adapters of the course scaffold are out of
scope."""
from __future__ import annotations

import os
from typing import Any

from bpe_transformer_b200 import train_bpe
from bpe_transformer_b200.tokenization.bpe_tokenizer import BPETokenizer


def get_tokenizer(vocab: dict[int, bytes], merges: list[tuple[bytes, bytes]], special_tokens: list[str] | None = None) -> Any:
    """A BPE tokenizer that uses the provided vocab, merges and special tokens."""
    return BPETokenizer(vocab=vocab, merges=merges, special_tokens=special_tokens)


def run_train_bpe(input_path: str | os.PathLike, vocab_size: int, special_tokens: list[str], **kwargs) -> tuple[dict[int, bytes], list[tuple[bytes, bytes]]]:
    """Train a BPE tokenizer on `input_path`; returns (vocab, merges)."""
    return train_bpe(input_path=input_path, vocab_size=vocab_size, special_tokens=special_tokens)
