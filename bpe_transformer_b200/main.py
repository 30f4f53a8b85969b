"""`train_bpe` entry point (reference: bpe_transformer/main.py:8-17)."""
from __future__ import annotations

from multiprocessing import cpu_count
from pathlib import Path

from .tokenization.bpe_trainer import BPETrainer

N_WORKERS = cpu_count()  # main.py:5 — only decides where a corpus WITHOUT special tokens is cut (at newlines)


def train_bpe(
    input_path: str | Path | None = None,
    vocab_size: int = 0,
    special_tokens: list[str] | None = None,
    *,
    corpus_path: str | Path | None = None,
) -> tuple[dict[int, bytes], list[tuple[bytes, bytes]]]:
    """Train a byte-level BPE tokenizer; returns (vocab, merges) exactly as the reference does.

    `corpus_path` is accepted as an alias of `input_path` (the reference's README uses that name).
    """
    if input_path is None:
        input_path = corpus_path
    if input_path is None:
        raise TypeError("train_bpe() needs input_path")
    special_tokens = list(special_tokens or [])
    if vocab_size < 255 + len(special_tokens):  # main.py:12-13
        raise ValueError("Input vocab_size is invalid: value too small.")
    bpe = BPETrainer(vocab_size=vocab_size, special_tokens=special_tokens)
    bpe.train(input_path=input_path, n_workers=N_WORKERS)
    return bpe.vocab, bpe.merges
