"""File formats either side of the tokenizer path.

  * model files: the reference's pickles (vocab.pkl / merges.pkl, bpe_trainer.py:447-472, bpe_tokenizer.py:292-337)
    are handled by BPETrainer.save_trainer and BPETokenizer.from_files; this module adds the GPT-2 text formats
    (vocab.json + merges.txt) that the reference's tests load (tests/test_tokenizer.py:39-74, tests/common.py:9-54);
  * token datasets: ids as a flat uint16 / uint32 array (.npy, memory-mappable) — the 1-D integer array the
    reference's data-loading adapter expects (tests/adapters.py:401-421).
"""
from __future__ import annotations

import json
import os
from pathlib import Path

import numpy as np


def gpt2_bytes_to_unicode() -> dict[int, str]:
    """GPT-2's printable alphabet: byte -> the character that stands for it in vocab.json / merges.txt."""
    keep = list(range(ord("!"), ord("~") + 1)) + list(range(0xA1, 0xAD)) + list(range(0xAE, 0x100))
    chars = keep[:]
    extra = 0
    for b in range(256):
        if b not in keep:
            keep.append(b)
            chars.append(256 + extra)
            extra += 1
    return {b: chr(c) for b, c in zip(keep, chars)}


def load_gpt2_files(vocab_path: str | os.PathLike, merges_path: str | os.PathLike, special_tokens: list[str] | None = None):
    """(vocab, merges) from GPT-2 style vocab.json + merges.txt; special tokens missing from the vocabulary are
    appended at id len(vocab), as the reference's loaders do."""
    decode = {c: b for b, c in gpt2_bytes_to_unicode().items()}
    with open(vocab_path, encoding="utf-8") as fh:
        vocab = {idx: bytes(decode[ch] for ch in tok) for tok, idx in json.load(fh).items()}
    merges: list[tuple[bytes, bytes]] = []
    with open(merges_path, encoding="utf-8") as fh:
        for line in fh:
            parts = line.rstrip().split(" ")
            if line.rstrip() and len(parts) == 2:
                merges.append((bytes(decode[ch] for ch in parts[0]), bytes(decode[ch] for ch in parts[1])))
    present = set(vocab.values())
    for s in special_tokens or []:
        raw = s.encode("utf-8")
        if raw not in present:
            vocab[len(vocab)] = raw
    return vocab, merges


def save_gpt2_files(vocab: dict[int, bytes], merges: list[tuple[bytes, bytes]], vocab_path, merges_path) -> None:
    """The inverse of load_gpt2_files (the format of the reference's train-bpe-reference-* fixtures)."""
    enc = gpt2_bytes_to_unicode()
    text = {"".join(enc[b] for b in tok): idx for idx, tok in vocab.items()}
    with open(vocab_path, "w", encoding="utf-8") as fh:
        json.dump(text, fh, ensure_ascii=False)
    with open(merges_path, "w", encoding="utf-8") as fh:
        for a, b in merges:
            fh.write("".join(enc[x] for x in a) + " " + "".join(enc[x] for x in b) + "\n")


def encode_file_to_npy(tokenizer, text_path, out_path, engine=None) -> np.ndarray:
    """Tokenize a UTF-8 text file line by line (the ids of tokenizer.encode_iterable(open(text_path))) and write
    them as a flat .npy array: uint16 when every id fits, else uint32.  Returns the array."""
    ids = tokenizer.encode_file_to_array(text_path, engine=engine)
    np.save(out_path, ids, allow_pickle=False)
    return ids


def load_token_dataset(path, mmap: bool = True) -> np.ndarray:
    """A token dataset written by encode_file_to_npy, memory-mapped by default."""
    return np.load(Path(path), mmap_mode="r" if mmap else None, allow_pickle=False)
