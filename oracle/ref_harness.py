"""Import the UNMODIFIED reference (milasd/BPE-Transformer) from /root/reference.  TEST INFRASTRUCTURE ONLY.

Used by oracle/gen_golden.py (to make tests/golden/*) and by the `not gpu` tests that pin the oracle while
/root/reference is mounted.  /root/reference does not exist on the GPU box: nothing that runs there may
import this module — `available()` says whether it can be used.

Two obstacles: bpe_transformer/__init__.py:4 asks importlib.metadata for the installed
version of a package that is not installed (answered with a throw-away dist-info on sys.path), and
tiktoken.get_encoding("gpt2") needs the network (answered by building the Encoding from the reference's
own gpt2_vocab.json fixture).
"""
from __future__ import annotations

import json
import os
import pathlib
import sys
import tempfile

REF_ROOT = pathlib.Path(os.environ.get("BPE_REFERENCE_ROOT", "/root/reference"))
FIXTURES = REF_ROOT / "tests" / "fixtures"
SAMPLE_MODEL = REF_ROOT / "notebooks" / "sample_data" / "bpe_tokenizer"

_loaded = None


def available() -> bool:
    return (REF_ROOT / "bpe_transformer" / "main.py").exists()


def load():
    """Returns the reference's `bpe_transformer` package (imported once)."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError(f"reference not mounted at {REF_ROOT}")
    shim = pathlib.Path(tempfile.mkdtemp(prefix="bpe_ref_shim_"))
    meta = shim / "bpe_transformer-0.5.0.dist-info"
    meta.mkdir()
    (meta / "METADATA").write_text("Metadata-Version: 2.1\nName: bpe_transformer\nVersion: 0.5.0\n")
    sys.dont_write_bytecode = True  # the reference tree is read-only
    sys.path.insert(0, str(shim))
    sys.path.insert(0, str(REF_ROOT))
    import bpe_transformer  # noqa: E402
    import bpe_transformer.tokenization.bpe_tokenizer  # noqa: F401,E402
    import bpe_transformer.tokenization.preprocessing.pretokenization  # noqa: F401,E402
    _loaded = bpe_transformer
    return _loaded


def gpt2_byte_decoder() -> dict[str, int]:
    """Inverse of GPT-2's printable-byte alphabet (the reference restates it in tests/common.py:9-54;
    this is the well-known construction, written out here so nothing is imported from the tests)."""
    keep = list(range(33, 127)) + list(range(161, 173)) + list(range(174, 256))
    chars = keep[:]
    extra = 0
    for b in range(256):
        if b not in keep:
            keep.append(b)
            chars.append(256 + extra)
            extra += 1
    return {chr(c): b for b, c in zip(keep, chars)}


def load_gpt2_model(vocab_json=None, merges_txt=None):
    """vocab dict[int, bytes] and merges list[(bytes, bytes)] from the GPT-2 fixture files
    (same filtering as the reference's tests/test_tokenizer.py:39-74)."""
    dec = gpt2_byte_decoder()
    with open(vocab_json or FIXTURES / "gpt2_vocab.json") as fh:
        raw = json.load(fh)
    vocab = {idx: bytes(dec[ch] for ch in tok) for tok, idx in raw.items()}
    merges = []
    with open(merges_txt or FIXTURES / "gpt2_merges.txt") as fh:
        for line in fh:
            parts = line.rstrip().split(" ")
            if line.rstrip() and len(parts) == 2:
                merges.append((bytes(dec[ch] for ch in parts[0]), bytes(dec[ch] for ch in parts[1])))
    return vocab, merges


def offline_tiktoken_gpt2():
    """tiktoken GPT-2 Encoding built without the network."""
    import tiktoken

    dec = gpt2_byte_decoder()
    with open(FIXTURES / "gpt2_vocab.json") as fh:
        raw = json.load(fh)
    ranks = {bytes(dec[ch] for ch in tok): idx for tok, idx in raw.items() if tok != "<|endoftext|>"}
    pat = r"""'(?:[sdmt]|ll|ve|re)| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)|\s+"""
    return tiktoken.Encoding("gpt2-offline", pat_str=pat, mergeable_ranks=ranks,
                             special_tokens={"<|endoftext|>": 50256})
