"""The C-ABI library loads and exports every symbol include/bpe_b200.h declares (no compute: runs without a GPU)."""
from __future__ import annotations

import ctypes
import pathlib
import re

ROOT = pathlib.Path(__file__).resolve().parent.parent


def declared_symbols() -> list[str]:
    text = (ROOT / "include" / "bpe_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(bpe_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as entry

    entry.build()
    from bpe_transformer_b200 import _lib

    lib = ctypes.CDLL(str(_lib.LIB_PATH))
    names = declared_symbols()
    assert len(names) >= 18
    for name in names:
        assert hasattr(lib, name), f"{name} is declared in include/bpe_b200.h but not exported"
    assert sorted(_lib.SYMBOLS) == [n for n in names if n in _lib.SYMBOLS]
    assert set(names) <= set(_lib.SYMBOLS) | {"bpe_kernel_time"}


def test_no_cpu_fallback_without_cuda():
    """On a box without a GPU the product path must fail loudly, not route around the CUDA library."""
    import torch

    if torch.cuda.is_available():
        return
    import pytest

    from bpe_transformer_b200 import train_bpe
    from bpe_transformer_b200.tokenization import BPETokenizer

    with pytest.raises(RuntimeError):
        train_bpe(str(ROOT / "README.md"), 300, ["<|endoftext|>"])
    with pytest.raises(RuntimeError):
        BPETokenizer({i: bytes([i]) for i in range(256)}, []).encode("abc")


def test_product_does_not_import_the_oracle():
    for path in (ROOT / "bpe_transformer_b200").rglob("*.py"):
        src = path.read_text()
        assert "oracle" not in src.replace("the oracle compares", ""), f"{path} mentions the oracle"
        assert "import regex" not in src and "from regex" not in src, f"{path} imports regex"


def test_gpt2_text_formats_round_trip(tmp_path):
    """Host-only: formats.save_gpt2_files / load_gpt2_files against the golden GPT-2 model."""
    from conftest import load_golden
    from bpe_transformer_b200 import formats

    g = load_golden("gpt2_model.json.gz")
    vocab = {i: bytes.fromhex(h) for i, h in enumerate(g["vocab"])}
    merges = [(bytes.fromhex(a), bytes.fromhex(b)) for a, b in g["merges"]]
    formats.save_gpt2_files(vocab, merges, tmp_path / "v.json", tmp_path / "m.txt")
    v2, m2 = formats.load_gpt2_files(tmp_path / "v.json", tmp_path / "m.txt")
    assert v2 == vocab and m2 == merges
    v3, _ = formats.load_gpt2_files(tmp_path / "v.json", tmp_path / "m.txt", ["<|new|>"])
    assert v3[len(vocab)] == b"<|new|>"
