"""Shared test plumbing: the `gpu` marker, golden-fixture loaders, the seeded synthetic-corpus generator.

`-m "not gpu"` runs here (no GPU): oracle vs. golden vectors, host logic, C-ABI symbol check.
`-m gpu` runs on a B200: CUDA path vs. oracle / goldens, through the C-ABI.
"""
from __future__ import annotations

import ctypes
import gzip
import hashlib
import json
import os
import pathlib
import subprocess
import sys
import tempfile

import pytest

ROOT = pathlib.Path(__file__).resolve().parent.parent
GOLDEN = ROOT / "tests" / "golden"
for p in (str(ROOT), str(ROOT / "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

EOT = "<|endoftext|>"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_golden(name: str):
    with gzip.open(GOLDEN / name, "rb") as fh:
        return json.load(fh)


_synth_lib = None


def synth_corpus(style: int, seed: int, nbytes: int, path: str | None = None) -> bytes | str:
    """tools/gen_corpus.c through ctypes.  Returns the bytes, or the path when `path` is given."""
    global _synth_lib
    if _synth_lib is None:
        so = ROOT / "tools" / "libgencorpus.so"
        src = ROOT / "tools" / "gen_corpus.c"
        if not so.exists() or so.stat().st_mtime < src.stat().st_mtime:
            subprocess.check_call(["gcc", "-O2", "-shared", "-fPIC", "-pthread", "-o", str(so), str(src)])
        _synth_lib = ctypes.CDLL(str(so))
        _synth_lib.gen_corpus.restype = ctypes.c_uint64
    keep = path is not None
    if path is None:
        fd, path = tempfile.mkstemp(suffix=".txt")
        os.close(fd)
    n = _synth_lib.gen_corpus(ctypes.c_int(style), ctypes.c_uint64(seed), ctypes.c_uint64(nbytes), str(path).encode(),
                              ctypes.c_int(min(8, os.cpu_count() or 1)))
    assert n > 0
    if keep:
        return path
    data = open(path, "rb").read()
    os.unlink(path)
    return data


@pytest.fixture(scope="session")
def texts():
    return {k: bytes.fromhex(v) for k, v in load_golden("texts.json.gz").items()}


@pytest.fixture(scope="session")
def synth_specs():
    return load_golden("synth_sha256.json.gz")


@pytest.fixture(scope="session")
def input_bytes(texts, synth_specs):
    cache: dict[str, bytes] = {}

    def get(key: str) -> bytes:
        if key in cache:
            return cache[key]
        if key.startswith("synth:"):
            spec = synth_specs[key]
            data = synth_corpus(spec["style"], spec["seed"], spec["bytes"])
            assert hashlib.sha256(data).hexdigest() == spec["sha256"], "synthetic generator is not reproducible"
        else:
            data = texts[key]
        cache[key] = data
        return data

    return get


@pytest.fixture(scope="session")
def gpt2_model():
    g = load_golden("gpt2_model.json.gz")
    vocab = {i: bytes.fromhex(h) for i, h in enumerate(g["vocab"])}
    merges = [(bytes.fromhex(a), bytes.fromhex(b)) for a, b in g["merges"]]
    return vocab, merges


@pytest.fixture(scope="session")
def ts10k_model():
    g = load_golden("ts10k_model.json.gz")
    vocab = {int(i): bytes.fromhex(h) for i, h in g["vocab"].items()}
    merges = [(bytes.fromhex(a), bytes.fromhex(b)) for a, b in g["merges"]]
    return vocab, merges


def with_specials(vocab, specials):
    """Append specials missing from the vocab at id len(vocab) (reference: bpe_tokenizer.py:311-320)."""
    v = dict(vocab)
    have = set(v.values())
    for s in specials or []:
        if s.encode("utf-8") not in have:
            v[len(v)] = s.encode("utf-8")
    return v


def expand_row(row: dict, key: str):
    """Resolve the "=other" de-duplication of tests/golden/encode.json.gz."""
    v = row[key]
    while isinstance(v, str):
        v = row[v[1:]]
    return v


@pytest.fixture()
def tmp_file(tmp_path):
    def make(data: bytes, name: str = "input.txt") -> str:
        p = tmp_path / name
        p.write_bytes(data)
        return str(p)

    return make
