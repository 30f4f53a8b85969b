"""ctypes front end of oracle/bpe_oracle.c.  TEST INFRASTRUCTURE ONLY (see the header of that file).

Builds oracle/liboracle.so with gcc on first use and exposes the four restated steps with Python types
that mirror the reference's: a Counter keyed by tuples of byte values, (vocab, merges), lists of ids.
"""
from __future__ import annotations

import ctypes as C
import pathlib
import subprocess
from collections import Counter

import numpy as np

HERE = pathlib.Path(__file__).resolve().parent
ROOT = HERE.parent
LIB = HERE / "liboracle.so"
SRC = HERE / "bpe_oracle.c"

CUT_AFTER_NEWLINE = 1
_lib = None


class InvalidUtf8(ValueError):
    pass


def build(force: bool = False) -> pathlib.Path:
    hdr = ROOT / "include" / "bpe_unicode_tables.h"
    if force or not LIB.exists() or LIB.stat().st_mtime < max(SRC.stat().st_mtime, hdr.stat().st_mtime):
        subprocess.check_call(["gcc", "-O2", "-shared", "-fPIC", "-I", str(ROOT / "include"), "-o", str(LIB), str(SRC)])
    return LIB


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(str(build()))
        _lib.orc_count.restype = C.c_int64
    return _lib


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def _specials(specials):
    """Priority order = longest first, stable (pretokenization.py:228)."""
    order = sorted([s.encode("utf-8") if isinstance(s, str) else bytes(s) for s in (specials or [])], key=len, reverse=True)
    blob = np.frombuffer(b"".join(order) + b"\0", dtype=np.uint8).copy()
    off = np.zeros(len(order) + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(s) for s in order], dtype=np.uint32) if order else []
    return order, blob, off


def scan(data, specials=None, cuts=None, flags: int = 0) -> np.ndarray:
    """Token-start bitmask (uint32 words, bit i of word i//32 = byte i starts a token).  `data`: bytes, or a
    contiguous uint8 array (no copy; orc_scan never reads past n)."""
    if isinstance(data, np.ndarray):
        t = np.ascontiguousarray(data, dtype=np.uint8)
        if t.size == 0:
            t = np.zeros(1, dtype=np.uint8)
    else:
        t = np.frombuffer(data + b"\0", dtype=np.uint8)
    n = len(data)
    order, blob, off = _specials(specials)
    cu = np.asarray(cuts if cuts is not None else [], dtype=np.uint64)
    out = np.zeros((n + 31) // 32 + 1, dtype=np.uint32)
    rc = lib().orc_scan(_p(t, C.c_uint8), C.c_uint64(n), _p(blob, C.c_uint8), _p(off, C.c_uint32), len(order),
                        _p(cu, C.c_uint64), C.c_uint64(len(cu)), C.c_uint32(flags), _p(out, C.c_uint32))
    if rc == 1:
        raise InvalidUtf8("invalid UTF-8")
    if rc:
        raise RuntimeError(f"orc_scan rc={rc}")
    return out[:(n + 31) // 32]


def count(data: bytes, specials=None, cuts=None, flags: int = 0) -> Counter:
    """The reference's pretoken Counter (training mode: specials dropped) for valid UTF-8 `data`."""
    n = len(data)
    start = scan(data, specials, cuts, flags)
    t = np.frombuffer(data + b"\0", dtype=np.uint8)
    order, blob, off = _specials(specials)
    cap = max(1024, n // 2 + 16)
    while True:
        o = np.zeros(cap, dtype=np.uint64)
        ln = np.zeros(cap, dtype=np.uint32)
        ct = np.zeros(cap, dtype=np.uint64)
        k = lib().orc_count(_p(t, C.c_uint8), C.c_uint64(n), _p(start, C.c_uint32), _p(blob, C.c_uint8), _p(off, C.c_uint32),
                            len(order), _p(o, C.c_uint64), _p(ln, C.c_uint32), _p(ct, C.c_uint64), C.c_uint64(cap))
        if k == -1:
            cap *= 2
            continue
        if k < 0:
            raise MemoryError
        break
    out: Counter = Counter()
    for i in range(k):
        a = int(o[i])
        out[tuple(data[a:a + int(ln[i])])] = int(ct[i])
    return out


def train(words: Counter, vocab_size: int, specials):
    """(vocab, merges) from a pretoken Counter, by SURVEY.md Appendix A.2."""
    sp = list(dict.fromkeys(specials or []))
    vocab = {b: bytes([b]) for b in range(256)}
    for k, s in enumerate(sp, start=256):
        vocab[k] = s.encode("utf-8")
    first_new = len(vocab)
    keys = [w for w in words if len(w) > 0]
    w_off = np.zeros(len(keys) + 1, dtype=np.uint64)
    w_off[1:] = np.cumsum([len(w) for w in keys], dtype=np.uint64) if keys else []
    sym = np.fromiter((b for w in keys for b in w), dtype=np.uint32, count=int(w_off[-1]))
    sym = np.concatenate([sym, np.zeros(1, dtype=np.uint32)])
    cnt = np.asarray([words[w] for w in keys] + [0], dtype=np.uint64)
    blob = b"".join(vocab[k] for k in range(first_new))
    tb = np.frombuffer(blob + b"\0", dtype=np.uint8)
    toff = np.zeros(first_new + 1, dtype=np.uint32)
    toff[1:] = np.cumsum([len(vocab[k]) for k in range(first_new)], dtype=np.uint32)
    cap = max(1, vocab_size - first_new)
    ma = np.zeros(cap, dtype=np.uint32)
    mb = np.zeros(cap, dtype=np.uint32)
    nm = C.c_uint32(0)
    has_pair = any(len(w) > 1 for w in keys)
    if not has_pair:
        raise AttributeError("no byte pairs to merge")  # bpe_trainer.py:237-240
    rc = lib().orc_train(C.c_uint64(len(keys)), _p(w_off, C.c_uint64), _p(sym, C.c_uint32), _p(cnt, C.c_uint64),
                         _p(tb, C.c_uint8), _p(toff, C.c_uint32), C.c_uint32(first_new), C.c_uint32(vocab_size),
                         _p(ma, C.c_uint32), _p(mb, C.c_uint32), C.byref(nm))
    if rc:
        raise RuntimeError(f"orc_train rc={rc}")
    merges = []
    for k in range(nm.value):
        a, b = vocab[int(ma[k])], vocab[int(mb[k])]
        merges.append((a, b))
        vocab[first_new + k] = a + b
    return vocab, merges


class COracleTokenizer:
    """Encoder half of the oracle: same constructor arguments as the reference's BPETokenizer."""

    def __init__(self, vocab, merges, specials=None):
        self.vocab = vocab
        self.specials = list(dict.fromkeys(specials or []))
        id_of = {v: k for k, v in vocab.items()}  # later id wins (bpe_tokenizer.py:85)
        self.id_of = id_of
        rank = {m: k for k, m in enumerate(merges)}  # later duplicate wins (:55)
        rows = sorted(rank.items(), key=lambda kv: kv[1])
        a, b, m = [], [], []
        for (x, y), _ in rows:
            if x in id_of and y in id_of:
                a.append(id_of[x]); b.append(id_of[y]); m.append(id_of.get(x + y, 0xFFFFFFFF))
        self.m_a = np.asarray(a + [0], dtype=np.uint32)
        self.m_b = np.asarray(b + [0], dtype=np.uint32)
        self.m_id = np.asarray(m + [0], dtype=np.uint32)
        self.n_merges = len(a)
        self.byte_to_id = np.asarray([id_of.get(bytes([k]), 0xFFFFFFFF) for k in range(256)], dtype=np.uint32)

    def encode_array(self, data, cuts=None, flags: int = 0) -> np.ndarray:
        """ids (uint32 array) of `data` (bytes or a contiguous uint8 array).  Thread-safe: callers may run several
        ranges side by side (ctypes releases the GIL around the C calls)."""
        order, blob, off = _specials(self.specials)
        sid = np.asarray([self.id_of[s] for s in order] + [0], dtype=np.uint32)  # KeyError like :175
        start = scan(data, self.specials, cuts, flags)
        t = np.ascontiguousarray(data, dtype=np.uint8) if isinstance(data, np.ndarray) else np.frombuffer(data + b"\0", dtype=np.uint8)
        n = len(data)
        present = np.flatnonzero(np.bincount(t[:n], minlength=256))
        if any(self.byte_to_id[b] == 0xFFFFFFFF for b in present):
            raise KeyError("byte missing from vocab")
        if n == 0:
            return np.zeros(0, dtype=np.uint32)
        out = np.zeros(n + 1, dtype=np.uint32)
        n_out = C.c_uint64(0)
        rc = lib().orc_encode(_p(t, C.c_uint8), C.c_uint64(n), _p(start, C.c_uint32), _p(blob, C.c_uint8),
                              _p(off, C.c_uint32), len(order), _p(sid, C.c_uint32), _p(self.byte_to_id, C.c_uint32),
                              _p(self.m_a, C.c_uint32), _p(self.m_b, C.c_uint32), _p(self.m_id, C.c_uint32),
                              C.c_uint32(self.n_merges), _p(out, C.c_uint32), C.c_uint64(len(out)), C.byref(n_out))
        if rc == 3:
            raise KeyError("merged bytes missing from vocab")
        if rc:
            raise RuntimeError(f"orc_encode rc={rc}")
        return out[:n_out.value]

    def encode_bytes(self, data: bytes, cuts=None, flags: int = 0) -> list[int]:
        return self.encode_array(data, cuts, flags).tolist()

    def encode(self, text: str) -> list[int]:
        return self.encode_bytes(text.encode("utf-8"))

    def encode_lines(self, text: str) -> list[int]:
        """encode_iterable over an iterable of lines (a file handle): a cut after every newline."""
        return self.encode_bytes(text.encode("utf-8"), flags=CUT_AFTER_NEWLINE)
