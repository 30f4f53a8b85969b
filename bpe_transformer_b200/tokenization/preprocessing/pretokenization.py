"""Pretokenisation on the GPU behind the reference's function names
(reference: bpe_transformer/tokenization/preprocessing/pretokenization.py:41-298).

What runs where:
  * host: read the file into pinned memory, copy it to HBM; for a corpus WITHOUT special tokens also compute
    the reference's worker chunk offsets (they are segment boundaries there and so change the result);
  * GPU (csrc/fused.cu; csrc/scan.cu for callers that want the start mask): special-token split, the GPT-2 pattern,
    the Counter of pretokens, in one pass over the text;
  * host: only when a Counter is asked for, the distinct pretokens (10^4..10^7 of them) are copied back and
    turned into the reference's `Counter[tuple[int, ...], int]`.
`n_workers` keeps its meaning for the chunk offsets only; the work itself is one GPU's.
"""
from __future__ import annotations

import os
from collections import Counter
from multiprocessing import cpu_count
from pathlib import Path
from typing import BinaryIO

import numpy as np

from ... import _lib
from ...settings import ENCODING_STD


_PIPE_BYTES = int(os.environ.get("BPE_B200_COUNT_PIPE_BYTES", str(256 << 20)))  # piece size of the load / count pipeline


# --------------------------------------------------------------------------------------------------------
# host helpers that mirror reference functions
# --------------------------------------------------------------------------------------------------------
def find_chunk_boundaries(file: BinaryIO, desired_num_chunks: int, special_tokens: list[str] | None = None) -> list[int]:
    """Offsets at which the reference hands byte ranges to its workers (pretokenization.py:114-168): uniform
    guesses moved forward, by 4 KiB probes, to the start of the next special token (b"\\n" if there are none)."""
    needles = [t.encode(ENCODING_STD) for t in special_tokens] if special_tokens else [b"\n"]
    file.seek(0, os.SEEK_END)
    size = file.tell()
    stride = size // desired_num_chunks
    offsets = [k * stride for k in range(desired_num_chunks + 1)]
    offsets[-1] = size
    for k in range(1, desired_num_chunks):
        pos = offsets[k]
        file.seek(pos)
        while True:
            window = file.read(4096)
            if not window:
                offsets[k] = size
                break
            found = [i for i in (window.find(nd) for nd in needles) if i != -1]
            if found:
                offsets[k] = pos + min(found)
                break
            pos += 4096
    return sorted(set(offsets))


def split_on_special_tokens(text: str, training: bool | None = True, special_tokens: list[str] | None = None) -> list[str]:
    """Host-side equivalent of pretokenization.py:211-235 for callers that want the parts as strings
    (the GPU path does this split inside scan.cu).  Longest special first, leftmost, non-overlapping;
    `training` drops the specials, otherwise they are kept as their own parts."""
    if not special_tokens:
        return [text]
    order = sorted(special_tokens, key=len, reverse=True)
    parts, seg, i, n = [], 0, 0, len(text)
    firsts = {t[0] for t in order if t}
    while i < n:
        if text[i] in firsts:
            hit = next((t for t in order if t and text.startswith(t, i)), None)
            if hit is not None:
                parts.append(text[seg:i])
                if not training:
                    parts.append(hit)
                i += len(hit)
                seg = i
                continue
        i += 1
    parts.append(text[seg:])
    return parts


def _clamp_workers(n_workers: int | None) -> int:
    if n_workers is None or n_workers == 0:  # pretokenization.py:89-93
        n_workers = 4
    return min(n_workers, cpu_count())


def read_file_pinned(path) -> "object":
    """Whole file into a pinned uint8 tensor (the H2D copy then runs at PCIe speed)."""
    import torch

    size = os.path.getsize(path)
    host = torch.empty(size, dtype=torch.uint8, pin_memory=torch.cuda.is_available())
    if size:
        view = memoryview(host.numpy())
        with open(path, "rb", buffering=0) as fh:
            got = 0
            while got < size:
                k = fh.readinto(view[got:])
                if not k:
                    raise OSError(f"short read on {path}")
                got += k
    return host


def _sanitised(data: bytes, cuts: list[int], strict: bool) -> tuple[bytes, list[int]]:
    """What the reference's workers see when the bytes are not valid UTF-8: every chunk is decoded on its own
    with errors="ignore" (pretokenization.py:194) — or strictly in the serial variant (:284), which raises."""
    edges = [0] + list(cuts) + [len(data)]
    out, new_cuts, total = [], [], 0
    for a, b in zip(edges[:-1], edges[1:]):
        piece = data[a:b].decode(ENCODING_STD) if strict else data[a:b].decode(ENCODING_STD, errors="ignore")
        raw = piece.encode(ENCODING_STD)
        out.append(raw)
        total += len(raw)
        new_cuts.append(total)
    return b"".join(out), new_cuts[:-1]


def count_into_engine(engine: "_lib.Engine", file_path, special_tokens: list[str] | None, n_chunks: int,
                      strict: bool = False, reset: bool = True) -> None:
    """scan + count the whole file on the engine's GPU (the table of distinct pretokens stays in the context)."""
    specials = _lib.sort_specials(special_tokens)
    engine.set_special_tokens(specials)
    if reset:
        engine.unique_reset()
    cuts: list[int] = []
    if not specials:
        # Without special tokens the reference cuts the file at newlines chosen by the worker count; a cut is a
        # segment boundary, so reproduce the same offsets.  (With special tokens the cuts sit at special-token
        # starts, where a segment ends anyway.)
        with open(file_path, "rb") as fh:
            cuts = find_chunk_boundaries(fh, n_chunks, None)[1:-1]
    if specials and os.path.getsize(file_path) > _PIPE_BYTES + _PIPE_BYTES // 2:
        while True:
            done = _count_file_pipelined(engine, file_path, special_tokens)
            if done == "counted":
                return
            engine.unique_reset()
            if done == "invalid-utf8":
                break  # whole-file path below, which repairs the bytes on the host
            if not engine.grow():  # "capacity"
                raise _lib.BpeError(_lib.ERR_CAPACITY, "table of distinct pretokens cannot grow further")
    d_text, n = engine.load_file(file_path)
    try:
        count_with_growth(engine, d_text, n, cuts, keep_counted=not reset)
    except _lib.InvalidUtf8:
        if not reset:
            raise  # counts of this file are already mixed into the caller's table: nothing sane to keep
        with open(file_path, "rb") as fh:
            clean, cuts = _sanitised(fh.read(), cuts, strict)
        engine.unique_reset()  # the one-pass kernel had counted some of the dirty bytes before it reported them
        d_text, n = engine.to_device(clean)
        count_with_growth(engine, d_text, n, cuts)


def _count_file_pipelined(engine: "_lib.Engine", file_path, special_tokens: list[str]) -> str:
    """Large corpora with special tokens: pieces cut at special-token starts (the reference's own chunk rule, so the
    cuts cannot change the result) stream into HBM in one pass of the I/O lanes while the GPU scans and counts the
    pieces that have landed.  Returns "counted", or why the caller has to start again: "capacity" / "invalid-utf8"."""
    size = os.path.getsize(file_path)
    with open(file_path, "rb") as fh:
        offs = find_chunk_boundaries(fh, max(2, -(-size // _PIPE_BYTES)), special_tokens)
    return _count_pieces_pipelined(engine, file_path, offs)


def _count_pieces_pipelined(engine: "_lib.Engine", file_path, offs: list[int]) -> str:
    """The byte ranges [offs[k], offs[k+1]) of the file (cut at special-token starts), streamed and counted."""
    load = engine.load_file_ranges(file_path, offs)  # one pass of the I/O lanes over the whole range, on its own thread
    outcome = "counted"
    try:
        for k in range(len(offs) - 1):
            d_text, n = load.piece(k)  # blocks until piece k has landed; later pieces keep arriving meanwhile
            if n:
                engine.pretokenize_count(d_text, n)
    except _lib.InvalidUtf8:
        outcome = "invalid-utf8"
    except _lib.BpeError as e:
        if e.code != _lib.ERR_CAPACITY:
            raise
        outcome = "capacity"
    finally:
        try:
            load.join()
        except BaseException:  # noqa: BLE001 - a failed load matters only if counting got that far
            if outcome == "counted":
                raise
    return outcome


def count_with_growth(engine: "_lib.Engine", d_text, n: int, cuts=None, keep_counted: bool = False) -> None:
    """bpe_pretokenize_count (split + hash + count in one pass); if the context's table of distinct pretokens or its
    byte pool is too small for this corpus, enlarge it (x4) and count again.  `keep_counted`: earlier shards are in
    the table and must not be dropped -> no retry."""
    while True:
        try:
            engine.pretokenize_count(d_text, n, cuts=cuts)
            return
        except _lib.BpeError as e:
            if e.code != _lib.ERR_CAPACITY or keep_counted or not engine.grow():
                raise


def counter_from_engine(engine: "_lib.Engine") -> Counter:
    """The context's table as the reference's Counter[tuple[int, ...], int]."""
    blob, off, ln, cnt = engine.unique_export()
    blob_h = blob.cpu().numpy().tobytes()
    off_h, ln_h, cnt_h = off.cpu().numpy(), ln.cpu().numpy(), cnt.cpu().numpy()
    out: Counter = Counter()
    for o, l, c in zip(off_h.tolist(), ln_h.tolist(), cnt_h.tolist()):
        out[tuple(blob_h[o:o + l])] = c
    return out


# --------------------------------------------------------------------------------------------------------
# the reference's public functions
# --------------------------------------------------------------------------------------------------------
def pretokenize(file_path: Path, training: bool | None = True, parallel_processing: bool | None = True,
                n_workers: int | None = 4, special_tokens: list[str] | None = None) -> Counter:
    """pretokenization.py:41-70."""
    if parallel_processing:
        return parallel_pretokenization(file_path=file_path, training=training, n_workers=n_workers, special_tokens=special_tokens)
    return serial_pretokenization(file_path=file_path, training=training, special_tokens=special_tokens)


def parallel_pretokenization(file_path: Path, n_workers: int | None = None, training: bool | None = True,
                             special_tokens: list[str] | None = None) -> Counter:
    """pretokenization.py:73-111 — Counter of pretokens (keys: tuples of byte values)."""
    return _pretokenize_file(file_path, _clamp_workers(n_workers), training, special_tokens, strict=False)


def serial_pretokenization(file_path: Path, training: bool | None = True, special_tokens: list[str] | None = None) -> Counter:
    """pretokenization.py:255-298 — four chunks, strict UTF-8 decoding."""
    return _pretokenize_file(file_path, 4, training, special_tokens, strict=True)


def _pretokenize_file(file_path, n_chunks: int, training, special_tokens, strict: bool) -> Counter:
    engine = _lib.get_engine()
    if training or not special_tokens:
        count_into_engine(engine, file_path, special_tokens, n_chunks, strict=strict)
        return counter_from_engine(engine)
    # training=False keeps each special token as a part of its own and then pretokenizes that part like any
    # other text (pretokenization.py:198-207): boundaries before and after every special-token match.
    with open(file_path, "rb") as fh:
        offsets = find_chunk_boundaries(fh, n_chunks, special_tokens)
        fh.seek(0)
        data = fh.read()
    pieces, cuts, total = [], [], 0
    for a, b in zip(offsets[:-1], offsets[1:]):
        text = data[a:b].decode(ENCODING_STD) if strict else data[a:b].decode(ENCODING_STD, errors="ignore")
        for part in split_on_special_tokens(text, training=False, special_tokens=special_tokens):
            raw = part.encode(ENCODING_STD)
            if raw:
                pieces.append(raw)
                total += len(raw)
                cuts.append(total)
    engine.set_special_tokens([])
    engine.unique_reset()
    d_text, n = engine.to_device(b"".join(pieces))
    count_with_growth(engine, d_text, n, cuts[:-1])
    return counter_from_engine(engine)


def pretokenize_text(text: str) -> list[bytes]:
    """pretokenization.py:238-252 — the pattern's matches of one string, as bytes."""
    engine = _lib.get_engine()
    data = text.encode(ENCODING_STD)
    if not data:
        return []
    engine.set_special_tokens([])
    d_text, n = engine.to_device(data)
    mask = engine.scan(d_text, n).cpu().numpy().view(np.uint32)
    bits = np.unpackbits(mask.view(np.uint8), bitorder="little")[:n]
    starts = np.flatnonzero(bits).tolist() + [n]
    return [data[a:b] for a, b in zip(starts[:-1], starts[1:])]
