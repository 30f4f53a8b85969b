"""BPETokenizer with the reference's interface (bpe_transformer/tokenization/bpe_tokenizer.py:9-423); encoding
runs in csrc/scan.cu, count.cu and encode.cu through the C-ABI (include/bpe_b200.h)."""
from __future__ import annotations

import io
import itertools
import os
import pickle
from collections.abc import Iterable, Iterator
from pathlib import Path

import numpy as np

from .. import _lib
from ..settings import ENCODING_STD
from .tokenizer import Tokenizer

_MISSING = 0xFFFFFFFF
_BATCH_BYTES = 64 << 20  # encode_iterable hands the GPU about this much text at a time


_PIPE_BYTES = int(os.environ.get("BPE_B200_PIPE_BYTES", str(128 << 20)))  # largest piece of the load / encode / store pipeline
_PIPE_MIN_BYTES = 16 << 20  # smallest: a rank's share of a file is cut into about eight pieces, so that the three stages overlap
_FIRST_BATCH_BYTES = 1 << 20  # encode_iterable over arbitrary chunks: first ids after ~1 MB of text, batches double up to _BATCH_BYTES
_serials = itertools.count(1)  # one number per tokenizer object, never reused (id() is: CPython recycles addresses)


def _cuts_after_newlines(path, lo: int, hi: int, piece: int) -> list[int]:
    """Offsets lo = o0 < o1 < ... = hi with every inner offset just behind the first newline at or after lo + k * piece."""
    offs = [lo]
    with open(path, "rb") as fh:
        target = lo + piece
        while target < hi:
            pos = max(target, offs[-1])
            fh.seek(pos)
            found = hi
            while pos < hi:
                window = fh.read(min(1 << 16, hi - pos))
                if not window:
                    break
                i = window.find(b"\n")
                if i != -1:
                    found = pos + i + 1
                    break
                pos += len(window)
            if found >= hi:
                break
            if found > offs[-1]:
                offs.append(found)
            target = max(target + piece, found)
    offs.append(hi)
    return offs


class BPETokenizer(Tokenizer):
    """Greedy BPE encoder / decoder.

    >>> tok = BPETokenizer.from_files(Path("vocab.pkl"), Path("merges.pkl"), special_tokens=["<|endoftext|>"])
    >>> ids = tok.encode("Hello world!")
    >>> tok.decode(ids)
    """

    def __init__(self, vocab: dict[int, bytes], merges: list[tuple[bytes, bytes]], special_tokens: list[str] | None = None):
        self._vocab = vocab
        self._merges = merges
        self._special_tokens = set(special_tokens) if special_tokens else set()
        self._bytes_to_id_cache = None
        self._merges_priority = {merge: i for i, merge in enumerate(merges)}  # later duplicate wins (:55)
        self._device_model = None
        self._serial = next(_serials)  # names this object's model in the engine: a loaded model is reused only for the same object

    # -- the reference's accessors ------------------------------------------------------------------------
    @property
    def vocab(self) -> dict[int, bytes]:
        return self._vocab

    @property
    def merges(self) -> list[tuple[bytes, bytes]]:
        return self._merges

    @property
    def special_tokens(self) -> list[str] | None:
        return list(self._special_tokens)

    @property
    def _bytes_to_id(self) -> dict[bytes, int]:
        if self._bytes_to_id_cache is None:
            self._bytes_to_id_cache = {v: k for k, v in self._vocab.items()}  # later id wins (:85)
        return self._bytes_to_id_cache

    @classmethod
    def from_files(cls, vocab_filepath: Path, merges_filepath: Path, special_tokens: list[str] | None = None) -> "BPETokenizer":
        return cls(vocab=cls.load_vocab(file_path=vocab_filepath, special_tokens=special_tokens),
                   merges=cls.load_merges(merges_filepath), special_tokens=special_tokens)

    @staticmethod
    def load_vocab(file_path: Path, special_tokens: list[str] | None) -> dict[int, bytes]:
        """Unpickle; special tokens missing from the file are appended at id len(vocab) (:292-320)."""
        with open(file_path, "rb") as f:
            vocab: dict[int, bytes] = pickle.load(f)
        if not special_tokens:
            return vocab
        present = set(vocab.values())
        for t in special_tokens:
            raw = t.encode(ENCODING_STD)
            if raw not in present:
                vocab[len(vocab)] = raw
        return vocab

    @staticmethod
    def load_merges(file_path: Path) -> list[tuple[bytes, bytes]]:
        with open(file_path, "rb") as f:
            return pickle.load(f)

    # -- device model -------------------------------------------------------------------------------------
    def _model(self):
        """Merge table over canonical ids: what the reference's two dicts resolve to (:55, :85, :274-290)."""
        if self._device_model is None:
            b2i = self._bytes_to_id
            rows = sorted(self._merges_priority.items(), key=lambda kv: kv[1])
            a, b, m = [], [], []
            for (x, y), _ in rows:
                ia, ib = b2i.get(x), b2i.get(y)
                if ia is None or ib is None:
                    continue  # can never match: every id in play comes out of bytes_to_id
                a.append(ia); b.append(ib); m.append(b2i.get(x + y, _MISSING))
            byte_to_id = [b2i.get(bytes([k]), _MISSING) for k in range(256)]
            specials = _lib.sort_specials(self._special_tokens)
            special_id = [b2i.get(s, _MISSING) for s in specials]
            wide = max(self._vocab.keys(), default=0) > 0xFFFF
            self._device_model = (np.asarray(a, dtype=np.uint32), np.asarray(b, dtype=np.uint32), np.asarray(m, dtype=np.uint32),
                                  np.asarray(byte_to_id, dtype=np.uint32), special_id, specials, wide)
        return self._device_model

    def _bind(self, engine: "_lib.Engine"):
        a, b, m, byte_to_id, special_id, specials, wide = self._model()
        engine.set_special_tokens(specials)
        engine.load_model((self._serial, len(self._merges), len(self._vocab)), a, b, m, byte_to_id, special_id)
        return wide

    def encode_bytes_to_device(self, data, cuts=None, flags: int = 0, engine: "_lib.Engine | None" = None, on_device=None):
        """UTF-8 bytes (bytes / pinned uint8 tensor) -> device tensor of ids (uint16 bit pattern in int16, or int32).
        `on_device` = (padded device tensor, length) skips the upload."""
        engine = engine or _lib.get_engine()
        d_text, n = on_device if on_device is not None else engine.to_device(data)
        while True:
            wide = self._bind(engine)
            try:
                return engine.encode(d_text, n, cuts=cuts, flags=flags, wide=wide)
            except _lib.BpeError as e:
                if e.code == _lib.ERR_MISSING_ID:  # bpe_tokenizer.py:175, :206, :290
                    raise KeyError("bytes without an id in the vocabulary") from e
                if e.code != _lib.ERR_CAPACITY or not engine.grow():  # more distinct pretokens than the cache holds: enlarge it
                    raise

    @staticmethod
    def _ids_to_numpy(ids) -> np.ndarray:
        h = ids.cpu().numpy()
        return h.view(np.uint16) if h.dtype == np.int16 else h.view(np.uint32)

    # -- the reference's methods ----------------------------------------------------------------------------
    def encode(self, text: str) -> list[int]:
        """Special tokens -> their id; everything else -> pretokens -> byte ids -> merges in rank order (:139-188)."""
        data = text.encode(ENCODING_STD)
        if not data:
            return []
        return self._ids_to_numpy(self.encode_bytes_to_device(data)).tolist()

    def decode(self, ids: list[int]) -> str:
        """Concatenate vocab[id] (U+FFFD for unknown ids), decode with errors="replace" (:119-137)."""
        replacement = b"\xef\xbf\xbd"
        get = self._vocab.get
        return b"".join(get(i, replacement) for i in ids).decode(ENCODING_STD, errors="replace")

    def decode_array(self, ids, engine: "_lib.Engine | None" = None, to_host: bool = True):
        """GPU decode of a device / numpy id array to the concatenated bytes (no UTF-8 repair): `bytes`, or the
        uint8 device tensor with to_host=False."""
        import torch

        engine = engine or _lib.get_engine()
        if not isinstance(ids, torch.Tensor):
            arr = np.ascontiguousarray(ids)
            arr = arr.astype(np.uint16).view(np.int16) if arr.dtype.itemsize <= 2 else arr.astype(np.uint32).view(np.int32)
            ids = torch.from_numpy(arr).to(f"cuda:{engine.device}")
        n_vocab = max(self._vocab.keys(), default=-1) + 1
        lens = np.zeros(n_vocab + 1, dtype=np.uint64)
        for k, v in self._vocab.items():
            lens[k + 1] = len(v)
        off = np.cumsum(lens, dtype=np.uint64)
        blob = np.frombuffer(b"".join(self._vocab.get(k, b"") for k in range(n_vocab)) + b"\0", dtype=np.uint8).copy()
        out = engine.decode(ids, blob, off)
        return out.cpu().numpy().tobytes() if to_host else out

    def encode_iterable(self, iterable: Iterable[str], n_workers: int | None = None) -> Iterator[int]:
        """Lazily encode an iterable of strings (:339-423).  Text is buffered, cut after the LAST newline of the
        buffer each time a chunk arrives, and every cut piece is encoded on its own — so over a file handle
        every line is a segment.  `n_workers` is accepted and ignored (one GPU does the work)."""
        if isinstance(iterable, io.TextIOBase):
            yield from self._encode_lines_of(iterable)
            return
        pieces: list[bytes] = []
        cuts: list[int] = []
        total = 0
        pending = ""
        batch = _FIRST_BATCH_BYTES  # small first batches keep the generator lazy; large ones later keep the GPU busy
        for chunk in iterable:
            pending += chunk
            cut = pending.rfind("\n")
            if cut == -1:
                continue
            raw = pending[:cut + 1].encode(ENCODING_STD)
            pending = pending[cut + 1:]
            pieces.append(raw)
            total += len(raw)
            cuts.append(total)
            if total >= batch:
                yield from self._ids_to_numpy(self.encode_bytes_to_device(b"".join(pieces), cuts=cuts[:-1])).tolist()
                pieces, cuts, total = [], [], 0
                batch = min(batch * 2, _BATCH_BYTES)
        if pending:
            raw = pending.encode(ENCODING_STD)
            pieces.append(raw)
            total += len(raw)
            cuts.append(total)
        if total:
            yield from self._ids_to_numpy(self.encode_bytes_to_device(b"".join(pieces), cuts=cuts[:-1])).tolist()

    def _encode_lines_of(self, fh: io.TextIOBase) -> Iterator[int]:
        """A text file handle iterates by lines, so every newline is a cut: read big blocks instead of lines and
        let the kernel cut after each newline (same ids, no per-line Python work)."""
        tail = ""
        batch = _FIRST_BATCH_BYTES
        while True:
            block = fh.read(batch)
            batch = min(batch * 2, _BATCH_BYTES)
            if not block:
                break
            block = tail + block
            cut = block.rfind("\n")
            if cut == -1:
                tail = block
                continue
            tail = block[cut + 1:]
            ids = self.encode_bytes_to_device(block[:cut + 1].encode(ENCODING_STD), flags=_lib.CUT_AFTER_NEWLINE)
            yield from self._ids_to_numpy(ids).tolist()
        if tail:
            yield from self._ids_to_numpy(self.encode_bytes_to_device(tail.encode(ENCODING_STD), flags=_lib.CUT_AFTER_NEWLINE)).tolist()

    def encode_file_to_array(self, path, cut_after_newline: bool = True, engine: "_lib.Engine | None" = None,
                             lo: int = 0, hi: int | None = None, out: np.ndarray | None = None) -> np.ndarray:
        """Array-returning sibling of encode_iterable(open(path)): the UTF-8 file (or its bytes [lo, hi), which must
        start and end at line ends) -> numpy uint16 / uint32 ids (the format a 1-D token dataset uses).  Files
        with '\\r' go through Python's universal-newline translation first, as open(path) would do.
        Large files move in pieces cut after a newline: one host thread reads the next piece into HBM and another
        copies the previous piece's ids out while the GPU encodes the current one.
        `out`: a 1-D C-contiguous array of the id dtype (uint16, or uint32 for vocabularies beyond 65535) to write
        into instead of a new array; the result is `out[:number of ids]`, ValueError if it is too small.  With a
        page-locked `out` (`_lib.pinned_array`, reused across calls) the ids arrive by DMA without any host copy."""
        engine = engine or _lib.get_engine()
        size = os.path.getsize(path)
        hi = size if hi is None else min(hi, size)
        if out is not None:
            want = np.uint32 if self._model()[6] else np.uint16
            if out.ndim != 1 or out.dtype != want or not out.flags.c_contiguous or not out.flags.writeable:
                raise ValueError(f"out must be a writable 1-D C-contiguous {np.dtype(want).name} array")
        piece = min(_PIPE_BYTES, max(_PIPE_MIN_BYTES, (hi - lo) // 8))
        if cut_after_newline and hi - lo > piece + piece // 2:
            ids = self._encode_file_pipelined(path, lo, hi, engine, out, piece)
            if ids is not None:
                return ids
        return self.encode_device_to_array(engine.load_file(path, lo, hi), engine, cut_after_newline, out)

    def _encode_file_pipelined(self, path, lo: int, hi: int, engine: "_lib.Engine", out: np.ndarray | None = None, piece: int = 0):
        """None if the file holds a '\\r' (the caller then takes the whole-file path with newline translation)."""
        import queue
        import threading

        import torch

        offs = _cuts_after_newlines(path, lo, hi, piece or _PIPE_BYTES)
        wide = self._bind(engine)
        if out is None:
            out = _lib.host_array(hi - lo, np.uint32 if wide else np.uint16)  # one id per byte at most; unused pages stay untouched
        to_store: queue.Queue = queue.Queue(maxsize=2)
        errors: list[BaseException] = []
        stop = threading.Event()

        def storer():
            failed = False
            torch.cuda.set_device(engine.device)
            while True:
                item = to_store.get()
                if item is None:
                    return
                if failed:
                    continue
                try:
                    ids, at = item
                    if at + int(ids.numel()) > out.size:
                        raise ValueError("out is too small for the ids of this file")
                    engine.store_into(ids, out[at:at + int(ids.numel())])
                except BaseException as e:  # noqa: BLE001 - re-raised on the calling thread
                    errors.append(e)
                    failed = True
                    stop.set()

        store_thread = threading.Thread(target=storer, daemon=True)
        store_thread.start()
        load = engine.load_file_ranges(path, offs)  # one pass of the I/O lanes over the whole range, on its own thread
        total, has_cr = 0, False
        try:
            for k in range(len(offs) - 1):
                if stop.is_set():
                    break
                d_text, n = load.piece(k)  # blocks until piece k has landed; later pieces keep arriving meanwhile
                if not n:
                    continue
                try:  # the '\r' check runs on the GPU inside the call: no small device-to-host copy queues behind the ids' DMA
                    ids = self.encode_bytes_to_device(None, flags=_lib.CUT_AFTER_NEWLINE | _lib.FAIL_ON_CR, engine=engine, on_device=(d_text, n))
                except _lib.CarriageReturn:
                    has_cr = True
                    break
                to_store.put((ids, total))
                total += int(ids.numel())
        except BaseException as e:  # noqa: BLE001
            errors.insert(0, e)
        finally:
            to_store.put(None)
            store_thread.join()
            try:
                load.join()
            except BaseException as e:  # noqa: BLE001
                errors.append(e)
        if errors:
            raise errors[0]
        return None if has_cr else out[:total]

    def encode_device_to_array(self, on_device, engine: "_lib.Engine", cut_after_newline: bool = True,
                               out: np.ndarray | None = None) -> np.ndarray:
        """(padded device text, length) -> NumPy ids, cutting after every newline like a file handle's line iteration."""
        d_text, n = on_device
        if n and bool((d_text[:n] == 13).any().item()):  # '\r': what text mode translates before the reference sees lines
            data = d_text[:n].cpu().numpy().tobytes().replace(b"\r\n", b"\n").replace(b"\r", b"\n")
            on_device = engine.to_device(data)
        ids = self.encode_bytes_to_device(None, flags=_lib.CUT_AFTER_NEWLINE if cut_after_newline else 0, engine=engine, on_device=on_device)
        if out is None:
            return engine.to_host_array(ids, np.uint16 if ids.element_size() == 2 else np.uint32)
        k = int(ids.numel())
        if k > out.size:
            raise ValueError("out is too small for the ids of this file")
        engine.store_into(ids, out[:k])
        return out[:k]
