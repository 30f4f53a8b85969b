"""ctypes binding of libbpe_b200.so (include/bpe_b200.h) and the per-GPU engine the Python API sits on.

There is NO CPU fallback: if the library is missing or no CUDA device is present, everything here raises.
PyTorch is used for device buffers, pinned host buffers and the current stream only.
"""
from __future__ import annotations

import ctypes as C
import os
import pathlib
import threading
import time
import warnings

import numpy as np

CSRC = pathlib.Path(__file__).resolve().parent / "csrc"
LIB_PATH = CSRC / "libbpe_b200.so"

TEXT_PAD = 64
CUT_AFTER_NEWLINE = 1
FAIL_ON_CR = 2  # bpe_encode: error 9 if the text holds a '\r' (a text-mode file handle would have translated it)

OK, ERR_CUDA, ERR_INVALID_UTF8, ERR_CAPACITY, ERR_ARGUMENT, ERR_HASH_COLLISION, ERR_MISSING_ID, ERR_NO_PAIRS, ERR_STATE, ERR_CARRIAGE_RETURN = range(10)

SYMBOLS = [
    "bpe_ctx_create", "bpe_ctx_destroy", "bpe_last_error", "bpe_ctx_configure", "bpe_set_special_tokens", "bpe_scan",
    "bpe_count", "bpe_unique_stats", "bpe_unique_export", "bpe_unique_import", "bpe_unique_reset", "bpe_train",
    "bpe_train_timing", "bpe_load_model", "bpe_encode", "bpe_decode", "bpe_kernel_times", "bpe_set_profiling",
    "bpe_decode_noalloc", "bpe_load_file", "bpe_store_to_host", "bpe_load_file_ranges", "bpe_pretokenize_count",
    "bpe_peer_alloc", "bpe_peer_connect", "bpe_peer_disconnect", "bpe_debug_key_mask",
]


class BpeError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"libbpe_b200 error {code}: {message}")
        self.code = code


class InvalidUtf8(BpeError):
    pass


class CarriageReturn(BpeError):
    pass


class KernelTime(C.Structure):
    _fields_ = [("name", C.c_char_p), ("ms", C.c_double), ("launches", C.c_uint64)]


_lib = None
_lib_lock = threading.Lock()


def load_library():
    """dlopen the in-tree library; fails loudly when it has not been built (python __graft_entry__.py build)."""
    global _lib
    with _lib_lock:
        if _lib is not None:
            return _lib
        if not LIB_PATH.exists():
            raise ImportError(f"{LIB_PATH} is missing: build it with `make -C {CSRC}` (or __graft_entry__.build()); "
                              "there is no CPU fallback")
        lib = C.CDLL(str(LIB_PATH))
        vp, u8p, u32p, u64p = C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p
        lib.bpe_ctx_create.argtypes = [C.c_int, C.POINTER(vp)]
        lib.bpe_ctx_destroy.argtypes = [vp]
        lib.bpe_ctx_destroy.restype = None
        lib.bpe_last_error.argtypes = [vp]
        lib.bpe_last_error.restype = C.c_char_p
        lib.bpe_ctx_configure.argtypes = [vp, C.c_int, C.c_uint64]
        lib.bpe_set_special_tokens.argtypes = [vp, u8p, u32p, C.c_int]
        lib.bpe_scan.argtypes = [vp, u8p, C.c_uint64, u64p, C.c_uint64, C.c_uint32, u32p, vp]
        lib.bpe_count.argtypes = [vp, u8p, C.c_uint64, u32p, vp]
        lib.bpe_pretokenize_count.argtypes = [vp, u8p, C.c_uint64, u64p, C.c_uint64, C.c_uint32, u32p, vp]
        lib.bpe_unique_stats.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        lib.bpe_unique_export.argtypes = [vp, u8p, u64p, u32p, u64p, vp]
        lib.bpe_unique_import.argtypes = [vp, u8p, u64p, u32p, u64p, C.c_uint64, vp]
        lib.bpe_unique_reset.argtypes = [vp]
        lib.bpe_train.argtypes = [vp, C.c_uint32, C.c_uint32, u8p, u32p, u32p, u32p, C.POINTER(C.c_uint32), vp]
        lib.bpe_train_timing.argtypes = [vp, C.POINTER(C.c_double)]
        lib.bpe_peer_alloc.argtypes = [vp, C.c_uint32, u8p]
        lib.bpe_peer_connect.argtypes = [vp, C.c_int, C.c_int, u8p]
        lib.bpe_peer_disconnect.argtypes = [vp]
        lib.bpe_debug_key_mask.argtypes = [vp, C.c_uint64]
        lib.bpe_load_model.argtypes = [vp, u32p, u32p, u32p, C.c_uint32, u32p, u32p]
        lib.bpe_encode.argtypes = [vp, u8p, C.c_uint64, u64p, C.c_uint64, C.c_uint32, u32p, vp, C.c_int, C.c_uint64,
                                   C.POINTER(C.c_uint64), vp]
        lib.bpe_decode.argtypes = [vp, vp, C.c_int, C.c_uint64, u8p, u64p, C.c_uint32, u8p, C.c_uint64, C.POINTER(C.c_uint64), vp]
        lib.bpe_decode_noalloc.argtypes = [vp, C.c_int, C.c_uint64, u8p, u64p, C.c_uint32, u8p, C.c_uint64, C.POINTER(C.c_uint64), vp,
                                           C.c_char_p, C.c_size_t]
        lib.bpe_load_file.argtypes = [vp, C.c_char_p, C.c_uint64, C.c_uint64, u8p, vp]
        lib.bpe_store_to_host.argtypes = [vp, vp, C.c_uint64, vp, vp]
        lib.bpe_load_file_ranges.argtypes = [vp, C.c_char_p, C.c_uint32, u64p, u64p, u64p, u8p, C.POINTER(C.c_uint64), vp]
        lib.bpe_kernel_times.argtypes = [vp, C.POINTER(KernelTime), C.c_int]
        lib.bpe_set_profiling.argtypes = [vp, C.c_int]
        for name in SYMBOLS:
            fn = getattr(lib, name)
            if name not in ("bpe_ctx_destroy", "bpe_last_error"):
                fn.restype = C.c_int
        _lib = lib
        return lib


def _np_ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def sort_specials(special_tokens) -> list[bytes]:
    """Match priority of the reference's split_on_special_tokens (pretokenization.py:228): longest first,
    ties in the given order (sorted() is stable)."""
    toks = [s.encode("utf-8") if isinstance(s, str) else bytes(s) for s in (special_tokens or [])]
    return sorted(toks, key=len, reverse=True)


class Engine:
    """One libbpe_b200 context on one GPU plus the torch buffers around it."""

    def __init__(self, device: int | None = None, table_log2: int | None = None, pool_bytes: int | None = None):
        import torch

        if not torch.cuda.is_available():
            raise RuntimeError("bpe_transformer_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
        self.torch = torch
        self.lib = load_library()
        self.device = torch.cuda.current_device() if device is None else int(device)
        torch.cuda.set_device(self.device)
        torch.cuda.init()
        torch.zeros(1, device=f"cuda:{self.device}")  # make sure the primary context exists before the library uses it
        ctx = C.c_void_p()
        rc = self.lib.bpe_ctx_create(self.device, C.byref(ctx))
        if rc:
            raise BpeError(rc, (self.lib.bpe_last_error(None) or b"").decode())
        self.ctx = ctx
        self._specials: list[bytes] | None = None
        self._model_token = None
        self.table_log2, self.pool_bytes = 22, 256 << 20
        if table_log2 is not None or pool_bytes is not None:
            self.configure(table_log2 or 22, pool_bytes or (256 << 20))

    # -- plumbing ----------------------------------------------------------------------------------------
    def close(self):
        if getattr(self, "ctx", None):
            self.lib.bpe_ctx_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc: int):
        if rc == OK:
            return
        msg = (self.lib.bpe_last_error(self.ctx) or b"").decode()
        if rc == ERR_INVALID_UTF8:
            raise InvalidUtf8(rc, msg)
        if rc == ERR_CARRIAGE_RETURN:
            raise CarriageReturn(rc, msg)
        raise BpeError(rc, msg)

    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream(self.device).cuda_stream)

    def configure(self, table_log2: int, pool_bytes: int):
        self._check(self.lib.bpe_ctx_configure(self.ctx, int(table_log2), int(pool_bytes)))
        self.table_log2, self.pool_bytes = int(table_log2), int(pool_bytes)
        self._model_token = None

    def grow(self) -> bool:
        """Four times the table slots and pool bytes (drops what was counted); False once 2^28 slots are reached."""
        if self.table_log2 >= 28:
            return False
        self.configure(self.table_log2 + 2, self.pool_bytes * 4)
        return True

    def set_profiling(self, on: bool):
        self._check(self.lib.bpe_set_profiling(self.ctx, 1 if on else 0))

    def kernel_times(self) -> dict[str, tuple[float, int]]:
        buf = (KernelTime * 32)()
        n = self.lib.bpe_kernel_times(self.ctx, buf, 32)
        return {buf[i].name.decode(): (buf[i].ms, int(buf[i].launches)) for i in range(n)}

    # -- buffers -----------------------------------------------------------------------------------------
    def to_device(self, data) -> tuple["object", int]:
        """bytes / bytearray / numpy uint8 / pinned torch uint8 -> padded device tensor and its length."""
        torch = self.torch
        if isinstance(data, torch.Tensor):
            host = data
        else:
            with warnings.catch_warnings():  # bytes are read-only; the tensor is only ever a copy source
                warnings.simplefilter("ignore", UserWarning)
                host = torch.frombuffer(memoryview(data), dtype=torch.uint8) if len(data) else torch.empty(0, dtype=torch.uint8)
        n = int(host.numel())
        dev = torch.empty(n + TEXT_PAD, dtype=torch.uint8, device=f"cuda:{self.device}")
        if n:
            dev[:n].copy_(host, non_blocking=True)
        return dev, n

    def load_file(self, path, lo: int = 0, hi: int | None = None) -> tuple["object", int]:
        """Bytes [lo, hi) of a file -> padded device tensor and its length (threaded reads through pinned staging)."""
        torch = self.torch
        size = os.path.getsize(path)  # FileNotFoundError like open()
        hi = size if hi is None else min(hi, size)
        n = max(0, hi - lo)
        dev = torch.empty(n + TEXT_PAD, dtype=torch.uint8, device=f"cuda:{self.device}")
        if n:
            self._check(self.lib.bpe_load_file(self.ctx, os.fsencode(path), lo, n, dev.data_ptr(), self._stream()))
        return dev, n

    def load_file_ranges(self, path, offsets: list[int]) -> "RangeLoad":
        """Starts loading the byte ranges [offsets[k], offsets[k+1]) of a file on a host thread, each range to its own
        16-byte aligned, padded place in ONE device buffer.  `RangeLoad.piece(k)` blocks until range k has landed."""
        return RangeLoad(self, path, offsets)

    def to_host_array(self, d_tensor, dtype) -> np.ndarray:
        """Device tensor -> new NumPy array of `dtype` (same bytes), through the pinned staging buffers."""
        nbytes = int(d_tensor.numel()) * d_tensor.element_size()
        out = host_array(nbytes // np.dtype(dtype).itemsize, dtype)
        if nbytes:
            self._check(self.lib.bpe_store_to_host(self.ctx, d_tensor.data_ptr(), nbytes, out.ctypes.data_as(C.c_void_p), self._stream()))
        return out

    def store_into(self, d_tensor, out: np.ndarray) -> None:
        """Device tensor -> an existing contiguous NumPy array of the same byte size."""
        nbytes = int(d_tensor.numel()) * d_tensor.element_size()
        if nbytes != out.nbytes or not out.flags.c_contiguous:
            raise ValueError("store_into: size or layout mismatch")
        if nbytes:
            self._check(self.lib.bpe_store_to_host(self.ctx, d_tensor.data_ptr(), nbytes, out.ctypes.data_as(C.c_void_p), self._stream()))

    def new_mask(self, n: int):
        return self.torch.empty(n // 32 + 2, dtype=self.torch.int32, device=f"cuda:{self.device}")

    # -- the C-ABI calls -----------------------------------------------------------------------------------
    def set_special_tokens(self, specials: list[bytes]):
        if self._specials == specials:
            return
        blob = np.frombuffer(b"".join(specials) + b"\0", dtype=np.uint8).copy()
        off = np.zeros(len(specials) + 1, dtype=np.uint32)
        if specials:
            off[1:] = np.cumsum([len(s) for s in specials])
        self._check(self.lib.bpe_set_special_tokens(self.ctx, _np_ptr(blob), _np_ptr(off), len(specials)))
        self._specials = list(specials)
        self._model_token = None  # special ids are part of the loaded model

    def scan(self, d_text, n: int, cuts=None, flags: int = 0):
        torch = self.torch
        mask = self.new_mask(n)
        d_cuts, n_cuts = None, 0
        if cuts is not None and len(cuts):
            d_cuts = torch.as_tensor(np.asarray(cuts, dtype=np.int64), device=f"cuda:{self.device}")
            n_cuts = int(d_cuts.numel())
        self._check(self.lib.bpe_scan(self.ctx, d_text.data_ptr(), n, d_cuts.data_ptr() if d_cuts is not None else None, n_cuts,
                                      flags, mask.data_ptr(), self._stream()))
        return mask

    def count(self, d_text, n: int, mask):
        self._check(self.lib.bpe_count(self.ctx, d_text.data_ptr(), n, mask.data_ptr(), self._stream()))

    def pretokenize_count(self, d_text, n: int, cuts=None, flags: int = 0, want_mask: bool = False):
        """scan + count in one pass (bpe_pretokenize_count); returns the start mask if asked for, else None."""
        torch = self.torch
        mask = self.new_mask(n) if want_mask else None
        d_cuts, n_cuts = None, 0
        if cuts is not None and len(cuts):
            d_cuts = torch.as_tensor(np.asarray(cuts, dtype=np.int64), device=f"cuda:{self.device}")
            n_cuts = int(d_cuts.numel())
        self._check(self.lib.bpe_pretokenize_count(self.ctx, d_text.data_ptr(), n, d_cuts.data_ptr() if d_cuts is not None else None, n_cuts,
                                                   flags, mask.data_ptr() if mask is not None else None, self._stream()))
        return mask

    def unique_stats(self) -> tuple[int, int, int]:
        a, b, c = C.c_uint64(), C.c_uint64(), C.c_uint64()
        self._check(self.lib.bpe_unique_stats(self.ctx, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def unique_reset(self):
        self._check(self.lib.bpe_unique_reset(self.ctx))

    def unique_export(self):
        """(blob uint8, off int64, len int32, cnt int64) device tensors."""
        torch = self.torch
        n_unique, n_bytes, _ = self.unique_stats()
        dev = f"cuda:{self.device}"
        blob = torch.zeros(n_bytes + TEXT_PAD, dtype=torch.uint8, device=dev)
        off = torch.zeros(max(n_unique, 1), dtype=torch.int64, device=dev)
        ln = torch.zeros(max(n_unique, 1), dtype=torch.int32, device=dev)
        cnt = torch.zeros(max(n_unique, 1), dtype=torch.int64, device=dev)
        if n_unique:
            self._check(self.lib.bpe_unique_export(self.ctx, blob.data_ptr(), off.data_ptr(), ln.data_ptr(), cnt.data_ptr(), self._stream()))
        return blob, off[:n_unique], ln[:n_unique], cnt[:n_unique]

    def unique_import(self, blob, off, ln, cnt):
        n = int(off.numel())
        if n:
            self._check(self.lib.bpe_unique_import(self.ctx, blob.data_ptr(), off.data_ptr(), ln.data_ptr(), cnt.data_ptr(), n, self._stream()))

    def train(self, vocab_size: int, init_tokens: list[bytes]) -> tuple[np.ndarray, np.ndarray]:
        first_new = len(init_tokens)
        blob = np.frombuffer(b"".join(init_tokens) + b"\0", dtype=np.uint8).copy()
        off = np.zeros(first_new + 1, dtype=np.uint32)
        off[1:] = np.cumsum([len(t) for t in init_tokens])
        cap = max(1, vocab_size - first_new)
        ma = np.zeros(cap, dtype=np.uint32)
        mb = np.zeros(cap, dtype=np.uint32)
        nm = C.c_uint32(0)
        self._check(self.lib.bpe_train(self.ctx, vocab_size, first_new, _np_ptr(blob), _np_ptr(off), _np_ptr(ma), _np_ptr(mb),
                                       C.byref(nm), self._stream()))
        return ma[:nm.value], mb[:nm.value]

    def debug_key_mask(self, mask: int):
        """Test hook: AND every pretoken key with `mask` (0xFFFF_FFFF_FFFF_FFFF = normal)."""
        self._check(self.lib.bpe_debug_key_mask(self.ctx, int(mask)))

    def peer_alloc(self, hits_per_sender: int) -> bytes:
        """This GPU's exchange buffer for shared heavy steps; returns the 64-byte handle the peers open it with."""
        h = np.zeros(64, dtype=np.uint8)
        self._check(self.lib.bpe_peer_alloc(self.ctx, int(hits_per_sender), _np_ptr(h)))
        return h.tobytes()

    def peer_connect(self, rank: int, handles: list[bytes]):
        blob = np.frombuffer(b"".join(handles), dtype=np.uint8).copy()
        self._check(self.lib.bpe_peer_connect(self.ctx, int(rank), len(handles), _np_ptr(blob)))

    def peer_disconnect(self):
        self._check(self.lib.bpe_peer_disconnect(self.ctx))

    def train_timing(self) -> tuple[float, float, int]:
        out = (C.c_double * 3)()
        self._check(self.lib.bpe_train_timing(self.ctx, out))
        return out[0], out[1], int(out[2])

    def load_model(self, token, m_a, m_b, m_id, byte_to_id, special_id):
        """`token` identifies the model; reloading the same one is skipped (it would drop the encoding cache)."""
        if self._model_token is not None and self._model_token == token:
            return
        m_a = np.ascontiguousarray(m_a, dtype=np.uint32)
        m_b = np.ascontiguousarray(m_b, dtype=np.uint32)
        m_id = np.ascontiguousarray(m_id, dtype=np.uint32)
        b2i = np.ascontiguousarray(byte_to_id, dtype=np.uint32)
        sid = np.ascontiguousarray(list(special_id) + [0], dtype=np.uint32)
        self._check(self.lib.bpe_load_model(self.ctx, _np_ptr(m_a), _np_ptr(m_b), _np_ptr(m_id), len(m_a), _np_ptr(b2i), _np_ptr(sid)))
        self._model_token = token

    def encode(self, d_text, n: int, cuts=None, flags: int = 0, wide: bool = False, mask=None):
        """Returns a device tensor of ids (int16 bit pattern = uint16, or int32)."""
        torch = self.torch
        dev = f"cuda:{self.device}"
        d_cuts, n_cuts = None, 0
        if cuts is not None and len(cuts):
            d_cuts = torch.as_tensor(np.asarray(cuts, dtype=np.int64), device=dev)
            n_cuts = int(d_cuts.numel())
        out = torch.empty(n + 1, dtype=torch.int32 if wide else torch.int16, device=dev)
        n_out = C.c_uint64(0)
        self._check(self.lib.bpe_encode(self.ctx, d_text.data_ptr(), n, d_cuts.data_ptr() if d_cuts is not None else None, n_cuts,
                                        flags, mask.data_ptr() if mask is not None else None, out.data_ptr(), 4 if wide else 2,
                                        n + 1, C.byref(n_out), self._stream()))
        return out[:n_out.value]

    def decode(self, d_ids, vocab_blob: np.ndarray, vocab_off: np.ndarray):
        """ids (device tensor, 2- or 4-byte elements) -> device tensor of the concatenated token bytes."""
        torch = self.torch
        n_ids = int(d_ids.numel())
        elem = d_ids.element_size()
        n_bytes = C.c_uint64(0)
        args = [self.ctx, d_ids.data_ptr(), elem, n_ids, _np_ptr(vocab_blob), _np_ptr(vocab_off), len(vocab_off) - 1]
        self._check(self.lib.bpe_decode(*args, None, 0, C.byref(n_bytes), self._stream()))  # size query
        out = torch.empty(n_bytes.value + 16, dtype=torch.uint8, device=f"cuda:{self.device}")
        self._check(self.lib.bpe_decode(*args, out.data_ptr(), n_bytes.value, C.byref(n_bytes), self._stream()))
        return out[:n_bytes.value]


class RangeLoad:
    """A running bpe_load_file_ranges call (one host thread) and the device buffer it fills."""

    def __init__(self, engine: Engine, path, offsets: list[int]):
        torch = engine.torch
        self.engine = engine
        lens = [b - a for a, b in zip(offsets[:-1], offsets[1:])]
        self.lens = lens
        dev_off, at = [], 0
        for n in lens:
            dev_off.append(at)
            at = (at + n + TEXT_PAD + 15) & ~15  # every piece: 16-byte aligned start, TEXT_PAD readable bytes behind it
        self.dev_off = dev_off
        self.buffer = torch.empty(at + TEXT_PAD, dtype=torch.uint8, device=f"cuda:{engine.device}")
        self._file_off = np.asarray(offsets[:-1], dtype=np.uint64)
        self._len = np.asarray(lens, dtype=np.uint64)
        self._dev_off = np.asarray(dev_off, dtype=np.uint64)
        self._done = C.c_uint64(0)
        self._rc: int | None = None
        self._error: BaseException | None = None
        self._path = os.fsencode(path)
        self._stream = engine._stream()
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()

    def _run(self):
        try:
            self.engine.torch.cuda.set_device(self.engine.device)
            self._rc = self.engine.lib.bpe_load_file_ranges(self.engine.ctx, self._path, len(self.lens), _np_ptr(self._file_off), _np_ptr(self._len),
                                                            _np_ptr(self._dev_off), self.buffer.data_ptr(), C.byref(self._done), self._stream)
        except BaseException as e:  # noqa: BLE001 - re-raised by piece() / join()
            self._error = e

    def _check(self):
        if self._error is not None:
            raise self._error
        if self._rc:
            self.engine._check(self._rc)

    def piece(self, k: int):
        """(device tensor view, length) of range k, once it is completely in device memory."""
        while self._done.value <= k:
            if not self._thread.is_alive():
                self._check()
                if self._done.value <= k:
                    raise RuntimeError("bpe_load_file_ranges ended before the range arrived")
                break
            time.sleep(0.0001)
        a, n = self.dev_off[k], self.lens[k]
        return self.buffer[a:a + n + TEXT_PAD], n

    def join(self):
        self._thread.join()
        self._check()


def pinned_array(n_items: int, dtype) -> np.ndarray:
    """Page-locked NumPy array (a view of a pinned torch tensor, which it keeps alive): as the `out=` of
    BPETokenizer.encode_file_to_array it receives the ids by DMA, with no host copy.  Allocate once, reuse across calls —
    page-locking costs far more than one copy."""
    import torch

    nbytes = int(n_items) * np.dtype(dtype).itemsize
    return torch.empty(nbytes, dtype=torch.uint8, pin_memory=True).numpy().view(dtype)


def host_array(n_items: int, dtype) -> np.ndarray:
    """Uninitialised NumPy array for results copied back from the GPU.  (Transparent huge pages were tried for the large
    ones and made the first touch by 16 copy threads five times slower on the test box: plain pages it is.)"""
    return np.empty(n_items, dtype=dtype)


_engines: dict[int, Engine] = {}


def get_engine(device: int | None = None) -> Engine:
    """The process-wide engine of a GPU (created on first use)."""
    import torch

    if not torch.cuda.is_available():
        raise RuntimeError("bpe_transformer_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    dev = torch.cuda.current_device() if device is None else int(device)
    eng = _engines.get(dev)
    if eng is None:
        eng = Engine(dev, table_log2=int(os.environ.get("BPE_B200_TABLE_LOG2", "22")),
                     pool_bytes=int(os.environ.get("BPE_B200_POOL_BYTES", str(256 << 20))))
        _engines[dev] = eng
    return eng
