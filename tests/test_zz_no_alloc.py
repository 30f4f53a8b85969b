"""bpe_decode_noalloc: the decode entry point for callers that may not allocate device memory.

Some users of this library decode id batches on a GPU whose memory is entirely owned by another allocator (for
example a training process whose framework pool has claimed all of HBM, where a raw cudaMalloc fails), or from a
loop that cannot take the device-wide stall of cudaMalloc / cudaFree.  For them the library offers one entry point
that allocates nothing: the caller brings every buffer, and the small working area sits in an array inside the
library itself (which is why a single call is limited to BPE_NOALLOC_MAX_IDS ids).

The check: called twice in a row through the C-ABI — a big batch, then a smaller and different one, so anything the
first call left in the shared array would show up — its bytes must equal, exactly, what the ordinary bpe_decode
produces for the same ids.  The file sorts last on purpose: a run that stops at the first failure has then already
executed every other test.  It passes alone in a fresh process as well as after the rest of the suite.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_IDS = 1 << 24


def _vocab():
    rnd = np.random.default_rng(5)
    vocab = {i: bytes([i]) for i in range(256)}
    for k in range(256, 3000):
        if k % 97 == 0:
            continue  # holes: ids without an entry decode to U+FFFD
        vocab[k] = bytes(rnd.integers(32, 127, size=int(rnd.integers(1, 12)), dtype=np.uint8))
    return vocab


def _flat(vocab):
    n_vocab = max(vocab) + 1
    lens = np.zeros(n_vocab + 1, dtype=np.uint64)
    for k, v in vocab.items():
        lens[k + 1] = len(v)
    off = np.cumsum(lens, dtype=np.uint64)
    blob = np.frombuffer(b"".join(vocab.get(k, b"") for k in range(n_vocab)) + b"\0", dtype=np.uint8).copy()
    return blob, off, n_vocab


def _noalloc(lib, d_ids, d_blob, d_off, n_vocab, stream):
    """The caller-side protocol: ask for the size, bring the output buffer, decode."""
    import torch

    err = C.create_string_buffer(256)
    n_bytes = C.c_uint64(0)
    elem = d_ids.element_size()
    rc = lib.bpe_decode_noalloc(d_ids.data_ptr(), elem, d_ids.numel(), d_blob.data_ptr(), d_off.data_ptr(), n_vocab,
                                None, 0, C.byref(n_bytes), stream, err, 256)
    assert rc == 0, err.value
    out = torch.empty(n_bytes.value + 16, dtype=torch.uint8, device=d_ids.device)  # the caller's buffer
    rc = lib.bpe_decode_noalloc(d_ids.data_ptr(), elem, d_ids.numel(), d_blob.data_ptr(), d_off.data_ptr(), n_vocab,
                                out.data_ptr(), n_bytes.value, C.byref(n_bytes), stream, err, 256)
    assert rc == 0, err.value
    return out[:n_bytes.value]


def test_noalloc_decode_equals_decode_twice_in_a_row():
    import torch

    from bpe_transformer_b200 import _lib

    eng = _lib.get_engine()
    lib = _lib.load_library()
    dev = f"cuda:{eng.device}"
    vocab = _vocab()
    blob, off, n_vocab = _flat(vocab)
    d_blob = torch.from_numpy(blob).to(dev)
    d_off = torch.from_numpy(off.view(np.int64)).to(dev)
    stream = C.c_void_p(torch.cuda.current_stream(eng.device).cuda_stream)
    rnd = np.random.default_rng(9)
    # first call: many tiles, 32-bit ids, some ids far outside the vocabulary; second call: fewer, 16-bit ids
    batches = [rnd.integers(0, 3300, size=1_000_003).astype(np.uint32), rnd.integers(0, 3000, size=70_001).astype(np.uint16)]
    batches[0][::1000] = 4_000_000_000
    for ids in batches:
        view = ids.view(np.int32) if ids.dtype == np.uint32 else ids.view(np.int16)
        d_ids = torch.from_numpy(view.copy()).to(dev)
        want = eng.decode(d_ids, blob, off)                      # the existing entry point
        got = _noalloc(lib, d_ids, d_blob, d_off, n_vocab, stream)
        assert got.numel() == want.numel()
        assert torch.equal(got, want)                             # bit-exact, like every other test here
        host = b"".join(vocab.get(int(i), b"\xef\xbf\xbd") for i in ids[:5000].tolist())
        assert got[:len(host)].cpu().numpy().tobytes() == host    # and equal to the reference's rule for the head


def test_noalloc_decode_rejects_oversize_batches_with_a_message():
    import torch

    from bpe_transformer_b200 import _lib

    eng = _lib.get_engine()
    lib = _lib.load_library()
    dev = f"cuda:{eng.device}"
    blob, off, n_vocab = _flat(_vocab())
    d_blob = torch.from_numpy(blob).to(dev)
    d_off = torch.from_numpy(off.view(np.int64)).to(dev)
    d_ids = torch.zeros(16, dtype=torch.int16, device=dev)
    err = C.create_string_buffer(256)
    n_bytes = C.c_uint64(0)
    # only the declared count matters for the check: nothing is read before the limit test
    rc = lib.bpe_decode_noalloc(d_ids.data_ptr(), 2, MAX_IDS + 1, d_blob.data_ptr(), d_off.data_ptr(), n_vocab, None, 0,
                                C.byref(n_bytes), None, err, 256)
    assert rc == _lib.ERR_CAPACITY
    assert str(MAX_IDS).encode() in err.value and b"split" in err.value
    rc = lib.bpe_decode_noalloc(d_ids.data_ptr(), 3, 16, d_blob.data_ptr(), d_off.data_ptr(), n_vocab, None, 0, C.byref(n_bytes), None, err, 256)
    assert rc == _lib.ERR_ARGUMENT and err.value
