"""Developer tool: BASELINE configs[4] shape on ONE GPU — OpenWebText-shaped synthetic corpus (tools/gen_corpus.c style 1),
train_bpe vocab_size=32000, then encode the whole file and check that GPU decode gives the bytes back.

    python tools/run_c5.py [bytes=11000000000] [vocab=32000]
No oracle finishes at this size: what is checked is size independent (merge count, id range, decode(encode(x)) == x,
two trainings give the same merges).
"""
import ctypes, faulthandler, os, pathlib, sys, time
faulthandler.dump_traceback_later(45, repeat=True)  # where the host is, should something take minutes
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import numpy as np
import torch
import bench
from bpe_transformer_b200 import _lib, train_bpe
from bpe_transformer_b200.tokenization import BPETokenizer

nbytes = int(sys.argv[1]) if len(sys.argv) > 1 else 11_000_000_000
vocab_size = int(sys.argv[2]) if len(sys.argv) > 2 else 32000
os.environ.setdefault("BPE_B200_TRAIN_VERBOSE", "1")
path = f"/dev/shm/bpe_b200_owt_seed5_{nbytes}.txt"
if not os.path.exists(path):
    so = ROOT / "tools" / "libgencorpus.so"
    lib = ctypes.CDLL(str(so))
    lib.gen_corpus.restype = ctypes.c_uint64
    t0 = time.perf_counter()
    n = lib.gen_corpus(ctypes.c_int(1), ctypes.c_uint64(5), ctypes.c_uint64(nbytes), path.encode(), ctypes.c_int(min(32, os.cpu_count() or 1)))
    print(f"corpus {path}: {n} B in {time.perf_counter() - t0:.1f} s", flush=True)
size = os.path.getsize(path)
eng = _lib.get_engine(0)
results = []
for rep in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    vocab, merges = train_bpe(path, vocab_size, [bench.EOT])
    dt = time.perf_counter() - t0
    n_unique, n_bytes, n_tokens = eng.unique_stats()
    print(f"train {rep}: {dt:.3f} s, {len(merges)} merges, {n_unique} distinct pretokens ({n_bytes} B), {n_tokens} pretokens, "
          f"device timing {eng.train_timing()}", flush=True)
    results.append(merges)
assert results[0] == results[1], "two trainings disagree"
assert len(vocab) == vocab_size and len(merges) == vocab_size - 257
tok = BPETokenizer(vocab, merges, [bench.EOT])
for rep in range(2):
    t0 = time.perf_counter()
    ids = tok.encode_file_to_array(path, engine=eng)
    dt = time.perf_counter() - t0
    print(f"encode {rep}: {dt:.3f} s = {size / dt / 1e9:.2f} GB/s, {len(ids)} ids ({size / len(ids):.2f} B per id), dtype {ids.dtype}, max id {int(ids.max())}", flush=True)
# roundtrip in pieces of 512M ids (GPU decode)
t0 = time.perf_counter()
pos = 0
with open(path, "rb") as fh:
    for a in range(0, len(ids), 1 << 29):
        back = tok.decode_array(ids[a:a + (1 << 29)], engine=eng)
        want = fh.read(len(back))
        assert back == want, f"roundtrip differs in the piece at id {a}"
        pos += len(back)
assert pos == size
print(f"roundtrip ok ({time.perf_counter() - t0:.1f} s)", flush=True)
