"""Developer tool: scan + count time on pieces of a style-1 (OpenWebText-shaped) corpus, per table size."""
import ctypes, faulthandler, os, pathlib, sys, time
faulthandler.dump_traceback_later(60, repeat=True)
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch
import bench
from bpe_transformer_b200 import _lib

nbytes = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000_000
path = f"/dev/shm/bpe_b200_owt_seed5_{nbytes}.txt"
if not os.path.exists(path):
    lib = ctypes.CDLL(str(ROOT / "tools" / "libgencorpus.so"))
    lib.gen_corpus.restype = ctypes.c_uint64
    lib.gen_corpus(ctypes.c_int(1), ctypes.c_uint64(5), ctypes.c_uint64(nbytes), path.encode(), ctypes.c_int(min(32, os.cpu_count() or 1)))
from bpe_transformer_b200.tokenization.preprocessing.pretokenization import find_chunk_boundaries
with open(path, "rb") as fh:
    offs = find_chunk_boundaries(fh, 8, [bench.EOT])
eng = _lib.get_engine(0)
eng.set_special_tokens(_lib.sort_specials([bench.EOT]))
for log2 in (24, 22):
    eng.configure(log2, 1 << 30)
    eng.unique_reset()
    for piece in range(len(offs) - 1):
        d, n = eng.load_file(path, offs[piece], offs[piece + 1])
        torch.cuda.synchronize(); t0 = time.perf_counter()
        mask = eng.scan(d, n)
        torch.cuda.synchronize(); t1 = time.perf_counter()
        try:
            eng.count(d, n, mask)
            err = ""
        except _lib.BpeError as e:
            err = str(e)
        torch.cuda.synchronize(); t2 = time.perf_counter()
        print(f"table 2^{log2} piece {piece}: scan {1e3 * (t1 - t0):.1f} ms, count {1e3 * (t2 - t1):.1f} ms, stats {eng.unique_stats()} {err}", flush=True)
