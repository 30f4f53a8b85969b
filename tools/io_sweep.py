"""Developer tool: time bpe_load_file / bpe_store_to_host on the bench corpus (thread count from BPE_B200_IO_THREADS)."""
import sys, pathlib, time, os
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import numpy as np, torch
import bench
from bpe_transformer_b200 import _lib
nbytes = int(sys.argv[1]) if len(sys.argv) > 1 else 2_100_000_000
path = bench.corpus_path(nbytes)
bench.make_corpus(path, nbytes)
eng = _lib.get_engine(0)
print("threads", os.environ.get("BPE_B200_IO_THREADS"), "cpus", os.cpu_count())
for r in range(4):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    d, n = eng.load_file(path)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(f"load {n / dt / 1e9:.1f} GB/s ({dt * 1e3:.1f} ms)", flush=True)
for r in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    h = eng.to_host_array(d[:n], np.uint8) if hasattr(eng, "to_host_array") else None
    dt = time.perf_counter() - t0
    print(f"store {n / dt / 1e9:.1f} GB/s ({dt * 1e3:.1f} ms)", flush=True)
