"""Developer tool: the one-pass pretokenize+count kernel (and optionally a warm encode) on the bench corpus — the command profiled by ncu.

    python tools/run_count_once.py [bytes] [reps] [encode]
"""
import sys, pathlib, time
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch
import bench
from bpe_transformer_b200 import _lib

nbytes = int(sys.argv[1]) if len(sys.argv) > 1 else 2_100_000_000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
path = bench.corpus_path(nbytes)
bench.make_corpus(path, nbytes)
eng = _lib.get_engine(0)
eng.set_special_tokens(_lib.sort_specials([bench.EOT]))
d_text, n = eng.load_file(path)
for _ in range(reps):
    eng.unique_reset()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    eng.pretokenize_count(d_text, n)
    torch.cuda.synchronize()
    print(f"pretokenize_count {n} B: {(time.perf_counter() - t0) * 1e3:.2f} ms, stats {eng.unique_stats()}", flush=True)
