"""Developer tool: where do the slow runs of the merge loop lose their time?  Trains `reps` times with the per-step trace on
and prints, per run: loop time, median step, time in steps slower than 100 us, and for slow runs where those steps are.

    python tools/outlier_hunt.py [reps=40]
"""
import os, sys, pathlib
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import numpy as np
import bench
from bpe_transformer_b200 import _lib
from bpe_transformer_b200.tokenization import BPETrainer

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 40
nbytes = 2_100_000_000
path = bench.corpus_path(nbytes)
bench.make_corpus(path, nbytes)
eng = _lib.get_engine(0)
eng.set_special_tokens(_lib.sort_specials([bench.EOT]))
d_text, n = eng.load_file(path)
out = "/tmp/hunt_trace.txt"
os.environ["BPE_B200_TRAIN_TRACE"] = out
rows = []
for rep in range(reps):
    eng.unique_reset()
    eng.pretokenize_count(d_text, n)
    tr = BPETrainer(vocab_size=10000, special_tokens=[bench.EOT])
    tr._merge_tokens(eng)
    t = np.loadtxt(out, dtype=np.int64)
    ns = t[:, 0]
    slow = ns > 100_000
    loop = tr.last_timing["merge_loop_ms"]
    rows.append(loop)
    line = (f"run {rep:2d}: loop {loop:7.2f} ms, steps sum {ns.sum() / 1e6:7.2f} ms, median {np.median(ns) / 1e3:.2f} us, "
            f"p90 {np.percentile(ns, 90) / 1e3:.1f} us, steps > 100 us: {int(slow.sum())} = {ns[slow].sum() / 1e6:.2f} ms")
    if loop > 1.05 * min(rows):
        idx = np.flatnonzero(slow)
        line += "  at " + ", ".join(f"{i}:{ns[i] // 1000}us(list {t[i, 1]}, all {t[i, 2]})" for i in idx[:12])
        q = [ns[a:b].sum() / 1e6 for a, b in ((0, 500), (500, 2000), (2000, 5000), (5000, len(ns)))]
        line += "  quarters " + "/".join(f"{x:.1f}" for x in q)
    print(line, flush=True)
print(f"min {min(rows):.2f} mean {sum(rows) / len(rows):.2f} max {max(rows):.2f}")
