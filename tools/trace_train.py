"""Developer tool: per-iteration timing of the merge loop (BPE_B200_TRAIN_TRACE)."""
import os, sys, pathlib
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import numpy as np
import bench
from bpe_transformer_b200 import _lib
from bpe_transformer_b200.tokenization import BPETrainer

nbytes = int(sys.argv[1]) if len(sys.argv) > 1 else 2_100_000_000
out = sys.argv[2] if len(sys.argv) > 2 else "gpurun_out/trace.txt"
path = bench.corpus_path(nbytes)
bench.make_corpus(path, nbytes)
eng = _lib.get_engine(0)
eng.set_special_tokens(_lib.sort_specials([bench.EOT]))
d_text, n = eng.load_file(path)
for rep in range(2):
    if rep == 1:
        os.environ["BPE_B200_TRAIN_TRACE"] = out
    eng.unique_reset()
    eng.pretokenize_count(d_text, n)
    tr = BPETrainer(vocab_size=10000, special_tokens=[bench.EOT])
    tr._merge_tokens(eng)
t = np.loadtxt(out, dtype=np.int64)
ns, ll, allw, matched = t[:, 0], t[:, 1], t[:, 2], t[:, 3]
print(f"iterations {len(t)} total {ns.sum() / 1e6:.1f} ms; all-words steps {int(allw.sum())} take {ns[allw == 1].sum() / 1e6:.1f} ms "
      f"(median {np.median(ns[allw == 1]) / 1e3:.1f} us); list steps {int((allw == 0).sum())} take {ns[allw == 0].sum() / 1e6:.1f} ms "
      f"(median {np.median(ns[allw == 0]) / 1e3:.1f} us)")
for lo, hi in [(0, 32), (32, 512), (512, 4096), (4096, 32768), (32768, 10**9)]:
    m = (allw == 0) & (ll >= lo) & (ll < hi)
    if m.any():
        print(f"  list length [{lo},{hi}): {int(m.sum())} steps, {ns[m].sum() / 1e6:.1f} ms, median {np.median(ns[m]) / 1e3:.1f} us, mean matched {matched[m].mean():.0f}")
for a, b in [(0, 500), (500, 2000), (2000, 5000), (5000, len(t))]:
    print(f"  iterations [{a},{b}): {ns[a:b].sum() / 1e6:.1f} ms, median {np.median(ns[a:b]) / 1e3:.1f} us")
print("slowest:", sorted(zip(ns.tolist(), ll.tolist(), allw.tolist(), matched.tolist()), reverse=True)[:8])
