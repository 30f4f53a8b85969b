"""Developer tool: timeline of the load / encode / store pipeline of BPETokenizer.encode_file_to_array."""
import sys, pathlib, time, threading
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch
import bench
from bpe_transformer_b200 import _lib
from bpe_transformer_b200.tokenization import BPETrainer, BPETokenizer

nbytes = int(sys.argv[1]) if len(sys.argv) > 1 else 2_100_000_000
path = bench.corpus_path(nbytes)
bench.make_corpus(path, nbytes)
eng = _lib.get_engine(0)
eng.set_special_tokens(_lib.sort_specials([bench.EOT]))
d_text, n = eng.load_file(path)
eng.unique_reset()
eng.pretokenize_count(d_text, n)
tr = BPETrainer(vocab_size=10000, special_tokens=[bench.EOT])
tr._merge_tokens(eng)
del d_text
tok = BPETokenizer(tr.vocab, tr.merges, [bench.EOT])
events = []
def wrap(name):
    f = getattr(eng, name)
    def g(*a, **k):
        t0 = time.perf_counter(); r = f(*a, **k); t1 = time.perf_counter()
        events.append((name, t0, t1)); return r
    setattr(eng, name, g)
for name in ("load_file", "encode", "store_into"):
    wrap(name)
_ranges = eng.load_file_ranges
class TimedLoad:
    def __init__(self, inner): self.inner = inner
    def piece(self, k):
        t0 = time.perf_counter(); r = self.inner.piece(k); events.append(("wait_piece", t0, time.perf_counter())); return r
    def join(self): return self.inner.join()
eng.load_file_ranges = lambda *a, **k: TimedLoad(_ranges(*a, **k))
import numpy as np
pinned = _lib.pinned_array(int(nbytes * 0.28) + 4096, np.uint16)
for rep in range(4):
    events.clear()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    ids = tok.encode_file_to_array(path, engine=eng, out=pinned)
    dt = time.perf_counter() - t0
    print(f"rep {rep}: {dt * 1e3:.1f} ms, {len(ids)} ids", flush=True)
busy = {}
for name, a, b in events:
    busy[name] = busy.get(name, 0) + (b - a)
print("busy ms:", {k: round(v * 1e3, 1) for k, v in busy.items()})
for name, a, b in sorted(events, key=lambda e: e[1])[:60]:
    print(f"  {name:10s} {1e3 * (a - t0):7.2f} -> {1e3 * (b - t0):7.2f}  ({1e3 * (b - a):.2f} ms)")
