"""Developer tool: one resident train step (scan, count, merge loop) on the bench corpus — the command profiled by ncu."""
import sys, pathlib
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench
from bpe_transformer_b200 import _lib
from bpe_transformer_b200.tokenization import BPETrainer

nbytes = int(sys.argv[1]) if len(sys.argv) > 1 else 2_100_000_000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
path = bench.corpus_path(nbytes)
bench.make_corpus(path, nbytes)
eng = _lib.get_engine(0)
eng.set_special_tokens(_lib.sort_specials([bench.EOT]))
d_text, n = eng.load_file(path)
for _ in range(reps):
    eng.unique_reset()
    eng.pretokenize_count(d_text, n)
    tr = BPETrainer(vocab_size=10000, special_tokens=[bench.EOT])
    tr._merge_tokens(eng)
    print(tr.last_timing, flush=True)
