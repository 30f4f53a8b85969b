"""Developer tool: merge-loop time over repeated trainings with different activity in between (variance hunt)."""
import sys, pathlib, time, os
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
os.environ["BPE_B200_TRAIN_VERBOSE"] = "1"
import torch
import bench
from bpe_transformer_b200 import _lib
from bpe_transformer_b200.tokenization import BPETrainer
from bpe_transformer_b200.tokenization.preprocessing.pretokenization import count_into_engine

nbytes = 2_100_000_000
path = bench.corpus_path(nbytes)
bench.make_corpus(path, nbytes)
eng = _lib.get_engine(0)
mode = sys.argv[1] if len(sys.argv) > 1 else "file"
eng.set_special_tokens(_lib.sort_specials([bench.EOT]))
d_text, n = eng.load_file(path)
for rep in range(8):
    if mode == "file":
        count_into_engine(eng, path, [bench.EOT], 4)
    elif mode == "resident":
        eng.unique_reset(); eng.count(d_text, n, eng.scan(d_text, n))
    elif mode == "sleep":
        eng.unique_reset(); eng.count(d_text, n, eng.scan(d_text, n)); torch.cuda.synchronize(); time.sleep(0.05)
    tr = BPETrainer(vocab_size=10000, special_tokens=[bench.EOT])
    tr._merge_tokens(eng)
