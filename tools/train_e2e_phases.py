"""Developer tool: where the host time of train_bpe(path, ...) goes (count phase, merge loop call, Python bookkeeping)."""
import sys, pathlib, time
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch
import bench
from bpe_transformer_b200 import _lib, train_bpe
from bpe_transformer_b200.tokenization import BPETrainer
from bpe_transformer_b200.tokenization.preprocessing.pretokenization import count_into_engine

nbytes = int(sys.argv[1]) if len(sys.argv) > 1 else 2_100_000_000
path = bench.corpus_path(nbytes)
bench.make_corpus(path, nbytes)
eng = _lib.get_engine(0)
for rep in range(4):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    tr = BPETrainer(vocab_size=10000, special_tokens=[bench.EOT])
    t1 = time.perf_counter()
    count_into_engine(eng, path, [bench.EOT], 4)
    t2 = time.perf_counter()
    m_a, m_b = eng.train(10000, [tr._vocab[i] for i in range(len(tr._vocab))])
    t3 = time.perf_counter()
    tr2 = BPETrainer(vocab_size=10000, special_tokens=[bench.EOT])
    tr2._merges, tr2._vocab = [], dict(tr._vocab)
    t4 = time.perf_counter()
    print(f"rep {rep}: ctor {1e3 * (t1 - t0):.1f} ms, load+scan+count {1e3 * (t2 - t1):.1f} ms, engine.train {1e3 * (t3 - t2):.1f} ms "
          f"(device: {eng.train_timing()})", flush=True)
    t0 = time.perf_counter()
    v, m = train_bpe(path, 10000, [bench.EOT])
    print(f"        train_bpe total {1e3 * (time.perf_counter() - t0):.1f} ms", flush=True)
