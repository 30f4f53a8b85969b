"""Developer tool: where the host time of train_bpe(path, ...) goes (count phase, merge loop call, Python bookkeeping),
optionally with bench.py's nvidia-smi clock sampler running beside it (argument "sampler")."""
import sys, pathlib, time
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch
import bench
from bpe_transformer_b200 import _lib, train_bpe
from bpe_transformer_b200.tokenization import BPETrainer
from bpe_transformer_b200.tokenization.preprocessing import pretokenization as pt

nbytes = 2_100_000_000
path = bench.corpus_path(nbytes)
bench.make_corpus(path, nbytes)
eng = _lib.get_engine(0)
sampler = None
if "sampler" in sys.argv:
    sampler = bench.ClockSampler(0); sampler.start()
marks = {}
orig_count, orig_train = pt.count_into_engine, eng.train
def count_wrapped(*a, **k):
    t = time.perf_counter(); r = orig_count(*a, **k); marks["count"] = time.perf_counter() - t; return r
def train_wrapped(*a, **k):
    t = time.perf_counter(); r = orig_train(*a, **k); marks["train"] = time.perf_counter() - t; return r
import bpe_transformer_b200.tokenization.bpe_trainer as bt
bt.count_into_engine = count_wrapped
eng.train = train_wrapped
for rep in range(6):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    v, m = train_bpe(path, 10000, [bench.EOT])
    dt = time.perf_counter() - t0
    print(f"rep {rep}: train_bpe {1e3 * dt:.1f} ms = count phase {1e3 * marks['count']:.1f} + engine.train {1e3 * marks['train']:.1f} "
          f"(merge loop {eng.train_timing()[1]:.1f}) + python {1e3 * (dt - marks['count'] - marks['train']):.1f}", flush=True)
if sampler:
    print(sampler.stop())
