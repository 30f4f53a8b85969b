"""Developer tool: encode the bench corpus a few times from HBM and print wall time and per-kernel device time per call."""
import sys, pathlib, time
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch
import bench
from bpe_transformer_b200 import _lib
from bpe_transformer_b200.tokenization import BPETrainer, BPETokenizer

nbytes = int(sys.argv[1]) if len(sys.argv) > 1 else 2_100_000_000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
prof = (sys.argv[3] != "0") if len(sys.argv) > 3 else True
path = bench.corpus_path(nbytes)
bench.make_corpus(path, nbytes)
eng = _lib.get_engine(0)
eng.set_special_tokens(_lib.sort_specials([bench.EOT]))
d_text, n = eng.load_file(path)
eng.unique_reset()
eng.pretokenize_count(d_text, n)
tr = BPETrainer(vocab_size=10000, special_tokens=[bench.EOT])
tr._merge_tokens(eng)
tok = BPETokenizer(tr.vocab, tr.merges, [bench.EOT])
for r in range(reps):
    eng.set_profiling(prof)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    ids = tok.encode_bytes_to_device(None, flags=_lib.CUT_AFTER_NEWLINE, engine=eng, on_device=(d_text, n))
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"call {r}: {dt * 1e3:.2f} ms wall, {ids.numel()} ids, kernels {eng.kernel_times()}", flush=True)
