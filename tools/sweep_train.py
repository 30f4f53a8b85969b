"""Developer tool: time the resident training step for several merge-loop grid sizes (BPE_B200_TRAIN_CTAS)."""
import os, sys, time, pathlib
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch
import bench
from bpe_transformer_b200 import _lib
from bpe_transformer_b200.tokenization import BPETrainer

nbytes = int(sys.argv[1]) if len(sys.argv) > 1 else 2_100_000_000
path = bench.corpus_path(nbytes)
bench.make_corpus(path, nbytes)
eng = _lib.get_engine(0)
eng.set_special_tokens(_lib.sort_specials([bench.EOT]))
d_text, n = eng.load_file(path)
ref = None
for ctas in sys.argv[2:] or ["0"]:
    if ctas == "0":
        os.environ.pop("BPE_B200_TRAIN_CTAS", None)
    else:
        os.environ["BPE_B200_TRAIN_CTAS"] = ctas
    for rep in range(2):
        eng.unique_reset()
        eng.set_profiling(True)
        eng.pretokenize_count(d_text, n)
        tr = BPETrainer(vocab_size=10000, special_tokens=[bench.EOT])
        tr._merge_tokens(eng)
        kt = eng.kernel_times()
    if ref is None:
        ref = tr.merges
    print(f"ctas={ctas}: merge loop {tr.last_timing['merge_loop_ms']:.1f} ms, {tr.last_timing['iterations']} iterations, "
          f"{tr.last_timing['merge_loop_ms'] * 1e3 / max(1, len(tr.merges)):.2f} us/merge, same={tr.merges == ref}, "
          f"count {kt['fused_count_kernel'][0]:.2f} ms", flush=True)
for mode in ("cluster", "grid"):
    os.environ["BPE_B200_TRAIN_MODE"] = mode
    os.environ.pop("BPE_B200_TRAIN_CTAS", None)
    for rep in range(2):
        eng.unique_reset()
        eng.pretokenize_count(d_text, n)
        tr = BPETrainer(vocab_size=10000, special_tokens=[bench.EOT])
        tr._merge_tokens(eng)
    print(f"mode={mode}: merge loop {tr.last_timing['merge_loop_ms']:.1f} ms, {tr.last_timing['iterations']} iterations, "
          f"{tr.last_timing['merge_loop_ms'] * 1e3 / max(1, len(tr.merges)):.2f} us/merge, same={tr.merges == ref}", flush=True)
