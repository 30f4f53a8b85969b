"""Developer tool: merge-loop time as a function of the SM the driver CTA runs on (BPE_B200_TRAIN_DRIVER_SM)."""
import sys, pathlib, os
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench
from bpe_transformer_b200 import _lib
from bpe_transformer_b200.tokenization import BPETrainer

nbytes = 2_100_000_000
path = bench.corpus_path(nbytes)
bench.make_corpus(path, nbytes)
eng = _lib.get_engine(0)
eng.set_special_tokens(_lib.sort_specials([bench.EOT]))
d_text, n = eng.load_file(path)
sms = [int(a) for a in sys.argv[1:]] or [-1, 0, 1, 2, 17, 36, 37, 72, 73, 74, 75, 110, 111, 144, 145, 146, 147]
ref = None
for sm in sms:
    os.environ["BPE_B200_TRAIN_DRIVER_SM"] = str(sm)
    times = []
    for rep in range(3):
        eng.unique_reset()
        eng.pretokenize_count(d_text, n)
        tr = BPETrainer(vocab_size=10000, special_tokens=[bench.EOT])
        tr._merge_tokens(eng)
        times.append(tr.last_timing["merge_loop_ms"])
    ref = ref or tr.merges
    print(f"driver on SM {sm:4d}: merge loop {min(times):.2f} ms (of {', '.join(f'{t:.1f}' for t in times)}), same merges {tr.merges == ref}", flush=True)
