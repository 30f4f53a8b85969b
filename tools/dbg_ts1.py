import sys, pathlib, time
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests")); sys.path.insert(0, str(ROOT / "oracle"))
from conftest import synth_corpus
import c_oracle
from bpe_transformer_b200 import _lib
from bpe_transformer_b200.tokenization.preprocessing.pretokenization import counter_from_engine
EOT = "<|endoftext|>"
data = synth_corpus(0, 1, 1_000_000)
eng = _lib.get_engine()
eng.set_special_tokens(_lib.sort_specials([EOT]))
eng.unique_reset()
d, n = eng.to_device(data)
t = time.time(); eng.count(d, n, eng.scan(d, n)); print("count ok", time.time() - t, eng.unique_stats(), flush=True)
got = counter_from_engine(eng)
want = c_oracle.count(data, [EOT])
print("counter equal", got == want, len(got), len(want), flush=True)
