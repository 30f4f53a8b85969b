import sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo/oracle")
from conftest import synth_corpus
from bpe_transformer_b200 import _lib
from bpe_transformer_b200.tokenization.preprocessing.pretokenization import counter_from_engine
eng = _lib.get_engine()
eng.set_special_tokens(_lib.sort_specials(["<|endoftext|>"]))
data = synth_corpus(1, 12, 8_000_000)
d_text, n = eng.to_device(data)
eng.unique_reset(); eng.count(d_text, n, eng.scan(d_text, n)); ref = counter_from_engine(eng)
for k in range(6):
    eng.unique_reset(); eng.pretokenize_count(d_text, n); c = counter_from_engine(eng)
    bad = [(bytes(t), v, ref.get(t)) for t, v in c.items() if ref.get(t) != v]
    miss = [bytes(t) for t in ref if t not in c]
    print(k, "equal" if c == ref else f"DIFF {len(bad)} wrong, {len(miss)} missing; e.g. {bad[:4]} {miss[:3]}", eng.unique_stats())
