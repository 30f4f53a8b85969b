"""Developer tool: merge-loop time over repeated trainings, with the SM the driver CTA ran on (run-to-run variance: does placement explain it?).

    python tools/train_variance.py [reps=24]
Prints one line per training (merge loop ms, iterations, heavy steps, driver SM) and a summary per SM.
"""
import sys, pathlib, os, re, subprocess
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
if len(sys.argv) > 1 and sys.argv[1] == "child":
    os.environ["BPE_B200_TRAIN_VERBOSE"] = "1"
    import bench
    from bpe_transformer_b200 import _lib
    from bpe_transformer_b200.tokenization import BPETrainer
    nbytes = 2_100_000_000
    path = bench.corpus_path(nbytes)
    bench.make_corpus(path, nbytes)
    eng = _lib.get_engine(0)
    eng.set_special_tokens(_lib.sort_specials([bench.EOT]))
    d_text, n = eng.load_file(path)
    for rep in range(int(sys.argv[2])):
        eng.unique_reset()
        eng.pretokenize_count(d_text, n)
        BPETrainer(vocab_size=10000, special_tokens=[bench.EOT])._merge_tokens(eng)
    sys.exit(0)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 24
out = subprocess.run([sys.executable, __file__, "child", str(reps)], capture_output=True, text=True).stderr  # the library prints to stderr
rows = [(float(m.group(1)), int(m.group(2))) for m in re.finditer(r"merge loop ([\d.]+) ms.*driver CTA on SM (\d+)", out)]
print(out)
by_sm = {}
for ms, sm in rows:
    print(f"merge loop {ms:.2f} ms, driver CTA on SM {sm}")
    by_sm.setdefault(sm, []).append(ms)
for sm, v in sorted(by_sm.items()):
    print(f"SM {sm}: {len(v)} runs, min {min(v):.2f} max {max(v):.2f} mean {sum(v) / len(v):.2f} ms")
if rows:
    v = [r[0] for r in rows]
    print(f"all: {len(v)} runs, min {min(v):.2f} max {max(v):.2f} mean {sum(v) / len(v):.2f} ms")
