"""Developer tool: merge loop time against the size of the hot-pair cache (BPE_B200_TRAIN_CACHE_FILL), on the bench corpus
(style 0, 2.1 GB, vocab 10000) or on the configs[4] corpus (style 1, 11 GB, vocab 32000: `c5`).

    python tools/cache_fill_sweep.py [c2|c5] [fill ...]      (fill 0 = the library's own choice)
"""
import hashlib, os, sys, pathlib, time
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
os.environ["BPE_B200_TRAIN_VERBOSE"] = "1"
import bench
from bpe_transformer_b200 import _lib, train_bpe

which = sys.argv[1] if len(sys.argv) > 1 else "c2"
fills = [int(v) for v in sys.argv[2:]] or [512, 1024, 2048, 3072]
if which == "c5":
    nbytes, vocab, style, seed = 11_000_000_000, 32000, 1, 5
    path = f"/dev/shm/bpe_b200_owt_seed{seed}_{nbytes}.txt"
    if not os.path.exists(path):
        import ctypes
        lib = ctypes.CDLL(str(ROOT / "tools" / "libgencorpus.so"))
        lib.gen_corpus.restype = ctypes.c_uint64
        lib.gen_corpus(ctypes.c_int(style), ctypes.c_uint64(seed), ctypes.c_uint64(nbytes), path.encode(), ctypes.c_int(min(32, os.cpu_count() or 1)))
else:
    nbytes, vocab = 2_100_000_000, 10000
    path = bench.corpus_path(nbytes)
    bench.make_corpus(path, nbytes)
ref = None
for fill in fills:
    if fill:
        os.environ["BPE_B200_TRAIN_CACHE_FILL"] = str(fill)
    else:
        os.environ.pop("BPE_B200_TRAIN_CACHE_FILL", None)  # 0 = the library's own choice
    for rep in range(2):
        t0 = time.perf_counter()
        vocab_d, merges = train_bpe(path, vocab, [bench.EOT])
        dt = time.perf_counter() - t0
    sha = hashlib.sha256(repr(merges).encode()).hexdigest()[:12]
    ref = ref or sha
    t = _lib.get_engine().train_timing()
    print(f"cache fill {fill}: train_bpe {dt:.3f} s, merge loop {t[1]:.1f} ms, {len(merges)} merges, same merges {sha == ref}", flush=True)
