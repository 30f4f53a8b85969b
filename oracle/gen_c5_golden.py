#!/usr/bin/env python3
"""Golden hashes for the OpenWebText-shaped configuration (BASELINE.json configs[4] at a size a CPU finishes).  TEST INFRASTRUCTURE ONLY.

Generates the seeded OWT-shaped corpus (tools/gen_corpus.c style 1), trains vocab_size 32000 with the port of the reference's
algorithm (oracle/bpe_ref_port.py, pinned to the reference by tests/test_oracle_pinned.py) on all host cores, encodes the corpus
line by line with the C oracle, and writes tests/golden/c5_owt_sha256.json.gz: sha256 of the ordered merges and of the uint16 ids.

    python oracle/gen_c5_golden.py [bytes] [seed]          (default 128 MB, seed 0; takes some minutes and a few GB of RAM)
"""
from __future__ import annotations

import ctypes
import gzip
import hashlib
import json
import os
import pathlib
import subprocess
import sys
import time

HERE = pathlib.Path(__file__).resolve().parent
ROOT = HERE.parent
sys.path.insert(0, str(HERE))
sys.path.insert(0, str(ROOT))
import bpe_ref_port as port  # noqa: E402

EOT = "<|endoftext|>"


def merges_sha256(merges) -> str:
    h = hashlib.sha256()
    for a, b in merges:
        h.update(a.hex().encode() + b" " + b.hex().encode() + b"\n")
    return h.hexdigest()


def main():
    nbytes = int(sys.argv[1]) if len(sys.argv) > 1 else 128_000_000
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    vocab_size = 32_000
    so = ROOT / "tools" / "libgencorpus.so"
    if not so.exists():
        subprocess.check_call(["gcc", "-O2", "-shared", "-fPIC", "-pthread", "-o", str(so), str(ROOT / "tools" / "gen_corpus.c")])
    lib = ctypes.CDLL(str(so))
    lib.gen_corpus.restype = ctypes.c_uint64
    path = f"/tmp/c5_owt_{seed}_{nbytes}.txt"
    n = lib.gen_corpus(ctypes.c_int(1), ctypes.c_uint64(seed), ctypes.c_uint64(nbytes), path.encode(), ctypes.c_int(os.cpu_count() or 1))
    assert n == os.path.getsize(path)
    data_sha = hashlib.sha256(open(path, "rb").read()).hexdigest()
    t0 = time.time()
    words = port.count_pretokens(path, os.cpu_count(), [EOT])
    n_distinct, n_tokens = len(words), sum(words.values())
    print(f"counted {n_tokens} pretokens, {n_distinct} distinct in {time.time() - t0:.0f} s", flush=True)
    t0 = time.time()
    vocab, merges = port.train_from_counts(words, vocab_size, [EOT])
    print(f"trained {len(merges)} merges in {time.time() - t0:.0f} s", flush=True)
    import bench  # noqa: E402  (the threaded C-oracle encoder lives there)

    t0 = time.time()
    ids_sha, n_ids = bench.c_oracle_ids_sha256(path, vocab, merges)
    print(f"encoded {n_ids} ids in {time.time() - t0:.0f} s", flush=True)
    out = {"style": 1, "seed": seed, "bytes": nbytes, "len": n, "corpus_sha256": data_sha, "vocab_size": vocab_size, "specials": [EOT],
           "distinct_pretokens": n_distinct, "pretokens": n_tokens, "merges": len(merges), "merges_sha256": merges_sha256(merges),
           "ids": n_ids, "ids_sha256": ids_sha, "last_merges": [[a.hex(), b.hex()] for a, b in merges[-5:]]}
    raw = json.dumps(out, indent=1).encode()
    with open(ROOT / "tests" / "golden" / "c5_owt_sha256.json.gz", "wb") as fh:
        with gzip.GzipFile(fileobj=fh, mode="wb", mtime=0) as gz:
            gz.write(raw)
    print(raw.decode())
    os.unlink(path)


if __name__ == "__main__":
    main()
