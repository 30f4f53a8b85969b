#!/usr/bin/env python3
"""Make tests/golden/* by running the UNMODIFIED reference from /root/reference.  TEST INFRASTRUCTURE ONLY.

Run in the build container (the reference is not on the GPU box):

    python oracle/gen_golden.py

What it writes (all gzip'd JSON, byte strings as hex):
  texts.json.gz        the reference's own test inputs (tests/fixtures/*.txt, corpus.en) and inline strings
  gpt2_model.json.gz   tests/fixtures/gpt2_vocab.json + gpt2_merges.txt decoded to bytes
  ts10k_model.json.gz  notebooks/sample_data/bpe_tokenizer/{vocab,merges}.pkl (10,000 entries / 9,743 merges)
  course_golden.json.gz  tests/fixtures/train-bpe-reference-{merges.txt,vocab.json} decoded to bytes
  pretok.json.gz       parallel_pretokenization Counters
  pretok_a25.json.gz   serial_pretokenization / pretokenize(training=False) Counters, strict-decode errors (`gen_golden.py a25` makes only this one)
  train.json.gz        (vocab, merges) of train_bpe / BPETrainer.train, incl. vocab sizes where the reference
                       departs from textbook BPE and seeded synthetic corpora
  encode.json.gz       BPETokenizer.encode / encode_iterable ids (plus offline-tiktoken ids for GPT-2 cases)
"""
from __future__ import annotations

import ctypes
import gzip
import hashlib
import json
import os
import pathlib
import pickle
import subprocess
import sys
import tempfile

HERE = pathlib.Path(__file__).resolve().parent
ROOT = HERE.parent
OUT = ROOT / "tests" / "golden"
sys.path.insert(0, str(HERE))
import ref_harness as H  # noqa: E402

EOT = "<|endoftext|>"


def dump(name, obj):
    OUT.mkdir(parents=True, exist_ok=True)
    raw = json.dumps(obj, separators=(",", ":")).encode()
    with open(OUT / name, "wb") as fh:
        with gzip.GzipFile(fileobj=fh, mode="wb", mtime=0) as gz:
            gz.write(raw)
    print(f"{name}: {len(raw)} B json, {(OUT / name).stat().st_size} B on disk")


def synth(style: int, seed: int, nbytes: int) -> bytes:
    so = ROOT / "tools" / "libgencorpus.so"
    subprocess.check_call(["gcc", "-O2", "-shared", "-fPIC", "-pthread", "-o", str(so), str(ROOT / "tools" / "gen_corpus.c")])
    lib = ctypes.CDLL(str(so))
    lib.gen_corpus.restype = ctypes.c_uint64
    with tempfile.NamedTemporaryFile(suffix=".txt", delete=False) as tf:
        path = tf.name
    n = lib.gen_corpus(ctypes.c_int(style), ctypes.c_uint64(seed), ctypes.c_uint64(nbytes), path.encode(), ctypes.c_int(4))
    data = open(path, "rb").read()
    os.unlink(path)
    assert n == len(data)
    return data


INLINE = {
    "empty": "",
    "single_s": "s",
    "emoji": "🙃",
    "ascii": "Hello, how are you?",
    "unicode": "Héllò hôw are ü? 🙃",
    "unicode_specials": "Héllò hôw <|endoftext|><|endoftext|> are ü? 🙃<|endoftext|>",
    "overlap_specials": "Hello, how <|endoftext|><|endoftext|> are you?<|endoftext|>",
    "notebook_known_answer": "Hello, I am encoding my text. How are you?",
    "contractions": "x's 's\n's  's !'s ''s x'lx x've'll I'm we'd THEY'RE don't 'tis",
    "whitespace": "a\n  b\n\n\nc  \t\td \r\n e  f 　g h   ",
    "runs": "aaaa aaaa aaaaaa bababab abab ....... !!!! ?!?!?! aaa aaaaa",
    "scripts": "Привет мир 你好世界 こんにちは ١٢٣ ½ ⅔ x² नमस्ते 한국어 é ​ zero",
    "no_trailing_newline": "first line\nsecond line<|endoftext|>third\n\nfourth",
}


def gen_a25(texts, synth_data):
    """pretok_a25.json.gz: the public wrappers of SURVEY.md §8 row A25 — serial_pretokenization (four chunks,
    STRICT decode, pretokenization.py:255-298) and pretokenize(..., training=False) (special tokens kept as parts and
    then pretokenized like any text, :198-207, :231) — run from the unmodified reference."""
    H.load()
    from bpe_transformer.tokenization.preprocessing.pretokenization import pretokenize, serial_pretokenization

    def get_bytes(key: str) -> bytes:
        if key.startswith("dirty:"):
            base = get_bytes(key[len("dirty:"):])
            return base[:1000] + b"\xff\xfe" + base[1000:2000] + b"\xe4\xbd" + base[2000:]
        return synth_data[key] if key.startswith("synth:") else bytes.fromhex(texts[key])

    rows = []
    for key, fn, training, parallel, specials, workers in [
        ("tinystories_sample.txt", "serial", True, None, [EOT], None),
        ("tinystories_sample.txt", "serial", True, None, [], None),
        ("corpus.en", "serial", True, None, [EOT], None),
        ("synth:ts:5:300000", "serial", True, None, [EOT], None),
        ("synth:owt:4:300000", "serial", False, None, [EOT], None),
        ("tinystories_sample.txt", "pretokenize", False, True, [EOT], 4),
        ("tinystories_sample.txt", "pretokenize", False, False, [EOT], 4),
        ("tinystories_sample.txt", "pretokenize", True, False, [EOT], 4),
        ("inline:overlap_specials", "pretokenize", False, True, [EOT, EOT + EOT], 1),
        ("inline:unicode_specials", "pretokenize", False, False, [EOT], 1),
        ("synth:ts:5:300000", "pretokenize", False, True, [EOT], 4),
        ("dirty:synth:ts:5:300000", "serial", True, None, [EOT], None),
        ("dirty:synth:ts:5:300000", "pretokenize", True, False, [EOT], 4),
        ("dirty:synth:ts:5:300000", "pretokenize", True, True, [EOT], 4),
        ("dirty:synth:ts:5:300000", "pretokenize", False, True, [EOT], 4),
    ]:
        tf = tempfile.NamedTemporaryFile(suffix=".txt", delete=False)
        tf.write(get_bytes(key))
        tf.close()
        row = {"input": key, "fn": fn, "training": training, "parallel": parallel, "specials": specials, "n_workers": workers}
        try:
            if fn == "serial":
                c = serial_pretokenization(file_path=tf.name, training=training, special_tokens=specials)
            else:
                c = pretokenize(file_path=tf.name, training=training, parallel_processing=parallel, n_workers=workers, special_tokens=specials)
            row["counts"] = sorted([bytes(k).hex(), v] for k, v in c.items())
        except UnicodeDecodeError:
            row["raises"] = "UnicodeDecodeError"
        os.unlink(tf.name)
        rows.append(row)
        print("a25", key, fn, training, parallel, row.get("raises", len(row.get("counts", []))))
    dump("pretok_a25.json.gz", rows)


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "a25":
        import gzip as _gz
        with _gz.open(OUT / "texts.json.gz", "rb") as fh:
            texts = json.load(fh)
        gen_a25(texts, {"synth:ts:5:300000": synth(0, 5, 300_000), "synth:owt:4:300000": synth(1, 4, 300_000)})
        return
    ref = H.load()
    from bpe_transformer.tokenization.bpe_tokenizer import BPETokenizer
    from bpe_transformer.tokenization.bpe_trainer import BPETrainer
    from bpe_transformer.tokenization.preprocessing.pretokenization import parallel_pretokenization

    # ---- inputs ------------------------------------------------------------------------------
    texts = {}
    for name in ["corpus.en", "address.txt", "german.txt", "tinystories_sample.txt",
                 "special_token_trailing_newlines.txt", "special_token_double_newlines_non_whitespace.txt"]:
        texts[name] = (H.FIXTURES / name).read_bytes().hex()
    for k, v in INLINE.items():
        texts["inline:" + k] = v.encode("utf-8").hex()
    dump("texts.json.gz", texts)

    gv, gm = H.load_gpt2_model()
    dump("gpt2_model.json.gz", {"vocab": [gv[i].hex() for i in range(len(gv))], "merges": [[a.hex(), b.hex()] for a, b in gm]})
    with open(H.SAMPLE_MODEL / "vocab.pkl", "rb") as fh:
        tv = pickle.load(fh)
    with open(H.SAMPLE_MODEL / "merges.pkl", "rb") as fh:
        tm = pickle.load(fh)
    dump("ts10k_model.json.gz", {"vocab": {str(k): v.hex() for k, v in tv.items()}, "merges": [[a.hex(), b.hex()] for a, b in tm]})

    dec = H.gpt2_byte_decoder()
    with open(H.FIXTURES / "train-bpe-reference-merges.txt", encoding="utf-8") as fh:
        cm = [[bytes(dec[c] for c in part).hex() for part in line.rstrip().split(" ")] for line in fh]
    with open(H.FIXTURES / "train-bpe-reference-vocab.json", encoding="utf-8") as fh:
        cv = {str(idx): bytes(dec[c] for c in tok).hex() for tok, idx in json.load(fh).items()}
    dump("course_golden.json.gz", {"merges": cm, "vocab": cv})

    # ---- synthetic inputs (regenerated by the tests from the spec; the hash pins the generator) ----
    synth_specs = {
        "synth:ts:1:1000000": (0, 1, 1_000_000),
        "synth:ts:2:6000000": (0, 2, 6_000_000),
        "synth:owt:3:2000000": (1, 3, 2_000_000),
        "synth:owt:4:300000": (1, 4, 300_000),
        "synth:ts:5:300000": (0, 5, 300_000),
    }
    synth_data = {k: synth(*v) for k, v in synth_specs.items()}
    dump("synth_sha256.json.gz", {k: {"style": v[0], "seed": v[1], "bytes": v[2], "len": len(synth_data[k]),
                                      "sha256": hashlib.sha256(synth_data[k]).hexdigest()} for k, v in synth_specs.items()})

    def as_file(data: bytes) -> str:
        tf = tempfile.NamedTemporaryFile(suffix=".txt", delete=False)
        tf.write(data)
        tf.close()
        return tf.name

    def get_bytes(key: str) -> bytes:
        return synth_data[key] if key.startswith("synth:") else bytes.fromhex(texts[key])

    # ---- pretoken Counters --------------------------------------------------------------------
    pretok = []
    for key, specials, workers in [("corpus.en", [EOT], 4), ("tinystories_sample.txt", [EOT], 4), ("german.txt", [EOT], 2),
                                   ("inline:whitespace", [EOT], 1), ("inline:scripts", [EOT], 1), ("inline:contractions", [], 1),
                                   ("inline:overlap_specials", [EOT, EOT + EOT], 1),
                                   ("synth:owt:4:300000", [EOT], 4), ("synth:ts:5:300000", [EOT], 4), ("synth:ts:5:300000", [], 4)]:
        path = as_file(get_bytes(key))
        c = parallel_pretokenization(file_path=path, n_workers=workers, special_tokens=specials)
        os.unlink(path)
        pretok.append({"input": key, "specials": specials, "n_workers": workers,
                       "counts": sorted([bytes(k).hex(), v] for k, v in c.items())})
    dump("pretok.json.gz", pretok)

    gen_a25(texts, synth_data)

    # ---- training -----------------------------------------------------------------------------
    train = []
    for key, vs, specials, workers in [
        ("corpus.en", 500, [EOT], None), ("corpus.en", 1200, [EOT], None), ("corpus.en", 4000, [EOT], None),
        ("inline:runs", 280, [EOT], None), ("inline:scripts", 300, [EOT], None), ("inline:no_trailing_newline", 270, [EOT], None),
        ("tinystories_sample.txt", 400, [EOT], None), ("tinystories_sample.txt", 300, [], 4),
        ("synth:ts:1:1000000", 1500, [EOT], None), ("synth:owt:3:2000000", 2000, [EOT], None),
        ("synth:ts:2:6000000", 6000, [EOT], None),
    ]:
        path = as_file(get_bytes(key))
        if workers is None:
            vocab, merges = ref.train_bpe(path, vs, specials)
        else:
            tr = BPETrainer(vocab_size=vs, special_tokens=specials)
            tr.train(input_path=path, n_workers=workers)
            vocab, merges = tr.vocab, tr.merges
        os.unlink(path)
        train.append({"input": key, "vocab_size": vs, "specials": specials, "n_workers": workers,
                      "merges": [[a.hex(), b.hex()] for a, b in merges], "vocab": {str(k): v.hex() for k, v in vocab.items()}})
        print("train", key, vs, len(merges))
    dump("train.json.gz", train)

    # ---- encoding -----------------------------------------------------------------------------
    tk = H.offline_tiktoken_gpt2()
    enc = []
    def with_specials(vocab, sp):
        """Append specials missing from the vocab at id len(vocab), as the reference's own loaders do
        (bpe_tokenizer.py:311-320, tests/test_tokenizer.py:62-66)."""
        v = dict(vocab)
        for s in sp:
            if s.encode("utf-8") not in set(v.values()):
                v[len(v)] = s.encode("utf-8")
        return v

    gpt2 = lambda sp: BPETokenizer(vocab=with_specials(gv, sp), merges=list(gm), special_tokens=sp)  # noqa: E731
    ts10k = lambda sp: BPETokenizer(vocab=with_specials(tv, sp), merges=list(tm), special_tokens=sp)  # noqa: E731
    for model, mk in [("gpt2", gpt2), ("ts10k", ts10k)]:
        for key in list(texts) + ["synth:owt:4:300000", "synth:ts:5:300000"]:
            sp = [EOT, EOT + EOT] if key == "inline:overlap_specials" else [EOT]
            text = get_bytes(key).decode("utf-8")
            tok = mk(sp)
            row = {"model": model, "input": key, "specials": sp, "encode": tok.encode(text)}
            # encode_iterable over a file handle == an iterable of lines (universal newlines)
            path = as_file(get_bytes(key))
            with open(path) as fh:
                row["iterable_lines"] = list(mk(sp).encode_iterable(fh))
            os.unlink(path)
            # encode_iterable over arbitrary chunks that ignore line structure
            chunks = [text[i:i + 997] for i in range(0, len(text), 997)]
            row["iterable_chunks_997"] = list(mk(sp).encode_iterable(chunks))
            if model == "gpt2" and len(sp) == 1:
                row["tiktoken"] = tk.encode(text, allowed_special={EOT})
            # de-duplicate identical id lists to keep the fixture small: "=key" means "same as row[key]"
            for k, base in [("iterable_chunks_997", "iterable_lines"), ("iterable_chunks_997", "encode"),
                            ("iterable_lines", "encode"), ("tiktoken", "encode")]:
                if k in row and isinstance(row[k], list) and row[k] == row[base]:
                    row[k] = "=" + base
            enc.append(row)
    dump("encode.json.gz", enc)


if __name__ == "__main__":
    main()
