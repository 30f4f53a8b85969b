"""Parity of the CUDA path with the oracle and the reference's golden vectors — runs on a B200 (`-m gpu`).

Everything here goes through the C-ABI of libbpe_b200.so (bpe_transformer_b200._lib.Engine is a thin ctypes
wrapper: one method per exported function) or through the reference-shaped Python API on top of it.  The bar is
bit-exact: identical start masks, Counters, ordered merges, vocab dicts and token ids.
"""
from __future__ import annotations

import os
import random
from collections import Counter

import numpy as np
import pytest

import bpe_ref_port as port
import c_oracle
from conftest import EOT, expand_row, load_golden, synth_corpus, with_specials

pytestmark = pytest.mark.gpu

PRETOK = load_golden("pretok.json.gz")
TRAIN = load_golden("train.json.gz")
ENCODE = load_golden("encode.json.gz")
A25 = load_golden("pretok_a25.json.gz")


@pytest.fixture(scope="module")
def engine():
    from bpe_transformer_b200 import _lib

    return _lib.get_engine()


def gpu_mask(engine, data: bytes, specials=None, cuts=None, flags=0) -> np.ndarray:
    """The start mask from bpe_scan, checked against the one the fused one-pass kernel (bpe_pretokenize_count) writes:
    two implementations of the same rule (a CTA per tile with shared memory there, a warp per window with shuffles here)."""
    from bpe_transformer_b200 import _lib

    engine.set_special_tokens(_lib.sort_specials(specials))
    d_text, n = engine.to_device(data)
    mask = engine.scan(d_text, n, cuts=cuts, flags=flags)
    out = mask.cpu().numpy().view(np.uint32)[:(n + 31) // 32]
    if n:
        engine.unique_reset()
        fused = engine.pretokenize_count(d_text, n, cuts=cuts, flags=flags, want_mask=True).cpu().numpy().view(np.uint32)[:(n + 31) // 32]
        assert np.array_equal(fused, out), "fused kernel's mask differs from bpe_scan's: " + first_diff(fused, out, data)
    return out


def gpu_counter(engine, data: bytes, specials=None, cuts=None, flags=0) -> Counter:
    """The Counter from bpe_scan + bpe_count, checked against the one from the fused one-pass kernel."""
    from bpe_transformer_b200 import _lib
    from bpe_transformer_b200.tokenization.preprocessing.pretokenization import counter_from_engine

    engine.set_special_tokens(_lib.sort_specials(specials))
    engine.unique_reset()
    d_text, n = engine.to_device(data)
    engine.count(d_text, n, engine.scan(d_text, n, cuts=cuts, flags=flags))
    two_pass = counter_from_engine(engine)
    stats = engine.unique_stats()
    engine.unique_reset()
    engine.pretokenize_count(d_text, n, cuts=cuts, flags=flags)
    fused = counter_from_engine(engine)
    if fused != two_pass:
        only = [(bytes(k), v, two_pass.get(k)) for k, v in fused.items() if two_pass.get(k) != v][:5]
        missing = [(bytes(k), v) for k, v in two_pass.items() if k not in fused][:5]
        raise AssertionError(f"fused counter differs: fused-vs-two-pass {only}, missing from fused {missing}")
    assert engine.unique_stats() == stats
    return two_pass


def first_diff(a: np.ndarray, b: np.ndarray, data: bytes) -> str:
    bad = np.flatnonzero(a != b)
    if not len(bad):
        return ""
    w = int(bad[0])
    lo = max(0, w * 32 - 24)
    return f"word {w}: gpu {int(a[w]):032b} oracle {int(b[w]):032b} text {data[lo:w * 32 + 40]!r}"


# ---- scan ---------------------------------------------------------------------------------------------------
ALPHABET = [" ", " ", "  ", "\n", "\t", "'", "'s", "'ll", "'ve", "'re", "'d", "'m", "'t", "a", "b", "s", "l", "v", "e", "r",
            "L", "1", "23", ".", "!", "?", "é", "ü", " ", "　", " ", "你", "好", "٣", "½", "🙃", "x", "\r", "-", "_",
            "\x1c", "\x85", " ", " ", "ǅ", "ʰ", "़", "​", "Ⅷ", "<|endoftext|>", "<|", "<", "|>"]


@pytest.mark.parametrize("seed", [0, 1, 2])
@pytest.mark.parametrize("specials", [[], [EOT], [EOT, EOT + EOT, "<|"]])
def test_scan_fuzz_matches_oracle(engine, seed, specials):
    """Many short random strings in one buffer, separated by caller cuts (every string is its own piece)."""
    rnd = random.Random(seed)
    pieces, cuts, total = [], [], 0
    for _ in range(4000):
        s = "".join(rnd.choice(ALPHABET) for _ in range(rnd.randint(1, 40))).encode("utf-8")
        pieces.append(s)
        total += len(s)
        cuts.append(total)
    data = b"".join(pieces)
    want = c_oracle.scan(data, specials, cuts[:-1])
    got = gpu_mask(engine, data, specials, cuts[:-1])
    assert np.array_equal(got, want), first_diff(got, want, data)
    # and as one long piece, and cut after newlines
    for flags in (0, c_oracle.CUT_AFTER_NEWLINE):
        want = c_oracle.scan(data, specials, None, flags)
        got = gpu_mask(engine, data, specials, None, flags)
        assert np.array_equal(got, want), first_diff(got, want, data)


@pytest.mark.parametrize("key", ["corpus.en", "german.txt", "tinystories_sample.txt", "inline:scripts", "inline:whitespace",
                                 "inline:contractions", "synth:owt:4:300000", "synth:ts:5:300000", "synth:owt:3:2000000"])
def test_scan_texts_match_oracle(engine, input_bytes, key):
    data = input_bytes(key)
    for specials in ([], [EOT]):
        for flags in (0, c_oracle.CUT_AFTER_NEWLINE):
            want = c_oracle.scan(data, specials, None, flags)
            got = gpu_mask(engine, data, specials, None, flags)
            assert np.array_equal(got, want), first_diff(got, want, data)


@pytest.mark.parametrize("n", [0, 1, 31, 32, 33, 63, 64, 65, 8191, 8192, 8193, 16384 + 5])
def test_scan_edge_sizes(engine, n):
    """Empty and ragged sizes around the 32-byte chunk and the 8 KiB tile."""
    data = synth_corpus(1, 77, 40000)[:n]
    while data and (data[-1] & 0xC0) == 0x80:
        data = data[:-1]
    if data and data[-1] >= 0xC0:
        data = data[:-1]
    if not data:
        assert gpu_mask(engine, data).size == 0
        return
    want = c_oracle.scan(data, [EOT])
    got = gpu_mask(engine, data, [EOT])
    assert np.array_equal(got, want), first_diff(got, want, data)


def test_scan_long_runs_and_tile_borders(engine):
    """Tokens far longer than a tile, specials and contractions sitting on chunk / tile borders."""
    parts = [b"a" * 20000, b" " * 9000, b"x", b"-" * 8190, EOT.encode(), b"'ll" * 3000, ("é" * 5000).encode(), b"\n" * 8193, b"7" * 100]
    data = b"".join(parts)
    for shift in range(0, 40, 7):
        d = b"z" * shift + data
        want = c_oracle.scan(d, [EOT])
        got = gpu_mask(engine, d, [EOT])
        assert np.array_equal(got, want), first_diff(got, want, d)


def test_scan_reports_invalid_utf8(engine):
    from bpe_transformer_b200 import _lib

    for bad in (b"abc \xff def", b"abc \xed\xa0\x80 def", b"abc \xe4\xbd", b"\x80abc", b"abc\xc0\xaf"):
        with pytest.raises(_lib.InvalidUtf8):
            gpu_mask(engine, bad * 3)


# ---- count --------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("case", PRETOK, ids=lambda c: f"{c['input']}-{len(c['specials'])}sp")
def test_counts_match_reference_golden(engine, case, input_bytes, tmp_file):
    from bpe_transformer_b200.tokenization.preprocessing import parallel_pretokenization

    data = input_bytes(case["input"])
    want = Counter({tuple(bytes.fromhex(h)): c for h, c in case["counts"]})
    if case["specials"] or case["n_workers"] == 1:
        assert gpu_counter(engine, data, case["specials"]) == want
    if not case["specials"] and case["n_workers"] > (os.cpu_count() or 1):
        pytest.skip("without special tokens the result depends on the worker count, which cpu_count() clamps on this box")
    got = parallel_pretokenization(file_path=tmp_file(data), n_workers=case["n_workers"], special_tokens=case["specials"])
    assert got == want


@pytest.mark.parametrize("case", A25, ids=lambda c: f"{c['input']}-{c['fn']}-{'train' if c['training'] else 'keep'}-{c['parallel']}-{len(c['specials'])}sp")
def test_public_pretokenize_wrappers_match_reference_golden(case, input_bytes, tmp_file):
    """serial_pretokenization (strict decode -> UnicodeDecodeError, pretokenization.py:284) and
    pretokenize(..., training=False) (special tokens kept as parts and pretokenized like text, :198-207, :231)."""
    from bpe_transformer_b200.tokenization.preprocessing import pretokenize, serial_pretokenization
    from test_oracle_pinned import a25_input

    path = tmp_file(a25_input(input_bytes, case["input"]))

    def run():
        if case["fn"] == "serial":
            return serial_pretokenization(file_path=path, training=case["training"], special_tokens=case["specials"])
        return pretokenize(file_path=path, training=case["training"], parallel_processing=case["parallel"],
                           n_workers=case["n_workers"], special_tokens=case["specials"])

    if case.get("raises"):
        with pytest.raises(UnicodeDecodeError):
            run()
    else:
        assert run() == Counter({tuple(bytes.fromhex(h)): c for h, c in case["counts"]})


@pytest.mark.parametrize("style,seed,size", [(0, 11, 8_000_000), (1, 12, 8_000_000)])
def test_counts_match_oracle_synthetic(engine, style, seed, size):
    data = synth_corpus(style, seed, size)
    assert gpu_counter(engine, data, [EOT]) == c_oracle.count(data, [EOT])


@pytest.mark.parametrize("seed", [0, 1])
@pytest.mark.parametrize("specials", [[], [EOT], [EOT, EOT + EOT, "<|"]])
def test_count_fuzz_matches_oracle(engine, seed, specials):
    """The fuzz strings of the scan test as one text, counted: with cuts, as one piece, cut after newlines."""
    rnd = random.Random(100 + seed)
    pieces, cuts, total = [], [], 0
    for _ in range(3000):
        s = "".join(rnd.choice(ALPHABET) for _ in range(rnd.randint(1, 60))).encode("utf-8")
        pieces.append(s)
        total += len(s)
        cuts.append(total)
    data = b"".join(pieces)
    assert gpu_counter(engine, data, specials, cuts[:-1]) == c_oracle.count(data, specials, cuts[:-1])
    for flags in (0, c_oracle.CUT_AFTER_NEWLINE):
        assert gpu_counter(engine, data, specials, None, flags) == c_oracle.count(data, specials, None, flags)


def test_count_long_tokens_and_window_borders(engine):
    """Pretokens far longer than a lane, a window, a CTA's range; special tokens, contractions and multi-byte characters on
    window borders; every shift of the text against the 928-byte windows."""
    parts = [b"a" * 20000, b" " * 9000, b"x", b"-" * 8190, EOT.encode(), b"'ll" * 3000, ("\u00e9" * 5000).encode(), b"\n" * 8193, b"7" * 100,
             b" word" * 500, b"q" * 70000, b" tail", EOT.encode(), b"z" * 927, b" ", b"y" * 929, b"!" * 31, b"w" * 33]
    data = b"".join(parts)
    for shift in list(range(0, 40, 7)) + [900, 927, 928, 929]:
        d = b"z" * shift + data
        assert gpu_counter(engine, d, [EOT]) == c_oracle.count(d, [EOT])
    huge = b"ab " + b"k" * 3_000_000 + b" cd " + b" " * 1_000_000 + b"e"
    assert gpu_counter(engine, huge, [EOT]) == c_oracle.count(huge, [EOT])


@pytest.mark.parametrize("n", [0, 1, 31, 32, 33, 63, 64, 65, 927, 928, 929, 960, 8191, 8192, 8193, 16 * 928, 16 * 928 + 1, 16384 + 5])
def test_count_edge_sizes(engine, n):
    data = synth_corpus(1, 78, 40000)[:n]
    while data and (data[-1] & 0xC0) == 0x80:
        data = data[:-1]
    if data and data[-1] >= 0xC0:
        data = data[:-1]
    for specials in ([], [EOT]):
        assert gpu_counter(engine, data, specials) == c_oracle.count(data, specials)


def test_key_collisions_are_reported_not_merged(engine, gpt2_model):
    """The table of distinct pretokens is found by 64-bit key but identified by BYTES.  With the keys narrowed to 10 bits (test
    hook) thousands of distinct pretokens share keys: counting, importing and encoding must raise the collision error, never
    hand back merged counts or another pretoken's ids.  With the full key again, everything is as before."""
    from bpe_transformer_b200 import _lib
    from adapters import get_tokenizer

    data = synth_corpus(1, 55, 2_000_000)
    want = c_oracle.count(data, [EOT])
    assert len(want) > 20_000
    engine.set_special_tokens(_lib.sort_specials([EOT]))
    d_text, n = engine.to_device(data)
    engine.unique_reset()
    engine.pretokenize_count(d_text, n)
    blob, off, ln, cnt = engine.unique_export()
    try:
        engine.debug_key_mask(0x3FF)
        for count in (lambda: engine.pretokenize_count(d_text, n), lambda: engine.count(d_text, n, engine.scan(d_text, n)),
                      lambda: engine.unique_import(blob, off, ln, cnt)):
            engine.unique_reset()
            with pytest.raises(_lib.BpeError) as err:
                count()
            assert err.value.code == _lib.ERR_HASH_COLLISION
        engine.unique_reset()
        vocab, merges = gpt2_model
        tok = get_tokenizer(with_specials(vocab, [EOT]), merges, [EOT])
        with pytest.raises(_lib.BpeError) as err:
            tok.encode(data[:200_000].decode("utf-8", errors="ignore"))
        assert err.value.code == _lib.ERR_HASH_COLLISION
    finally:
        engine.debug_key_mask(0xFFFFFFFFFFFFFFFF)
        engine.unique_reset()
    assert gpu_counter(engine, data, [EOT]) == want


def test_count_accumulates_over_shards(engine):
    """bpe_count per shard (cut at special-token starts) == one call over the whole text."""
    data = synth_corpus(0, 21, 3_000_000)
    whole = gpu_counter(engine, data, [EOT])
    from bpe_transformer_b200 import _lib
    from bpe_transformer_b200.tokenization.preprocessing.pretokenization import counter_from_engine

    engine.set_special_tokens(_lib.sort_specials([EOT]))
    engine.unique_reset()
    cuts = [0]
    for k in range(1, 5):
        cuts.append(data.index(EOT.encode(), k * len(data) // 5))
    cuts.append(len(data))
    for a, b in zip(cuts[:-1], cuts[1:]):
        d_text, n = engine.to_device(data[a:b])
        engine.count(d_text, n, engine.scan(d_text, n))
    assert counter_from_engine(engine) == whole
    # export -> import into a fresh table doubles every count
    blob, off, ln, cnt = engine.unique_export()
    engine.unique_import(blob, off, ln, cnt)
    doubled = counter_from_engine(engine)
    assert doubled == Counter({k: 2 * v for k, v in whole.items()})


# ---- train --------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("case", TRAIN, ids=lambda c: f"{c['input']}-{c['vocab_size']}")
def test_train_matches_reference_golden(case, input_bytes, tmp_file):
    from bpe_transformer_b200.tokenization import BPETrainer
    from adapters import run_train_bpe

    data = input_bytes(case["input"])
    want_merges = [(bytes.fromhex(a), bytes.fromhex(b)) for a, b in case["merges"]]
    want_vocab = {int(k): bytes.fromhex(v) for k, v in case["vocab"].items()}
    path = tmp_file(data)
    if case["n_workers"] is None:
        vocab, merges = run_train_bpe(path, case["vocab_size"], case["specials"])
    else:
        tr = BPETrainer(vocab_size=case["vocab_size"], special_tokens=case["specials"])
        tr.train(input_path=path, n_workers=case["n_workers"])
        vocab, merges = tr.vocab, tr.merges
    bad = next((i for i, (x, y) in enumerate(zip(merges, want_merges)) if x != y), None)
    assert bad is None, f"merge {bad}: gpu {merges[bad]} reference {want_merges[bad]}"
    assert merges == want_merges
    assert vocab == want_vocab


def test_train_course_golden_c1(input_bytes, tmp_file):
    """BASELINE config 1: the reference's own test (tests/test_train_bpe.py:27-63)."""
    from adapters import run_train_bpe

    g = load_golden("course_golden.json.gz")
    vocab, merges = run_train_bpe(tmp_file(input_bytes("corpus.en")), 500, [EOT])
    assert merges == [(bytes.fromhex(a), bytes.fromhex(b)) for a, b in g["merges"]]
    assert set(vocab.keys()) == {int(k) for k in g["vocab"]}
    assert set(vocab.values()) == {bytes.fromhex(v) for v in g["vocab"].values()}


@pytest.mark.parametrize("style,seed,size,vocab_size", [(0, 31, 4_000_000, 3000), (1, 32, 1_500_000, 1200)])
def test_train_matches_port_on_fresh_corpus(style, seed, size, vocab_size, tmp_path):
    from adapters import run_train_bpe

    path = synth_corpus(style, seed, size, str(tmp_path / "c.txt"))
    words = port.count_pretokens(path, 8, [EOT])
    want_vocab, want_merges = port.train_from_counts(words, vocab_size, [EOT])
    vocab, merges = run_train_bpe(path, vocab_size, [EOT])
    bad = next((i for i, (x, y) in enumerate(zip(merges, want_merges)) if x != y), None)
    assert bad is None, f"merge {bad}: gpu {merges[bad]} port {want_merges[bad]}"
    assert (vocab, merges) == (want_vocab, want_merges)


@pytest.mark.parametrize("style,seed,size,vocab_size", [(1, 33, 16_000_000, 2500), (0, 34, 32_000_000, 6000)])
def test_train_matches_c_oracle_on_larger_corpora(style, seed, size, vocab_size, tmp_path):
    """OpenWebText-shaped (2*10^5 distinct pretokens, many light steps with deferred updates) and TinyStories-shaped
    corpora, thousands of merges: ordered merges and vocab equal the C oracle's."""
    from adapters import run_train_bpe

    path = synth_corpus(style, seed, size, str(tmp_path / "c.txt"))
    want_vocab, want_merges = c_oracle.train(c_oracle.count(open(path, "rb").read(), [EOT]), vocab_size, [EOT])
    vocab, merges = run_train_bpe(path, vocab_size, [EOT])
    bad = next((i for i, (x, y) in enumerate(zip(merges, want_merges)) if x != y), None)
    assert bad is None, f"merge {bad}: gpu {merges[bad]} oracle {want_merges[bad]}"
    assert (vocab, merges) == (want_vocab, want_merges)


def test_train_and_encode_owt_shaped_32k_vocab_golden(tmp_path):
    """BASELINE configs[4] at a size a CPU finishes: OpenWebText-shaped text (128 MB, 8*10^5 distinct pretokens), vocab_size
    32000.  Ordered merges and line-by-line ids against hashes made by oracle/gen_c5_golden.py (the port of the reference's
    algorithm, pinned to the reference by tests/test_oracle_pinned.py; ids by the C oracle)."""
    import hashlib

    from adapters import get_tokenizer, run_train_bpe

    g = load_golden("c5_owt_sha256.json.gz")
    path = synth_corpus(g["style"], g["seed"], g["bytes"], str(tmp_path / "owt.txt"))
    data = open(path, "rb").read()
    assert len(data) == g["len"] and hashlib.sha256(data).hexdigest() == g["corpus_sha256"], "synthetic generator is not reproducible"
    vocab, merges = run_train_bpe(path, g["vocab_size"], g["specials"])
    assert len(merges) == g["merges"]
    assert [[a.hex(), b.hex()] for a, b in merges[-5:]] == g["last_merges"]
    h = hashlib.sha256()
    for a, b in merges:
        h.update(a.hex().encode() + b" " + b.hex().encode() + b"\n")
    assert h.hexdigest() == g["merges_sha256"]
    assert vocab[256] == EOT.encode() and all(vocab[257 + k] == a + b for k, (a, b) in enumerate(merges))
    tok = get_tokenizer(vocab, merges, g["specials"])
    ids = tok.encode_file_to_array(path)
    assert ids.dtype == np.uint16 and len(ids) == g["ids"]
    assert hashlib.sha256(ids.tobytes()).hexdigest() == g["ids_sha256"]
    assert tok.decode_array(ids) == data


def test_train_ignores_invalid_utf8_like_the_reference(tmp_file):
    """pretokenization.py:194 decodes with errors="ignore"."""
    from adapters import run_train_bpe

    data = synth_corpus(0, 41, 200_000)
    dirty = data[:1000] + b"\xff\xfe" + data[1000:50_000] + b"\xe4\xbd" + data[50_000:]
    path = tmp_file(dirty)
    want = port.train_from_counts(port.count_pretokens(path, 4, [EOT]), 600, [EOT])
    assert run_train_bpe(path, 600, [EOT]) == want


def test_train_error_conventions(tmp_file):
    from bpe_transformer_b200 import train_bpe
    from bpe_transformer_b200.tokenization import BPETrainer

    with pytest.raises(ValueError):
        train_bpe(tmp_file(b"abc"), 200, [EOT])  # main.py:12-13
    with pytest.raises(ValueError):
        BPETrainer(vocab_size=100)  # bpe_trainer.py:82-83
    with pytest.raises(AttributeError):
        train_bpe(tmp_file(b"a\nb\nc"), 300, [EOT])  # single-byte pretokens only: no pairs (:237-240)
    with pytest.raises(AttributeError):
        train_bpe(tmp_file(b""), 300, [EOT])
    with pytest.raises(FileNotFoundError):
        train_bpe("/nonexistent/corpus.txt", 300, [EOT])
    vocab, merges = train_bpe(tmp_file(b"ab ab ab"), 257, [EOT])  # room for nothing: initial vocab comes back
    assert len(vocab) == 257 and merges == []


def test_save_and_load_roundtrip(tmp_path, input_bytes, tmp_file):
    from bpe_transformer_b200.tokenization import BPETokenizer, BPETrainer

    tr = BPETrainer(vocab_size=400, special_tokens=[EOT])
    tr.train(input_path=tmp_file(input_bytes("tinystories_sample.txt")), n_workers=2)
    tr.save_trainer(output_dir=tmp_path / "tok")
    tok = BPETokenizer.from_files(tmp_path / "tok" / "vocab.pkl", tmp_path / "tok" / "merges.pkl", special_tokens=[EOT])
    assert tok.vocab == tr.vocab and tok.merges == tr.merges
    text = input_bytes("tinystories_sample.txt").decode()
    assert tok.decode(tok.encode(text)) == text


# ---- encode -------------------------------------------------------------------------------------------------
def _models(gpt2_model, ts10k_model):
    return {"gpt2": gpt2_model, "ts10k": ts10k_model}


@pytest.mark.parametrize("row", ENCODE, ids=lambda r: f"{r['model']}-{r['input']}")
def test_encode_matches_reference_golden(row, input_bytes, gpt2_model, ts10k_model, tmp_file):
    from adapters import get_tokenizer

    vocab, merges = _models(gpt2_model, ts10k_model)[row["model"]]
    tok = get_tokenizer(with_specials(vocab, row["specials"]), merges, row["specials"])
    data = input_bytes(row["input"])
    text = data.decode("utf-8")
    want = expand_row(row, "encode")
    got = tok.encode(text)
    bad = next((i for i, (x, y) in enumerate(zip(got, want)) if x != y), None)
    assert bad is None and len(got) == len(want), f"id {bad}: gpu {got[bad:bad + 5] if bad is not None else len(got)} reference {want[bad:bad + 5] if bad is not None else len(want)}"
    assert tok.decode(got) == text
    with open(tmp_file(data)) as fh:  # a file handle: an iterable of lines
        assert list(tok.encode_iterable(fh)) == expand_row(row, "iterable_lines")
    translated = text.replace("\r\n", "\n").replace("\r", "\n")
    assert list(tok.encode_iterable(translated.splitlines(keepends=True))) == expand_row(row, "iterable_lines")
    chunks = [text[i:i + 997] for i in range(0, len(text), 997)]
    assert list(tok.encode_iterable(chunks)) == expand_row(row, "iterable_chunks_997")
    if len(data) < 50_000:
        assert tok.decode_array(np.asarray(want)) == b"".join(tok.vocab[i] for i in want)


@pytest.mark.parametrize("model", ["gpt2", "ts10k"])
@pytest.mark.parametrize("style,seed", [(0, 51), (1, 52)])
def test_encode_matches_oracle_synthetic(model, style, seed, gpt2_model, ts10k_model, tmp_path):
    from adapters import get_tokenizer

    vocab, merges = _models(gpt2_model, ts10k_model)[model]
    vocab = with_specials(vocab, [EOT])
    path = synth_corpus(style, seed, 3_000_000, str(tmp_path / "c.txt"))
    data = open(path, "rb").read()
    tok = get_tokenizer(vocab, merges, [EOT])
    ct = c_oracle.COracleTokenizer(vocab, merges, [EOT])
    want = np.asarray(ct.encode_lines(data.decode()), dtype=np.uint32)
    got = tok.encode_file_to_array(path)
    assert got.dtype == np.uint16
    assert np.array_equal(got.astype(np.uint32), want)
    want_whole = ct.encode(data.decode())
    assert tok.encode(data.decode()) == want_whole
    assert tok.decode_array(got) == data


def test_encode_with_a_caller_made_start_mask_equals_the_one_pass_path(engine, ts10k_model, tmp_path):
    """bpe_encode takes an optional start mask made by bpe_scan: that path (mask kernel, count, emit over the mask) and the
    one-pass path (no mask) must give the same ids, cold and warm."""
    from adapters import get_tokenizer
    from bpe_transformer_b200 import _lib

    vocab, merges = ts10k_model
    tok = get_tokenizer(with_specials(vocab, [EOT]), merges, [EOT])
    data = synth_corpus(1, 53, 2_000_000) + ("tail without newline " + EOT + " x" * 700 + "q" * 5000).encode()
    for flags in (0, _lib.CUT_AFTER_NEWLINE):
        engine.unique_reset()
        tok._bind(engine)
        d_text, n = engine.to_device(data)
        one_pass = [engine.encode(d_text, n, flags=flags).cpu().numpy().view(np.uint16) for _ in range(2)]
        engine.unique_reset()
        tok._device_model = None
        engine._model_token = None
        tok._bind(engine)
        mask = engine.scan(d_text, n, flags=flags)
        with_mask = [engine.encode(d_text, n, flags=flags, mask=mask).cpu().numpy().view(np.uint16) for _ in range(2)]
        want = np.asarray(c_oracle.COracleTokenizer(tok.vocab, merges, [EOT]).encode_bytes(data, flags=flags), dtype=np.uint16)
        for got in one_pass + with_mask:
            assert np.array_equal(got, want)


def test_encode_fail_on_cr_flag(engine, ts10k_model):
    """BPE_FAIL_ON_CR: bpe_encode refuses a text that holds a '\\r' anywhere (first byte, last byte, last byte of a 16-byte
    group, with or without a caller's mask) and encodes as usual when it holds none — also when a '\\r' sits just behind
    the end, in the buffer's padding."""
    from adapters import get_tokenizer
    from bpe_transformer_b200 import _lib

    vocab, merges = ts10k_model
    tok = get_tokenizer(with_specials(vocab, [EOT]), merges, [EOT])
    tok._bind(engine)
    base = synth_corpus(0, 5, 100_003)
    want = np.asarray(c_oracle.COracleTokenizer(tok.vocab, merges, [EOT]).encode_bytes(base, flags=1), dtype=np.uint16)
    flags = _lib.CUT_AFTER_NEWLINE | _lib.FAIL_ON_CR
    for cut in (len(base), 4096, 4097, 4111, 15, 1):
        d_text, n = engine.to_device(base[:cut] + b"\r\r\r")  # the three '\r' are beyond n: padding, not text
        got = engine.encode(d_text, cut, flags=flags).cpu().numpy().view(np.uint16)
        if cut == len(base):
            assert np.array_equal(got, want)
    for at in (0, 1, 15, 16, 4095, len(base) - 1):
        data = bytearray(base)
        data[at] = 13
        d_text, n = engine.to_device(bytes(data))
        for mask in (None, engine.scan(d_text, n, flags=1)):
            with pytest.raises(_lib.CarriageReturn):
                engine.encode(d_text, n, flags=flags, mask=mask)
        assert engine.encode(d_text, n, flags=1).numel() > 0  # without the flag the byte is ordinary whitespace
    d_text, n = engine.to_device(base)
    assert np.array_equal(engine.encode(d_text, n, flags=flags).cpu().numpy().view(np.uint16), want)  # the status word was cleared


def test_encode_with_a_freshly_trained_vocab(tmp_path):
    """train on the GPU, encode with the result, check both against the port."""
    from adapters import get_tokenizer, run_train_bpe

    path = synth_corpus(0, 61, 2_000_000, str(tmp_path / "c.txt"))
    vocab, merges = run_train_bpe(path, 2000, [EOT])
    text = open(path, encoding="utf-8").read()[:300_000]
    tok = get_tokenizer(vocab, merges, [EOT])
    assert tok.encode(text) == port.PortTokenizer(vocab, merges, [EOT]).encode(text)


def test_encode_error_conventions(gpt2_model):
    from adapters import get_tokenizer

    vocab, merges = gpt2_model
    tok = get_tokenizer(dict(vocab), merges, ["<|missing|>"])  # special token without an id: KeyError when met (:175)
    assert tok.encode("no special here") == get_tokenizer(dict(vocab), merges, None).encode("no special here")
    with pytest.raises(KeyError):
        tok.encode("a <|missing|> b")
    small = {k: v for k, v in vocab.items() if v != b"z"}
    with pytest.raises(KeyError):
        get_tokenizer(small, merges, None).encode("zebra")  # byte without an id (:206)
    assert get_tokenizer(dict(vocab), merges, None).decode([0, 10**9, 1]) == port.PortTokenizer(vocab, merges).decode([0, 10**9, 1])


def test_encode_dense_windows_and_warm_cache(gpt2_model):
    """Windows of 1 KiB with more tokens than the emit pass keeps on chip (one-byte tokens), and texts that bring
    new pretokens to a cache that earlier calls have filled: same ids as the oracle, call after call."""
    from adapters import get_tokenizer

    vocab, merges = gpt2_model
    vocab = with_specials(vocab, [EOT])
    tok = get_tokenizer(vocab, merges, [EOT])
    ct = c_oracle.COracleTokenizer(vocab, merges, [EOT])
    rng = np.random.default_rng(7)
    dense = "".join(rng.choice(list("a.b,c;d!e?f:g(h)1-2+"), size=40_000).tolist())
    words = " ".join("w%d" % v for v in rng.integers(0, 5000, size=20_000))
    later = " ".join("later%dx" % v for v in rng.integers(0, 3000, size=10_000)) + EOT + dense[:5000]
    for text in (dense, words, dense + EOT + words, later, dense, later + "\n" + words):
        assert tok.encode(text) == ct.encode(text)
    long_ids = "z" * 3000 + " " + "ab" * 2000  # two pretokens with hundreds of ids each
    assert tok.encode(long_ids) == ct.encode(long_ids)


def test_encode_with_ids_beyond_16_bits(gpt2_model, tmp_path):
    """A vocabulary whose ids do not fit uint16: ids come back as uint32, the same as the oracle's."""
    from adapters import get_tokenizer

    vocab, merges = gpt2_model
    shift = 70_000
    wide = {(k if k < 256 else k + shift): v for k, v in vocab.items()}
    wide[max(wide) + 1] = EOT.encode()
    tok = get_tokenizer(wide, merges, [EOT])
    ct = c_oracle.COracleTokenizer(wide, merges, [EOT])
    path = synth_corpus(0, 81, 1_000_000, str(tmp_path / "c.txt"))
    text = open(path, encoding="utf-8").read()
    want = ct.encode(text)
    assert max(want) > 0xFFFF
    assert tok.encode(text) == want
    got = tok.encode_file_to_array(path)
    assert got.dtype == np.uint32 and got.tolist() == ct.encode_lines(text)
    assert tok.decode(want[:5000]) == ct_decode(wide, want[:5000])
    assert tok.decode_array(got) == open(path, "rb").read()


def ct_decode(vocab, ids):
    return b"".join(vocab[i] for i in ids).decode("utf-8", errors="replace")


def test_two_tokenizers_of_equal_size_do_not_share_a_loaded_model(tmp_path):
    """The engine keeps the last loaded model and skips reloading it for the same tokenizer object.  Two tokenizers with
    equal vocabulary and merge counts but different merges, created one after the other (the first one freed, so that
    CPython may hand its address to the second), must each encode with their own merges."""
    import gc

    from adapters import get_tokenizer, run_train_bpe

    texts = []
    models = []
    for seed in (101, 102):
        path = synth_corpus(1, seed, 400_000, str(tmp_path / f"c{seed}.txt"))
        models.append(run_train_bpe(path, 700, [EOT]))
        texts.append(open(path, encoding="utf-8").read()[:100_000])
    assert models[0][1] != models[1][1] and len(models[0][1]) == len(models[1][1])
    probe = texts[0][:50_000] + texts[1][:50_000]

    def encode_with(k):
        tok = get_tokenizer(*models[k], [EOT])
        return tok.encode(probe)

    for _ in range(3):
        got = []
        for k in (0, 1):
            got.append(encode_with(k))
            gc.collect()
        for k in (0, 1):
            assert got[k] == port.PortTokenizer(*models[k], [EOT]).encode(probe)


def test_token_byte_pool_grows_when_tokens_get_long(tmp_file, monkeypatch):
    """Runs of one character double the length of the newest token with every merge; the trainer's pool of token
    byte strings is then too small, the loop stops and training runs again with a larger pool: same merges as the oracle."""
    from adapters import run_train_bpe

    data = (b"a" * 3000 + b" " + b"ab" * 700 + b"\n") * 3 + b"the end of it"
    want = c_oracle.train(c_oracle.count(data, [EOT]), 290, [EOT])
    assert max(len(v) for v in want[0].values()) > 512
    monkeypatch.setenv("BPE_B200_TRAIN_BLOB_PER_TOKEN", "4")
    assert run_train_bpe(tmp_file(data), 290, [EOT]) == want


def test_unicode_tables_match_the_regex_on_this_box():
    """The parity target is the `regex` module installed where the CPU baseline runs: the class
    tables compiled into the library must be the ones that module gives, on the GPU box too."""
    from test_oracle_pinned import test_unicode_tables_match_installed_regex

    test_unicode_tables_match_installed_regex()


def test_encode_keyerror_is_raised_every_time(gpt2_model):
    """A pretoken that cannot be encoded is not cached as if it had been: the reference raises on every call."""
    from adapters import get_tokenizer

    vocab, merges = gpt2_model
    small = {k: v for k, v in vocab.items() if v != b"z"}
    tok = get_tokenizer(small, merges, None)
    fine = tok.encode("a plain sentence")
    for k in range(3):
        with pytest.raises(KeyError):
            tok.encode("a plain zebra sentence")
        assert tok.encode("a plain sentence") == fine
        # pretokens the cache has not seen, after a failed call: the failed one must not block or corrupt them
        fresh = f"brand new words number{k} arrive quietly"
        assert tok.encode(fresh) == c_oracle.COracleTokenizer(small, merges, None).encode(fresh)
        with pytest.raises(KeyError):
            tok.encode(fresh + " zoo")


def test_pipelined_file_paths_match_whole_file_paths(tmp_path, ts10k_model, monkeypatch):
    """Large files are read, processed and stored piece by piece on three host threads; with a small piece size the
    same code must give what the whole-file path gives: pretoken counts, trained model, ids, '\\r' handling."""
    from adapters import get_tokenizer, run_train_bpe
    from bpe_transformer_b200.tokenization import bpe_tokenizer as tk
    from bpe_transformer_b200.tokenization.preprocessing import pretokenization as pt

    path = synth_corpus(0, 71, 6_000_000, str(tmp_path / "c.txt"))
    whole_model = run_train_bpe(path, 1500, [EOT])
    whole_counts = pt.parallel_pretokenization(path, special_tokens=[EOT])
    vocab, merges = ts10k_model
    tok = get_tokenizer(with_specials(vocab, [EOT]), merges, [EOT])
    whole_ids = tok.encode_file_to_array(path)
    monkeypatch.setattr(pt, "_PIPE_BYTES", 1 << 20)
    monkeypatch.setattr(tk, "_PIPE_BYTES", 1 << 20)
    assert pt.parallel_pretokenization(path, special_tokens=[EOT]) == whole_counts
    assert run_train_bpe(path, 1500, [EOT]) == whole_model
    ids = tok.encode_file_to_array(path)
    assert ids.dtype == whole_ids.dtype and np.array_equal(ids, whole_ids)
    # into a caller's array: page-locked (ids arrive by DMA) and ordinary, piece by piece and whole-file
    from bpe_transformer_b200 import _lib
    for buf in (_lib.pinned_array(len(whole_ids) + 10, np.uint16), np.zeros(len(whole_ids), dtype=np.uint16)):
        for pipe in (1 << 20, 1 << 40):
            monkeypatch.setattr(tk, "_PIPE_BYTES", pipe)
            buf[:] = 0
            got = tok.encode_file_to_array(path, out=buf)
            assert np.shares_memory(got, buf) and np.array_equal(got, whole_ids)
    monkeypatch.setattr(tk, "_PIPE_BYTES", 1 << 20)
    with pytest.raises(ValueError):
        tok.encode_file_to_array(path, out=np.zeros(len(whole_ids) - 1, dtype=np.uint16))
    with pytest.raises(ValueError):
        tok.encode_file_to_array(path, out=np.zeros(len(whole_ids), dtype=np.uint32))
    size = os.path.getsize(path)
    cuts = tk._cuts_after_newlines(path, 0, size, size // 3)
    parts = [tok.encode_file_to_array(path, lo=a, hi=b) for a, b in zip(cuts[:-1], cuts[1:])]
    assert np.array_equal(np.concatenate(parts), whole_ids)
    # a carriage return anywhere sends the file through the newline-translating path
    data = open(path, "rb").read()
    cr = tmp_path / "cr.txt"
    cr.write_bytes(data[:3_000_000] + b"one\r\ntwo\rthree\n" + data[3_000_000:])
    with open(cr, encoding="utf-8") as fh:
        want = list(tok.encode_iterable(fh))
    assert tok.encode_file_to_array(str(cr)).tolist() == want
    # invalid UTF-8 in a late piece: counted like the reference's errors="ignore" decoding
    bad = tmp_path / "bad.txt"
    bad.write_bytes(data[:4_000_000] + b" caf\xc3 \xff\xfe ok " + data[4_000_000:])
    got = pt.parallel_pretokenization(str(bad), special_tokens=[EOT])
    monkeypatch.setattr(pt, "_PIPE_BYTES", 1 << 40)
    assert got == pt.parallel_pretokenization(str(bad), special_tokens=[EOT])


def test_tables_grow_when_a_corpus_outgrows_them(engine, tmp_path, gpt2_model, monkeypatch):
    """A context configured far too small for the corpus: counting, training and encoding notice (capacity error
    from the kernels, quickly — not after walking a full table), enlarge the tables and give the usual results."""
    import time

    from adapters import get_tokenizer, run_train_bpe
    from bpe_transformer_b200.tokenization.preprocessing import pretokenization as pt

    path = synth_corpus(1, 91, 6_000_000, str(tmp_path / "c.txt"))
    data = open(path, "rb").read()
    want = c_oracle.count(data, [EOT])
    assert len(want) > 50_000
    before = (engine.table_log2, engine.pool_bytes)
    try:
        engine.configure(13, 1 << 16)  # 8192 slots, 64 KiB of pretoken bytes
        t0 = time.perf_counter()
        got = pt.parallel_pretokenization(path, special_tokens=[EOT])
        assert time.perf_counter() - t0 < 20
        assert got == Counter({tuple(k): v for k, v in want.items()})
        assert engine.table_log2 > 13
        engine.configure(13, 1 << 16)
        monkeypatch.setattr(pt, "_PIPE_BYTES", 1 << 20)  # the piece-by-piece path of large files grows the same way
        assert pt.parallel_pretokenization(path, special_tokens=[EOT]) == got
        assert engine.table_log2 > 13
        monkeypatch.setattr(pt, "_PIPE_BYTES", 1 << 40)
        engine.configure(13, 1 << 16)
        vocab, merges = run_train_bpe(path, 600, [EOT])
        assert (vocab, merges) == c_oracle.train(want, 600, [EOT])
        engine.configure(13, 1 << 16)
        gv, gm = gpt2_model
        gv = with_specials(gv, [EOT])
        tok = get_tokenizer(gv, gm, [EOT])
        text = data[:2_000_000].decode("utf-8", errors="ignore")
        assert tok.encode(text) == c_oracle.COracleTokenizer(gv, gm, [EOT]).encode(text)
    finally:
        engine.configure(*before)


def test_model_and_dataset_formats(tmp_path, gpt2_model, input_bytes, tmp_file):
    """GPT-2 text formats round trip; a tokenized file lands in a .npy that holds encode_iterable's ids."""
    from adapters import get_tokenizer
    from bpe_transformer_b200 import formats

    vocab, merges = gpt2_model
    formats.save_gpt2_files(vocab, merges, tmp_path / "vocab.json", tmp_path / "merges.txt")
    v2, m2 = formats.load_gpt2_files(tmp_path / "vocab.json", tmp_path / "merges.txt", [EOT])
    assert m2 == merges and v2 == with_specials(vocab, [EOT])
    tok = get_tokenizer(v2, m2, [EOT])
    path = tmp_file(input_bytes("tinystories_sample.txt"))
    ids = formats.encode_file_to_npy(tok, path, tmp_path / "ids.npy")
    with open(path) as fh:
        want = list(tok.encode_iterable(fh))
    back = formats.load_token_dataset(tmp_path / "ids.npy")
    assert back.dtype == np.uint16 and back.tolist() == want == ids.tolist()


def test_large_roundtrip_and_conservation(tmp_path, ts10k_model):
    """Size-independent properties on a larger corpus (256 MB): occurrences == start bits outside specials,
    decode(encode(x)) == x through the GPU decoder, ids below the vocab size."""
    from adapters import get_tokenizer
    from bpe_transformer_b200 import _lib

    path = synth_corpus(0, 71, 256_000_000, str(tmp_path / "big.txt"))
    data = open(path, "rb").read()
    eng = _lib.get_engine()
    eng.set_special_tokens(_lib.sort_specials([EOT]))
    eng.unique_reset()
    d_text, n = eng.to_device(data)
    mask = eng.scan(d_text, n)
    eng.count(d_text, n, mask)
    n_unique, n_bytes, n_tokens = eng.unique_stats()
    bits = int(np.unpackbits(mask.cpu().numpy().view(np.uint8)).sum())
    assert n_tokens == bits - data.count(EOT.encode())
    vocab, merges = ts10k_model
    tok = get_tokenizer(with_specials(vocab, [EOT]), merges, [EOT])
    ids = tok.encode_file_to_array(path)
    assert int(ids.max()) < len(tok.vocab)
    assert tok.decode_array(ids) == data
    # the first 2 MB against the C oracle (cut at a line end)
    cut = data.rindex(b"\n", 0, 2_000_000) + 1
    want = c_oracle.COracleTokenizer(tok.vocab, merges, [EOT]).encode_lines(data[:cut].decode())
    assert ids[:len(want)].tolist() == want
