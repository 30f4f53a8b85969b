"""Pins the oracle (oracle/bpe_ref_port.py and oracle/bpe_oracle.c) to the reference.

Golden vectors in tests/golden/ were produced by the UNMODIFIED reference (oracle/gen_golden.py); they
include the reference's own known-answer fixtures (tests/fixtures/train-bpe-reference-merges.txt,
the tiktoken-differential inputs of tests/test_tokenizer.py) and vocab sizes > 500 where the reference's
pair accounting departs from textbook BPE.  Runs without a GPU.
"""
from __future__ import annotations

import random
from collections import Counter

import pytest
import regex

import bpe_ref_port as port
import c_oracle
from conftest import EOT, expand_row, load_golden, with_specials

PRETOK = load_golden("pretok.json.gz")
TRAIN = load_golden("train.json.gz")
ENCODE = load_golden("encode.json.gz")
A25 = load_golden("pretok_a25.json.gz")


def a25_input(input_bytes, key: str) -> bytes:
    """Inputs of tests/golden/pretok_a25.json.gz ("dirty:" = invalid UTF-8 spliced in, as oracle/gen_golden.py does)."""
    if key.startswith("dirty:"):
        base = input_bytes(key[len("dirty:"):])
        return base[:1000] + b"\xff\xfe" + base[1000:2000] + b"\xe4\xbd" + base[2000:]
    return input_bytes(key)


def _counter(rows):
    return Counter({tuple(bytes.fromhex(h)): c for h, c in rows})


@pytest.mark.parametrize("case", PRETOK, ids=lambda c: f"{c['input']}-{len(c['specials'])}sp")
def test_pretoken_counts_match_reference(case, input_bytes, tmp_file):
    data = input_bytes(case["input"])
    want = _counter(case["counts"])
    got_port = port.count_pretokens(tmp_file(data), case["n_workers"], case["specials"])
    assert got_port == want
    if case["specials"] or case["n_workers"] == 1:
        # the C oracle has no worker chunking; with specials (or one worker) chunk cuts cannot matter
        assert c_oracle.count(data, case["specials"]) == want


@pytest.mark.parametrize("case", [c for c in A25 if c["fn"] == "serial" and c["training"]],
                         ids=lambda c: f"{c['input']}-{len(c['specials'])}sp")
def test_serial_pretokenization_matches_reference(case, input_bytes, tmp_file):
    """four chunks, strict decode (pretokenization.py:255-298)."""
    path = tmp_file(a25_input(input_bytes, case["input"]))
    if case.get("raises"):
        with pytest.raises(UnicodeDecodeError):
            port.count_pretokens_serial(path, case["specials"])
    else:
        assert port.count_pretokens_serial(path, case["specials"]) == _counter(case["counts"])


@pytest.mark.parametrize("case", TRAIN, ids=lambda c: f"{c['input']}-{c['vocab_size']}")
def test_train_matches_reference(case, input_bytes, tmp_file):
    data = input_bytes(case["input"])
    want_merges = [(bytes.fromhex(a), bytes.fromhex(b)) for a, b in case["merges"]]
    want_vocab = {int(k): bytes.fromhex(v) for k, v in case["vocab"].items()}
    path = tmp_file(data)
    if case["n_workers"] is None:
        words = port.count_pretokens(path, 4, case["specials"])
    else:
        words = port.count_pretokens(path, case["n_workers"], case["specials"])
    vocab, merges = port.train_from_counts(words.copy(), case["vocab_size"], case["specials"])
    assert merges == want_merges
    assert vocab == want_vocab
    if len(data) <= 2_500_000:  # the brute-force C trainer is quadratic-ish; keep the CPU suite short
        vocab_c, merges_c = c_oracle.train(words, case["vocab_size"], case["specials"])
        assert merges_c == want_merges
        assert vocab_c == want_vocab


def test_course_golden_c1():
    """The reference's own known-answer vector (tests/test_train_bpe.py:27-63): ordered merges equal,
    vocab key set and value set equal (ids are not pinned by that fixture)."""
    g = load_golden("course_golden.json.gz")
    case = next(c for c in TRAIN if c["input"] == "corpus.en" and c["vocab_size"] == 500)
    assert case["merges"] == g["merges"]
    assert {int(k) for k in case["vocab"]} == {int(k) for k in g["vocab"]}
    assert set(case["vocab"].values()) == set(g["vocab"].values())


def _models(gpt2_model, ts10k_model):
    return {"gpt2": gpt2_model, "ts10k": ts10k_model}


@pytest.mark.parametrize("row", ENCODE, ids=lambda r: f"{r['model']}-{r['input']}")
def test_encode_matches_reference(row, input_bytes, gpt2_model, ts10k_model):
    vocab, merges = _models(gpt2_model, ts10k_model)[row["model"]]
    vocab = with_specials(vocab, row["specials"])
    data = input_bytes(row["input"])
    text = data.decode("utf-8")
    want = expand_row(row, "encode")
    big = len(data) > 50_000
    ct = c_oracle.COracleTokenizer(vocab, merges, row["specials"])
    assert ct.encode(text) == want
    # a file handle in text mode applies universal-newline translation before the tokenizer sees lines
    translated = text.replace("\r\n", "\n").replace("\r", "\n")
    assert ct.encode_lines(translated) == expand_row(row, "iterable_lines")
    if "tiktoken" in row:
        assert expand_row(row, "tiktoken") == want  # the reference's own differential tests
    if not big or row["input"].startswith("synth:ts"):
        pt = port.PortTokenizer(vocab, merges, row["specials"])
        assert pt.encode(text) == want
        assert list(pt.encode_iterable(translated.splitlines(keepends=True))) == expand_row(row, "iterable_lines")
        chunks = [text[i:i + 997] for i in range(0, len(text), 997)]
        assert list(pt.encode_iterable(chunks)) == expand_row(row, "iterable_chunks_997")
        assert pt.decode(want) == text


def test_notebook_known_answer(ts10k_model):
    """notebooks/3_bpe_tokenization_encode_decode.ipynb cells 9-11."""
    vocab, merges = ts10k_model
    want = [1183, 44, 338, 740, 835, 1262, 5686, 622, 799, 983, 46, 2687, 483, 349, 63]
    s = "Hello, I am encoding my text. How are you?"
    assert c_oracle.COracleTokenizer(vocab, merges, [EOT]).encode(s) == want
    assert port.PortTokenizer(vocab, merges, [EOT]).encode(s) == want


ALPHABET = [" ", " ", "  ", "\n", "\t", "'", "'s", "'ll", "'ve", "'re", "'d", "'m", "'t", "a", "b", "s", "l", "v", "e", "r",
            "L", "1", "23", ".", "!", "?", "é", "ü", " ", "　", " ", "你", "好", "٣", "½", "🙃", "x", "\r", "-", "_",
            "\x1c", "\x85", " ", " ", "ǅ", "ʰ", "़", "​", "Ⅷ"]


def test_c_scanner_matches_regex_on_fuzz():
    """The sequential scanner (Appendix A.1) against regex.finditer on the reference's pattern."""
    rnd = random.Random(7)
    pat = regex.compile(port.GPT2_SPLIT)
    for _ in range(20000):
        s = "".join(rnd.choice(ALPHABET) for _ in range(rnd.randint(0, 16)))
        data = s.encode("utf-8")
        want = [m.group().encode("utf-8") for m in pat.finditer(s)]
        if not data:
            continue
        mask = c_oracle.scan(data)
        starts = [i for i in range(len(data)) if (int(mask[i >> 5]) >> (i & 31)) & 1]
        got = [data[a:b] for a, b in zip(starts, starts[1:] + [len(data)])]
        assert got == want, repr(s)


def test_unicode_tables_match_installed_regex():
    """include/bpe_unicode_tables.h was generated from `regex`; check every code point against the module
    installed where the tests run (never unicodedata)."""
    import importlib.util
    import pathlib

    spec = importlib.util.spec_from_file_location("gen_tables", pathlib.Path(__file__).resolve().parent.parent / "tools" / "gen_unicode_tables.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    cls = mod.classes_from_regex()
    n1, stage1, nb, stage2 = mod.build_tables(cls)
    header = mod.HEADER.read_text()
    assert mod.render(cls) == header, "include/bpe_unicode_tables.h is stale for the installed regex"
    for cp in list(range(0, 0x3500)) + list(range(0xA000, 0xAC00, 7)) + list(range(0x10000, 0x35000, 13)) + [0x10FFFF, 0xE0001]:
        assert mod.lookup(cp, n1, stage1, stage2) == cls[cp]


def test_invalid_utf8_is_reported():
    with pytest.raises(c_oracle.InvalidUtf8):
        c_oracle.scan(b"abc \xff def")
    with pytest.raises(c_oracle.InvalidUtf8):
        c_oracle.scan(b"abc \xed\xa0\x80 def")  # surrogate
    with pytest.raises(c_oracle.InvalidUtf8):
        c_oracle.scan(b"abc \xe4\xbd")  # truncated
