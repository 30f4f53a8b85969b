"""Host-side helpers of the product that need no GPU: offsets at which files are cut for the I/O pipelines and for
sharding, the reference's own chunk rule, the host split on special tokens."""
from __future__ import annotations

import io
import random

import regex

from bpe_transformer_b200.distributed import line_shard_offsets, shard_offsets
from bpe_transformer_b200.tokenization.bpe_tokenizer import _cuts_after_newlines
from bpe_transformer_b200.tokenization.preprocessing.pretokenization import find_chunk_boundaries, split_on_special_tokens
from conftest import EOT


def _random_lines(rng: random.Random, n: int, newline_free: bool = False) -> bytes:
    data = bytes(rng.choice(b"ab \n") if rng.random() < 0.9 else 0x78 for _ in range(n))
    return data.replace(b"\n", b"x") if newline_free else data


def test_pieces_of_the_encode_pipeline_end_after_newlines(tmp_path):
    rng = random.Random(1)
    for trial in range(120):
        data = _random_lines(rng, rng.randint(0, 3000), newline_free=trial % 5 == 0)
        path = tmp_path / "f.txt"
        path.write_bytes(data)
        lo = rng.randint(0, len(data) // 3) if data else 0
        offs = _cuts_after_newlines(path, lo, len(data), rng.choice([1, 7, 64, 500, 5000]))
        assert offs[0] == lo and offs[-1] == len(data)
        assert all(a < b for a, b in zip(offs[:-1], offs[1:])) or lo == len(data)
        assert all(data[o - 1:o] == b"\n" for o in offs[1:-1])  # every inner cut sits just behind a newline
        assert b"".join(data[a:b] for a, b in zip(offs[:-1], offs[1:])) == data[lo:]


def test_line_shards_cover_the_file_in_whole_lines(tmp_path):
    rng = random.Random(2)
    for trial in range(60):
        data = _random_lines(rng, rng.randint(1, 4000), newline_free=trial % 7 == 0)
        path = tmp_path / "f.txt"
        path.write_bytes(data)
        world = rng.choice([1, 2, 3, 8])
        offs = line_shard_offsets(path, world)
        assert len(offs) == world + 1 and offs[0] == 0 and offs[-1] == len(data)
        assert all(a <= b for a, b in zip(offs[:-1], offs[1:]))
        assert all(o in (0, len(data)) or data[o - 1:o] == b"\n" for o in offs)
        lines = [ln for a, b in zip(offs[:-1], offs[1:]) for ln in data[a:b].splitlines(keepends=True)]
        assert lines == data.splitlines(keepends=True)


def test_chunk_boundaries_follow_the_reference_rule(tmp_path):
    """pretokenization.py:114-168: guesses k * (size // n) moved forward in 4 KiB probes to the START of the next special
    token (or newline when there are none); EOF when nothing is found; duplicates collapse."""
    rng = random.Random(3)
    words = [b"one ", b"two ", b"three\n", EOT.encode(), b"<|other|>"]
    for trial in range(40):
        data = b"".join(rng.choice(words) for _ in range(rng.randint(1, 6000)))
        specials = [EOT] if trial % 3 else [EOT, "<|other|>"]
        if trial % 4 == 0:
            specials = None
        n = rng.choice([1, 2, 4, 16])
        got = find_chunk_boundaries(io.BytesIO(data), n, specials)
        needles = [s.encode() for s in specials] if specials else [b"\n"]
        want = {0, len(data)}
        for k in range(1, n):
            pos, found = k * (len(data) // n), len(data)
            while pos < len(data):
                window = data[pos:pos + 4096]
                hits = [i for i in (window.find(nd) for nd in needles) if i != -1]
                if hits:
                    found = pos + min(hits)
                    break
                pos += 4096
            want.add(found)
        assert got == sorted(want)
        path = tmp_path / "f.txt"
        path.write_bytes(data)
        if specials:
            offs = shard_offsets(path, n, specials)
            assert len(offs) >= n + 1 and offs[0] == 0 and offs[-1] == len(data) and offs == sorted(offs)


def test_host_split_on_special_tokens_matches_regex_split():
    """pretokenization.py:211-235 with the third-party `regex`: longest special first, kept only when not training."""
    rng = random.Random(4)
    specials = [EOT, EOT + EOT, "<|", "a|b"]
    atoms = ["x", " ", "<|", "endoftext", "|>", EOT, "a|b", "|", "a"]
    pattern = "|".join(regex.escape(s) for s in sorted(specials, key=len, reverse=True))
    for _ in range(300):
        text = "".join(rng.choice(atoms) for _ in range(rng.randint(0, 30)))
        assert split_on_special_tokens(text, training=True, special_tokens=specials) == regex.split(pattern, text)
        assert split_on_special_tokens(text, training=False, special_tokens=specials) == regex.split(f"({pattern})", text)
    assert split_on_special_tokens("abc", special_tokens=None) == ["abc"]


def test_shard_pieces_are_cut_at_special_token_starts(tmp_path):
    from bpe_transformer_b200.distributed import _pieces_at_special_starts

    rng = random.Random(5)
    words = [b"one ", b"two ", b"three\n", EOT.encode(), b"<|other|>", b"<", b"<|"]
    for trial in range(60):
        data = b"".join(rng.choice(words) for _ in range(rng.randint(1, 5000)))
        path = tmp_path / "f.txt"
        path.write_bytes(data)
        lo = rng.randint(0, len(data) // 2)
        hi = rng.randint(lo, len(data))
        specials = [EOT] if trial % 2 else [EOT, "<|other|>"]
        offs = _pieces_at_special_starts(str(path), lo, hi, rng.choice([1, 50, 700, 100000]), specials)
        assert offs[0] == lo and offs[-1] == hi and offs == sorted(offs)
        assert len(set(offs[:-1])) == len(offs[:-1])
        needles = [s.encode() for s in specials]
        assert all(any(data[o:o + len(nd)] == nd for nd in needles) for o in offs[1:-1])


def test_packed_table_exchange_layout_round_trips():
    """pack_table / unpack_tables without a process group: two ranks' payloads laid out as the all-gather would."""
    import numpy as np
    import torch

    from bpe_transformer_b200 import distributed as D

    def table(words):
        blob = b"".join(words)
        lens = np.asarray([len(w) for w in words], dtype=np.int32)
        off = np.zeros(len(words), dtype=np.int64)
        off[1:] = np.cumsum(lens[:-1])
        return (torch.from_numpy(np.frombuffer(blob, dtype=np.uint8).copy()), torch.from_numpy(off), torch.from_numpy(lens),
                torch.arange(1, len(words) + 1, dtype=torch.int64))

    tables = [table([b" the", b"a", b" longer word"]), table([b"x" * 37])]
    packed = [D.pack_table(*t) for t in tables]
    sizes = [(n, nb) for _, n, nb in packed]
    stride = (max(D.payload_bytes(*sz) for sz in sizes) + 63) // 64 * 64 + 64
    gathered = torch.zeros(stride * 2, dtype=torch.uint8)
    for r, (payload, _, _) in enumerate(packed):
        gathered[r * stride:r * stride + payload.numel()] = payload
    blob, off, ln, cnt = D.unpack_tables(gathered, stride, sizes)
    raw = blob.numpy().tobytes()
    got = [(raw[o:o + l], c) for o, l, c in zip(off.tolist(), ln.tolist(), cnt.tolist())]
    assert got == [(b" the", 1), (b"a", 2), (b" longer word", 3), (b"x" * 37, 1)]
