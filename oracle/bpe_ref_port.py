"""CPU restatement ("port") of the reference's BPE hot path.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module, and only as the checker or the timed CPU baseline.  Nothing under bpe_transformer_b200/ imports it.

It restates, step for step and with the same third-party scanner (`regex`, pinned 2025.7.34 in the
reference's uv.lock:1275, 2026.3.32 installed in this image), what the reference (pure Python) does:

  pretokenisation   bpe_transformer/tokenization/preprocessing/pretokenization.py:73-252, settings.py:8
  trainer           bpe_transformer/tokenization/bpe_trainer.py:115-426
  tokenizer         bpe_transformer/tokenization/bpe_tokenizer.py:38-423

Parity pinned: tests/test_oracle_pinned.py checks this module against (a) the reference's own golden
vectors (train-bpe-reference-merges.txt / -vocab.json, the tiktoken-differential fixtures) and (b) outputs
of the reference itself imported from /root/reference (tests/golden/*.json.gz, made by
oracle/gen_golden.py), including vocab sizes > 500 where the reference departs from textbook BPE.

It keeps the reference's data structures (Counter keyed by tuples, heapq with lazy invalidation, Pool)
so that timing it says something about the reference's CPU cost; it is written as free functions over a
small state object rather than as the reference's classes.
"""
from __future__ import annotations

import heapq
import os
from collections import Counter
from multiprocessing import Pool, cpu_count

import regex

# settings.py:8 — the GPT-2 split pattern the reference hands to regex.finditer
GPT2_SPLIT = r"""'(?:[sdmt]|ll|ve|re)| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)|\s+"""
_SPLIT_RE = regex.compile(GPT2_SPLIT)


# --------------------------------------------------------------------------------------------------
# pretokenisation
# --------------------------------------------------------------------------------------------------
def chunk_offsets(fh, want: int, specials) -> list[int]:
    """pretokenization.py:114-168: uniform guesses, each moved forward (4 KiB probes) to the start of
    the next special token (b"\\n" when there are none), EOF if none is found; sorted, de-duplicated."""
    needles = [s.encode("utf-8") for s in specials] if specials else [b"\n"]
    fh.seek(0, os.SEEK_END)
    size = fh.tell()
    step = size // want
    cuts = [k * step for k in range(want + 1)]
    cuts[-1] = size
    for k in range(1, want):
        at = cuts[k]
        fh.seek(at)
        while True:
            probe = fh.read(4096)
            if not probe:
                cuts[k] = size
                break
            hits = [h for h in (probe.find(nd) for nd in needles) if h >= 0]
            if hits:
                cuts[k] = at + min(hits)
                break
            at += 4096
    return sorted(set(cuts))


def split_specials(text: str, specials, keep: bool) -> list[str]:
    """pretokenization.py:211-235: regex.split on the escaped specials, longest first (stable);
    `keep` (encoding) wraps the alternation in a capturing group so the specials stay as parts."""
    if not specials:
        return [text]
    alt = "|".join(regex.escape(s) for s in sorted(specials, key=len, reverse=True))
    return regex.split(f"({alt})" if keep else alt, text)


def split_pretokens(text: str) -> list[bytes]:
    """pretokenization.py:238-252."""
    return [m.group().encode("utf-8") for m in _SPLIT_RE.finditer(text)]


def count_range(path, lo: int, hi: int, specials) -> Counter:
    """pretokenization.py:171-208 (training mode): lossy decode, drop specials, count tuples of bytes."""
    with open(path, "rb") as fh:
        fh.seek(lo)
        text = fh.read(hi - lo).decode("utf-8", errors="ignore")
    tally: Counter = Counter()
    for part in split_specials(text, specials or [], keep=False):
        if part:
            for tok in split_pretokens(part):
                tally[tuple(tok)] += 1
    return tally


def count_pretokens(path, n_workers: int | None, specials) -> Counter:
    """pretokenization.py:73-111: clamp workers, chunk, Pool.starmap, fold the Counters."""
    if not n_workers:
        n_workers = 4
    n_workers = min(n_workers, cpu_count())
    with open(path, "rb") as fh:
        cuts = chunk_offsets(fh, n_workers, specials)
    jobs = [(path, a, b, specials) for a, b in zip(cuts[:-1], cuts[1:])]
    with Pool(processes=n_workers) as pool:
        parts = pool.starmap(count_range, jobs)
    total: Counter = Counter()
    for p in parts:
        total = total + p
    return total


def count_pretokens_serial(path, specials) -> Counter:
    """pretokenization.py:255-298: four chunks, strict decode, one process."""
    total: Counter = Counter()
    with open(path, "rb") as fh:
        cuts = chunk_offsets(fh, 4, specials or [])
        for a, b in zip(cuts[:-1], cuts[1:]):
            fh.seek(a)
            text = fh.read(b - a).decode("utf-8")
            for part in split_specials(text, specials or [], keep=False):
                if part:
                    for tok in split_pretokens(part):
                        total[tuple(tok)] += 1
    return total


# --------------------------------------------------------------------------------------------------
# trainer
# --------------------------------------------------------------------------------------------------
class _Cand:
    """bpe_trainer.py:36-67: heap entry; 'smaller' == higher count, ties → greater (bytes, bytes)."""
    __slots__ = ("n", "pair", "key")

    def __init__(self, n, pair, vocab):
        self.n = n
        self.pair = pair
        self.key = (vocab[pair[0]], vocab[pair[1]])

    def __lt__(self, other):
        return self.n > other.n if self.n != other.n else self.key > other.key


def initial_vocab(specials) -> dict[int, bytes]:
    """bpe_trainer.py:115-129: 256 single bytes, then the specials from id 256 (set iteration order)."""
    vocab = {b: bytes([b]) for b in range(256)}
    for k, s in enumerate(set(specials) if specials else set(), start=256):
        vocab[k] = s.encode("utf-8")
    return vocab


def rewrite_word(word: tuple, pair: tuple, new_id: int):
    """bpe_trainer.py:361-426: left-to-right, non-overlapping replacement.  Returns the new word, the
    positions of new_id in it, and the neighbour pairs 'removed' — read from the ORIGINAL word, so
    back-to-back matches double-count (the quirk SURVEY.md §0.2 describes)."""
    out, where, gone = [], [], Counter()
    hit = False
    i = 1
    n = len(word)
    while i < n:
        if (word[i - 1], word[i]) != pair:
            out.append(word[i - 1])
            i += 1
            continue
        hit = True
        out.append(new_id)
        where.append(len(out) - 1)
        if i - 2 >= 0:
            gone[(word[i - 2], word[i - 1])] += 1
        if i + 1 < n:
            gone[(word[i], word[i + 1])] += 1
        i += 2
    if i == n:
        out.append(word[-1])
    return (tuple(out), where, gone) if hit else None


def neighbours_of(word: tuple, where) -> Counter:
    """bpe_trainer.py:327-359: +1 for the pair left and the pair right of every new-token position."""
    made = Counter()
    for p in where:
        if p >= 1:
            made[(word[p - 1], word[p])] += 1
        if p + 1 < len(word):
            made[(word[p], word[p + 1])] += 1
    return made


def train_from_counts(words: Counter, vocab_size: int, specials):
    """bpe_trainer.py:169-325 on an already-built pretoken Counter (which it consumes)."""
    vocab = initial_vocab(specials)
    merges: list[tuple[bytes, bytes]] = []

    # :169-208 pair histogram, inverted index, heap
    tally: Counter = Counter()
    where_used: dict[tuple, set] = {}
    for w, c in words.items():
        if len(w) == 1:
            continue
        for k in range(len(w) - 1):
            pr = (w[k], w[k + 1])
            tally[pr] += c
            where_used.setdefault(pr, set()).add(w)
    heap = [_Cand(c, pr, vocab) for pr, c in tally.items()]
    heapq.heapify(heap)
    if not heap:  # :237-240
        raise AttributeError("no byte pairs to merge")

    new_id = len(vocab)
    while len(vocab) < vocab_size and heap:  # :242
        top = heapq.heappop(heap)
        pr = top.pair
        if tally[pr] != top.n:  # stale entry (:251)
            continue
        touched = set()
        merged_any = False
        for w in list(where_used.get(pr, ())):
            if w not in words:
                continue
            res = rewrite_word(w, pr, new_id)
            if res is None:
                continue
            merged_any = True
            nw, pos, gone = res
            c = words[w]
            for q, k in neighbours_of(nw, pos).items():
                tally[q] += k * c
                touched.add(q)
            for q, k in gone.items():
                tally[q] -= k * c
                touched.add(q)
            for k in range(len(w) - 1):
                s = where_used.get((w[k], w[k + 1]))
                if s is not None:
                    s.discard(w)
            for k in range(len(nw) - 1):
                where_used.setdefault((nw[k], nw[k + 1]), set()).add(nw)
            words[nw] += c
            del words[w]
        if merged_any:
            merges.append((vocab[pr[0]], vocab[pr[1]]))
            vocab[new_id] = vocab[pr[0]] + vocab[pr[1]]
            new_id += 1
            for q in touched:
                if tally[q] > 0:
                    heapq.heappush(heap, _Cand(tally[q], q, vocab))
    return vocab, merges


def train_bpe(path, vocab_size: int, specials, n_workers: int | None = None):
    """main.py:8-17 + bpe_trainer.py:69-93,141-167."""
    if vocab_size < 255 + len(specials):
        raise ValueError("vocab_size too small")
    if vocab_size < 256:
        raise ValueError("vocab_size must be at least 256")
    words = count_pretokens(path, n_workers or cpu_count(), list(set(specials)) if specials else [])
    return train_from_counts(words, vocab_size, specials)


# --------------------------------------------------------------------------------------------------
# tokenizer
# --------------------------------------------------------------------------------------------------
class PortTokenizer:
    """bpe_tokenizer.py:38-423 restated: merges are looked up by their BYTES pair, ids by bytes."""

    def __init__(self, vocab, merges, specials=None):
        self.vocab = vocab
        self.merges = merges
        self.specials = set(specials) if specials else set()
        self.rank = {m: k for k, m in enumerate(merges)}  # later duplicate wins (:55)
        self.id_of = {v: k for k, v in vocab.items()}  # later id wins (:85)
        self.memo: dict[tuple, tuple] = {}

    def _merge_once(self, ids: list[int]):
        """:257-290: the minimum-rank adjacent pair, leftmost on ties; merge that ONE occurrence."""
        best, at = len(self.merges) + 1, -1
        for k in range(1, len(ids)):
            r = self.rank.get((self.vocab[ids[k - 1]], self.vocab[ids[k]]))
            if r is not None and r < best:
                best, at = r, k - 1
        if at < 0:
            return None
        joined = self.vocab[ids[at]] + self.vocab[ids[at + 1]]
        return ids[:at] + [self.id_of[joined]] + ids[at + 2:]

    def _word(self, tok: bytes) -> tuple:
        start = tuple(self.id_of[bytes([b])] for b in tok)  # :190-207
        hit = self.memo.get(start)
        if hit is not None:
            return hit
        ids = list(start)
        while len(ids) >= 2:  # :240-249
            nxt = self._merge_once(ids)
            if not nxt:
                break
            ids = nxt
        self.memo[start] = tuple(ids)
        return self.memo[start]

    def encode(self, text: str) -> list[int]:
        """:139-188."""
        out: list[int] = []
        sp = list(self.specials)
        for part in (split_specials(text, sp, keep=True) if sp else [text]):
            if part in self.specials:
                out.append(self.id_of[part.encode("utf-8")])
                continue
            for tok in split_pretokens(part):
                out.extend(self._word(tok))
        return out

    def encode_iterable(self, chunks, n_workers=None):
        """:339-423: buffer, cut after the LAST newline of the buffer, encode each cut on its own."""
        if n_workers is None or n_workers <= 1:
            pending = ""
            for ch in chunks:
                pending += ch
                cut = pending.rfind("\n")
                if cut != -1:
                    head, pending = pending[:cut + 1], pending[cut + 1:]
                    yield from self.encode(head)
            if pending:
                yield from self.encode(pending)
            return
        pending, batch = "", []
        with Pool(processes=n_workers) as pool:
            for ch in chunks:
                pending += ch
                cut = pending.rfind("\n")
                if cut != -1:
                    batch.append(pending[:cut + 1])
                    pending = pending[cut + 1:]
                    if len(batch) >= n_workers * 10:
                        for ids in pool.map(self.encode, batch, chunksize=5):
                            yield from ids
                        batch = []
            if batch:
                for ids in pool.map(self.encode, batch, chunksize=5):
                    yield from ids
        if pending:
            yield from self.encode(pending)

    def decode(self, ids) -> str:
        """:119-137."""
        return b"".join(self.vocab.get(i, b"\xef\xbf\xbd") for i in ids).decode("utf-8", errors="replace")
