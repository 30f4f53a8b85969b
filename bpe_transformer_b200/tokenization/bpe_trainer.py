"""BPETrainer with the reference's interface (bpe_transformer/tokenization/bpe_trainer.py:10-472); the work
is done by csrc/scan.cu, count.cu and train.cu through the C-ABI (include/bpe_b200.h)."""
from __future__ import annotations

import pickle
from pathlib import Path

from .. import _lib
from ..settings import DEFAULT_OUTPUT_DIR
from .preprocessing.pretokenization import _clamp_workers, count_into_engine


class BPETrainer:
    """Greedy byte-pair-encoding trainer.

    >>> trainer = BPETrainer(vocab_size=10000, special_tokens=["<|endoftext|>"])
    >>> trainer.train(input_path=Path("data.txt"), n_workers=4)
    >>> trainer.save_trainer(output_dir=Path("./tokenizer"))
    """

    def __init__(self, vocab_size: int, special_tokens: list[str] | None = None):
        if vocab_size < 256:  # bpe_trainer.py:82-83
            raise ValueError("Invalid Vocab size: must be at least 256")
        self.encoding = "utf-8"
        self._vocab_size = vocab_size
        self._merges: list[tuple[bytes, bytes]] = []
        self._special_tokens: set = set(special_tokens) if special_tokens else set()  # a set, as in :91
        self._vocab = self._build_initial_vocab()
        self._initial_vocab_size = len(self._vocab)
        self.last_timing: dict | None = None

    @property
    def vocab(self) -> dict[int, bytes]:
        return self._vocab

    @property
    def merges(self) -> list[tuple[bytes, bytes]]:
        return self._merges

    @property
    def special_tokens(self) -> set[str]:
        return self._special_tokens

    @property
    def vocab_size(self) -> int:
        return self._vocab_size

    def _build_initial_vocab(self) -> dict[int, bytes]:
        """ids 0..255 = the bytes, special tokens from id 256 in the set's iteration order (:115-129)."""
        vocab = {i: bytes([i]) for i in range(256)}
        for idx, token in enumerate(self.special_tokens, start=256):
            vocab[idx] = token.encode(self.encoding)
        return vocab

    def set_special_token_order(self, order: list[str]) -> None:
        """Rebuild the initial vocabulary with the special tokens at ids 256.. in the given order (the reference's own
        order is a set's, i.e. arbitrary per process; ranks that train together must agree on one).  Before training only."""
        if set(order) != self._special_tokens or self._merges:
            raise ValueError("set_special_token_order: not the trainer's special tokens, or training has started")
        vocab = {i: bytes([i]) for i in range(256)}
        for idx, token in enumerate(order, start=256):
            vocab[idx] = token.encode(self.encoding)
        self._vocab = vocab

    def add_new_vocab(self, id: int, new_value: bytes) -> None:
        self._vocab[id] = new_value

    def train(self, input_path: Path, n_workers: int | None = None, engine: "_lib.Engine | None" = None) -> None:
        """Pretokenize `input_path`, build the pair histogram, run the merge loop (:141-167) — on the GPU."""
        engine = engine or _lib.get_engine()
        count_into_engine(engine, input_path, list(self._special_tokens), _clamp_workers(n_workers))
        self._merge_tokens(engine)

    def _merge_tokens(self, engine: "_lib.Engine") -> None:
        """The merge loop over the pretokens counted in `engine` (:210-325)."""
        first_new = len(self._vocab)
        init_tokens = [self._vocab[i] for i in range(first_new)]
        try:
            m_a, m_b = engine.train(self._vocab_size, init_tokens)
        except _lib.BpeError as e:
            if e.code == _lib.ERR_NO_PAIRS:  # :237-240
                raise AttributeError("Warning: self._vocab_pairs_heap was not initialized. No pairs to start merging process.") from e
            raise
        vocab, merges = self._vocab, self._merges  # locals: this loop is ~10^4 iterations of pure bookkeeping
        for k, (a, b) in enumerate(zip(m_a.tolist(), m_b.tolist()), start=first_new):
            left, right = vocab[a], vocab[b]
            merges.append((left, right))  # :310
            vocab[k] = left + right       # add_new_vocab, :314-316
        t = engine.train_timing()
        self.last_timing = {"pair_init_ms": t[0], "merge_loop_ms": t[1], "iterations": t[2], "merges": len(m_a)}

    def save_trainer(self, output_dir: Path = Path(DEFAULT_OUTPUT_DIR / "tokenizer" / "bpe_trainer")) -> None:
        """vocab.pkl and merges.pkl, the reference's pickle format (:447-472)."""
        output_dir = Path(output_dir)
        output_dir.mkdir(parents=True, exist_ok=True)
        with open(output_dir / "vocab.pkl", "wb") as f:
            pickle.dump(self._vocab, f)
        with open(output_dir / "merges.pkl", "wb") as f:
            pickle.dump(self._merges, f)
