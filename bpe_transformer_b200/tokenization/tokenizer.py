"""What every tokenizer of this package offers — the reference's abstract `Tokenizer`
(bpe_transformer/tokenization/tokenizer.py:6-31), so that code written against it accepts the GPU-backed one.

The contract, in the reference's terms:
  decode(ids)                 ids -> text; unknown ids become U+FFFD, bytes are decoded with errors="replace"
  encode(text)                text -> ids; special tokens map to their own id, everything else goes through the
                              GPT-2 split pattern and the merges in rank order
  encode_iterable(iterable)   the same, lazily, over pieces of text (a file handle yields lines)
  merges / vocab              the ordered merges and the id -> bytes table the tokenizer was built from
  from_files(...)             alternative constructor; concrete classes define their own arguments
"""
import abc
from collections.abc import Iterable, Iterator

Merges = list[tuple[bytes, bytes]]
Vocab = dict[int, bytes]


class Tokenizer(abc.ABC):
    """Abstract base: subclasses implement all five members below."""

    @abc.abstractmethod
    def decode(self, ids: list[int]) -> str:
        raise NotImplementedError

    @abc.abstractmethod
    def encode(self, text: str) -> list[int]:
        raise NotImplementedError

    @abc.abstractmethod
    def encode_iterable(self, iterable: Iterable[str]) -> Iterator[int]:
        raise NotImplementedError

    @property
    @abc.abstractmethod
    def merges(self) -> Merges:
        raise NotImplementedError

    @property
    @abc.abstractmethod
    def vocab(self) -> Vocab:
        raise NotImplementedError

    @classmethod
    def from_files(cls) -> None:
        """Placeholder, as in the reference: the base class loads nothing."""
        return None
