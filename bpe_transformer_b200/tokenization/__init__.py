"""Reference: bpe_transformer/tokenization/__init__.py:1-4 (BPETrainer, Tokenizer); BPETokenizer and train_bpe
are exported too because the reference's README imports them from here."""
from .bpe_tokenizer import BPETokenizer
from .bpe_trainer import BPETrainer
from .tokenizer import Tokenizer


def train_bpe(*args, **kwargs):
    from ..main import train_bpe as _train_bpe

    return _train_bpe(*args, **kwargs)


__all__ = ["Tokenizer", "BPETrainer", "BPETokenizer", "train_bpe"]
