"""Reference: bpe_transformer/tokenization/preprocessing/__init__.py:1-3."""
from .pretokenization import parallel_pretokenization, pretokenize, serial_pretokenization

__all__ = ["parallel_pretokenization", "serial_pretokenization", "pretokenize"]
