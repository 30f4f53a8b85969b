"""Pretokenisation entry points under the names the reference exports (its preprocessing/__init__.py): the three
file-level functions.  The helpers (chunk offsets, host split, GPU counting) live in `.pretokenization`."""
from . import pretokenization as _impl

pretokenize = _impl.pretokenize
parallel_pretokenization = _impl.parallel_pretokenization
serial_pretokenization = _impl.serial_pretokenization

__all__ = ["pretokenize", "parallel_pretokenization", "serial_pretokenization"]
