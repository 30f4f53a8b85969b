"""Constants of the tokenizer path (reference: bpe_transformer/settings.py:4-10)."""
from pathlib import Path

ENCODING_STD = "utf-8"

# GPT-2 split pattern (Radford et al., 2019).  The CUDA scanner (csrc/scan.cu) implements exactly this pattern
# as a local rule; the string is kept for documentation and for callers that import it.
PAT = r"""'(?:[sdmt]|ll|ve|re)| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)|\s+"""

# The reference writes Path(__file__) / "output" (a path *under a file*, settings.py:10), which cannot be
# created; the directory next to this file is what was meant.
DEFAULT_OUTPUT_DIR = Path(__file__).resolve().parent / "output"
