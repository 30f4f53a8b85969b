"""bpe_transformer_b200 — the BPE tokenizer hot path of milasd/BPE-Transformer on B200 GPUs.

Same Python surface as the reference package (`bpe_transformer/__init__.py:1-6`): `train_bpe`, plus
`tokenization.{BPETrainer, BPETokenizer, Tokenizer}` and `tokenization.preprocessing.*`.  Underneath, Python
calls hand-written sm_100a CUDA kernels through the C-ABI in include/bpe_b200.h (ctypes).  There is no CPU
fallback: importing is cheap, but the first call needs the built library and a CUDA device.
"""
from .main import train_bpe

__version__ = "0.1.0"

__all__ = ["train_bpe"]
