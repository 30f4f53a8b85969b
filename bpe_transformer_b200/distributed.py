"""One process per GPU: shard a corpus at special-token starts, count per shard, exchange the tables of distinct
pretokens, run the merge loop replicated; encode shards with no communication.

The reference's only parallelism is a multiprocessing.Pool over byte ranges cut at special-token starts whose
Counters are summed (pretokenization.py:96-109); the exchange below is that sum across GPUs.  torch.distributed
(NCCL on GPUs, gloo in the CPU tests) is the plumbing; the tables travel as four flat tensors per rank.
"""
from __future__ import annotations

import os

import numpy as np

from . import _lib
from .tokenization.preprocessing.pretokenization import find_chunk_boundaries


def shard_offsets(path, world: int, special_tokens: list[str]) -> list[int]:
    """Byte offsets of the shards: the reference's own chunking rule (pretokenization.py:114-168), so a shard
    always begins at the start of a special token (or at 0) and sharding cannot change the result."""
    if not special_tokens:
        raise ValueError("sharded training needs at least one special token to cut the corpus at")
    with open(path, "rb") as fh:
        offs = find_chunk_boundaries(fh, world, special_tokens)
    size = os.path.getsize(path)
    # fewer distinct offsets than ranks (tiny file): trailing ranks get empty shards
    while len(offs) < world + 1:
        offs.append(size)
    return offs


def line_shard_offsets(path, world: int) -> list[int]:
    """Shards for encoding a file line by line: cut after the first newline at or behind k * size / world."""
    size = os.path.getsize(path)
    offs = [0]
    with open(path, "rb") as fh:
        for k in range(1, world):
            pos = max(offs[-1], k * size // world)
            fh.seek(pos)
            found = size
            while pos < size:
                window = fh.read(1 << 16)
                if not window:
                    break
                i = window.find(b"\n")
                if i != -1:
                    found = pos + i + 1
                    break
                pos += len(window)
            offs.append(found)
    offs.append(size)
    return offs


def all_gather_tables(blob, off, ln, cnt, group=None):
    """All-gather of variable-size tables (blob uint8, off int64, len int32, cnt int64) -> list per rank.
    Device-agnostic: NCCL with CUDA tensors, gloo with CPU tensors."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    dev = blob.device
    sizes = torch.tensor([int(off.numel()), int(blob.numel())], dtype=torch.int64, device=dev)
    all_sizes = [torch.zeros(2, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(all_sizes, sizes, group=group)
    all_sizes = [tuple(int(v) for v in s.tolist()) for s in all_sizes]
    max_n = max(1, max(s[0] for s in all_sizes))
    max_b = max(1, max(s[1] for s in all_sizes))

    def padded(t, n, dtype):
        out = torch.zeros(n, dtype=dtype, device=dev)
        out[:t.numel()] = t
        return out

    gathered = []
    for t, n, dtype in ((blob, max_b, torch.uint8), (off, max_n, torch.int64), (ln, max_n, torch.int32), (cnt, max_n, torch.int64)):
        bufs = [torch.zeros(n, dtype=dtype, device=dev) for _ in range(world)]
        dist.all_gather(bufs, padded(t, n, dtype), group=group)
        gathered.append(bufs)
    out = []
    for r, (n_u, n_b) in enumerate(all_sizes):
        out.append((gathered[0][r][:max(n_b, 0)], gathered[1][r][:n_u], gathered[2][r][:n_u], gathered[3][r][:n_u]))
    return out


def exchange_unique(engine: "_lib.Engine", group=None) -> None:
    """After this every rank's table holds the counts of the whole corpus."""
    import torch.distributed as dist

    rank = dist.get_rank(group)
    blob, off, ln, cnt = engine.unique_export()
    for r, (b, o, l, c) in enumerate(all_gather_tables(blob, off, ln, cnt, group)):
        if r != rank and o.numel():
            engine.unique_import(b.contiguous(), o.contiguous(), l.contiguous(), c.contiguous())


def count_shard(engine: "_lib.Engine", path, lo: int, hi: int, special_tokens: list[str], reset: bool = True):
    """scan + count bytes [lo, hi) of the file on this rank's GPU.  Returns (device text, length) for reuse."""
    import torch

    from .tokenization.preprocessing.pretokenization import _sanitised

    engine.set_special_tokens(_lib.sort_specials(special_tokens))
    if reset:
        engine.unique_reset()
    d_text, n = engine.load_file(path, lo, hi)
    try:
        mask = engine.scan(d_text, n)
    except _lib.InvalidUtf8:
        with open(path, "rb") as fh:
            fh.seek(lo)
            clean, _ = _sanitised(fh.read(hi - lo), [], strict=False)
        d_text, n = engine.to_device(clean)
        mask = engine.scan(d_text, n)
    from .tokenization.preprocessing.pretokenization import count_with_growth

    count_with_growth(engine, d_text, n, mask, keep_counted=not reset)
    return d_text, n


def train_bpe_sharded(input_path, vocab_size: int, special_tokens: list[str], group=None):
    """train_bpe over the ranks of `group`: same (vocab, merges) on every rank, identical to one GPU's."""
    import torch.distributed as dist

    from .tokenization.bpe_trainer import BPETrainer

    if vocab_size < 255 + len(special_tokens):
        raise ValueError("Input vocab_size is invalid: value too small.")
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    engine = _lib.get_engine()
    offs = shard_offsets(input_path, world, special_tokens)
    count_shard(engine, input_path, offs[rank], offs[rank + 1], special_tokens)
    exchange_unique(engine, group)
    trainer = BPETrainer(vocab_size=vocab_size, special_tokens=special_tokens)
    trainer._merge_tokens(engine)
    return trainer.vocab, trainer.merges


def encode_file_sharded(tokenizer, path, group=None, out: np.ndarray | None = None) -> tuple[np.ndarray, int, int]:
    """Every rank encodes its share of the file's lines; returns (ids of this rank, offset of its first id in the
    whole file's id sequence, total ids).  Concatenating the ranks' arrays in rank order gives
    list(tokenizer.encode_iterable(open(path))).  `out`: see BPETokenizer.encode_file_to_array."""
    import torch
    import torch.distributed as dist

    world, rank = dist.get_world_size(group), dist.get_rank(group)
    engine = _lib.get_engine()
    offs = line_shard_offsets(path, world)
    lo, hi = offs[rank], offs[rank + 1]
    ids = tokenizer.encode_file_to_array(path, engine=engine, lo=lo, hi=hi, out=out)
    counts = [torch.zeros(1, dtype=torch.int64, device=f"cuda:{engine.device}") for _ in range(world)]
    dist.all_gather(counts, torch.tensor([len(ids)], dtype=torch.int64, device=f"cuda:{engine.device}"), group=group)
    counts = [int(c.item()) for c in counts]
    return ids, sum(counts[:rank]), sum(counts)
