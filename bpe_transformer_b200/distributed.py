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

_SHARD_PIPE_MIN = int(os.environ.get("BPE_B200_SHARD_PIPE_MIN", str(96 << 20)))  # shards beyond this stream in pieces while the GPU counts


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


def pack_table(blob, off, ln, cnt):
    """One table of distinct pretokens as ONE flat uint8 tensor: [cnt int64 x n][off int64 x n][len int32 x n][blob],
    so that it crosses the wire in a single collective.  Returns (payload, n_entries, n_blob_bytes)."""
    import torch

    n, nb = int(off.numel()), int(blob.numel())
    parts = [cnt.contiguous().view(torch.uint8), off.contiguous().view(torch.uint8), ln.contiguous().view(torch.uint8), blob.contiguous()]
    return (torch.cat(parts) if n else torch.empty(0, dtype=torch.uint8, device=blob.device)), n, nb


def payload_bytes(n: int, nb: int) -> int:
    return 20 * n + nb


def all_gather_tables(blob, off, ln, cnt, group=None):
    """All-gather of variable-size tables (blob uint8, off int64, len int32, cnt int64): one tiny collective for the
    sizes, one for the packed payloads (padded to the largest).  Returns (gathered, stride, sizes): `gathered` is one
    flat uint8 tensor holding rank r's payload at r * stride, sizes[r] = (entries, blob bytes).
    Device-agnostic: NCCL with CUDA tensors, gloo with CPU tensors."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    dev = blob.device
    payload, n, nb = pack_table(blob, off, ln, cnt)
    all_sizes = torch.zeros(2 * world, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(all_sizes, torch.tensor([n, nb], dtype=torch.int64, device=dev), group=group)
    flat = all_sizes.tolist()  # the one host sync of the exchange: buffer sizes
    sizes = [(flat[2 * r], flat[2 * r + 1]) for r in range(world)]
    stride = (max(payload_bytes(*sz) for sz in sizes) + 63) // 64 * 64 + 64  # 16-byte aligned payloads, readable padding behind each
    mine = torch.zeros(stride, dtype=torch.uint8, device=dev)
    mine[:payload.numel()] = payload
    gathered = torch.empty(stride * world, dtype=torch.uint8, device=dev)
    dist.all_gather_into_tensor(gathered, mine, group=group)
    return gathered, stride, sizes


def unpack_tables(gathered, stride: int, sizes):
    """(blob, off, len, cnt) over ALL ranks' entries: `blob` is the gathered buffer itself, `off` holds absolute
    offsets into it — what one bpe_unique_import call takes."""
    import torch

    offs, lens, cnts = [], [], []
    for r, (n, nb) in enumerate(sizes):
        if not n:
            continue
        base = r * stride
        cnts.append(gathered[base:base + 8 * n].view(torch.int64))
        offs.append(gathered[base + 8 * n:base + 16 * n].view(torch.int64) + (base + 20 * n))
        lens.append(gathered[base + 16 * n:base + 20 * n].view(torch.int32))
    if not offs:
        e64, e32 = torch.empty(0, dtype=torch.int64, device=gathered.device), torch.empty(0, dtype=torch.int32, device=gathered.device)
        return gathered, e64, e32, e64
    return gathered, torch.cat(offs), torch.cat(lens), torch.cat(cnts)


def exchange_unique(engine: "_lib.Engine", group=None) -> None:
    """After this every rank's table holds the counts of the whole corpus: the tables are all-gathered, the context's
    table is emptied (and enlarged first when the union of all ranks may not fit — every rank takes the same decision
    from the same sizes) and all ranks' entries, this rank's included, are imported by ONE kernel launch."""
    blob, off, ln, cnt = engine.unique_export()
    gathered, stride, sizes = all_gather_tables(blob, off, ln, cnt, group)
    total_n, total_b = sum(s[0] for s in sizes), sum(s[1] for s in sizes)  # upper bounds of the union
    log2, pool = engine.table_log2, engine.pool_bytes
    while (1 << log2) * 5 // 8 < total_n and log2 < 28:
        log2 += 2
    while pool < total_b + (1 << 20):
        pool *= 4
    if (log2, pool) != (engine.table_log2, engine.pool_bytes):
        engine.configure(log2, pool)
    else:
        engine.unique_reset()
    engine.unique_import(*unpack_tables(gathered, stride, sizes))


def _pieces_at_special_starts(path, lo: int, hi: int, piece: int, special_tokens: list[str]) -> list[int]:
    """lo = o0 < o1 < ... = hi with every inner offset at the start of a special token (any such cut leaves the
    pretoken counts unchanged: a special token ends a segment anyway), about `piece` bytes apart."""
    needles = [t.encode("utf-8") for t in special_tokens]
    longest = max(len(nd) for nd in needles)
    offs = [lo]
    with open(path, "rb") as fh:
        target = lo + piece
        while target < hi:
            pos, found = target, hi
            while pos < hi:
                fh.seek(pos)
                window = fh.read(min((1 << 16) + longest, hi - pos))
                hits = [i for i in (window.find(nd) for nd in needles) if i != -1]
                if hits:
                    found = pos + min(hits)
                    break
                pos += 1 << 16
            if found >= hi:
                break
            if found > offs[-1]:
                offs.append(found)
            target = max(target + piece, found + 1)
    offs.append(hi)
    return offs


def count_shard(engine: "_lib.Engine", path, lo: int, hi: int, special_tokens: list[str], reset: bool = True):
    """scan + count bytes [lo, hi) of the file on this rank's GPU; large shards stream into HBM in pieces cut at
    special-token starts while the GPU counts the pieces that have landed."""
    from .tokenization.preprocessing import pretokenization as pt

    engine.set_special_tokens(_lib.sort_specials(special_tokens))
    if reset:
        engine.unique_reset()
    if reset and hi - lo > _SHARD_PIPE_MIN:
        piece = min(pt._PIPE_BYTES, max(_SHARD_PIPE_MIN // 3, (hi - lo) // 4))
        offs = _pieces_at_special_starts(path, lo, hi, piece, special_tokens)
        while True:
            done = pt._count_pieces_pipelined(engine, path, offs)
            if done == "counted":
                return
            engine.unique_reset()
            if done == "invalid-utf8":
                break  # whole-shard path below, which repairs the bytes on the host
            if not engine.grow():  # "capacity"
                raise _lib.BpeError(_lib.ERR_CAPACITY, "table of distinct pretokens cannot grow further")
    d_text, n = engine.load_file(path, lo, hi)
    try:
        if not reset:
            engine.scan(d_text, n)  # earlier shards are in the table: find invalid UTF-8 before anything of this shard is counted
        pt.count_with_growth(engine, d_text, n, keep_counted=not reset)
    except _lib.InvalidUtf8:
        with open(path, "rb") as fh:
            fh.seek(lo)
            clean, _ = pt._sanitised(fh.read(hi - lo), [], strict=False)
        if reset:
            engine.unique_reset()  # the one-pass kernel had counted some of the dirty bytes before it reported them
        d_text, n = engine.to_device(clean)
        pt.count_with_growth(engine, d_text, n, keep_counted=not reset)


def merge_tokens_together(trainer, engine: "_lib.Engine", group=None) -> None:
    """The merge loop with its heavy steps shared between the ranks' GPUs (one box, NVLink): every rank holds all
    pretokens already (exchange_unique); exchange buffers are set up, every rank runs the same loop, and in a step with
    a long candidate list each GPU filters only the words it owns and stores the keys of its hits into its peers' memory."""
    import torch.distributed as dist

    world, rank = dist.get_world_size(group), dist.get_rank(group)
    n_unique = engine.unique_stats()[0]
    handle = engine.peer_alloc(2 * n_unique // world + 4096)
    handles: list = [None] * world
    dist.all_gather_object(handles, handle, group=group)
    engine.peer_connect(rank, handles)
    try:
        dist.barrier(group=group)  # every buffer is open everywhere before the first kernel stores into a peer
        trainer._merge_tokens(engine)
    finally:
        dist.barrier(group=group)  # nobody frees a buffer a peer may still be writing to
        engine.peer_disconnect()


def train_bpe_sharded(input_path, vocab_size: int, special_tokens: list[str], group=None, merge_loop: str = "replicated"):
    """train_bpe over the ranks of `group`: same (vocab, merges) on every rank, identical to one GPU's.
    merge_loop: "replicated" — every GPU runs the whole loop on its own; "shared" — the filter of heavy steps is split
    between the GPUs, matches travel over NVLink.  Both are measured in DESIGN.md section 5: replicated was faster in every
    configuration tried (2.1 GB / 10k vocabulary on 2, 4, 8 GPUs; 11 GB / 32k vocabulary on 8), hence the default."""
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
    if len(trainer.special_tokens) > 1:
        # The reference gives the special tokens ids 256.. in the iteration order of a set (bpe_trainer.py:91, :126), which
        # str hashing randomises per process: every rank takes the order of the group's first rank, so all ranks agree.
        order = [list(trainer.special_tokens)]
        dist.broadcast_object_list(order, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        trainer.set_special_token_order(order[0])
    if merge_loop == "shared" and world > 1:
        merge_tokens_together(trainer, engine, group)
    elif merge_loop in ("replicated", "shared"):
        trainer._merge_tokens(engine)
    else:
        raise ValueError('merge_loop must be "replicated" or "shared"')
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
