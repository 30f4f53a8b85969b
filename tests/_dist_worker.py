"""Worker of tests/test_distributed_cpu.py: one rank of a gloo group exercising the host side of
bpe_transformer_b200.distributed (shard offsets, variable-size table all-gather) on CPU tensors."""
from __future__ import annotations

import json
import os
import pathlib
import sys
from collections import Counter

import numpy as np
import torch
import torch.distributed as dist

ROOT = pathlib.Path(__file__).resolve().parent.parent
for p in (str(ROOT), str(ROOT / "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

import c_oracle  # noqa: E402
from bpe_transformer_b200 import distributed as D  # noqa: E402

EOT = "<|endoftext|>"


def table_tensors(counter: Counter):
    """A Counter in the flat exchange layout of bpe_unique_export (blob, off, len, cnt)."""
    keys = list(counter)
    blob = b"".join(bytes(k) for k in keys)
    lens = np.asarray([len(k) for k in keys], dtype=np.int32)
    off = np.zeros(len(keys), dtype=np.int64)
    if len(keys):
        off[1:] = np.cumsum(lens[:-1])
    cnt = np.asarray([counter[k] for k in keys], dtype=np.int64)
    return (torch.from_numpy(np.frombuffer(blob + b"\0" * 64, dtype=np.uint8).copy()), torch.from_numpy(off),
            torch.from_numpy(lens), torch.from_numpy(cnt))


def main():
    path, out = sys.argv[1], sys.argv[2]
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    data = open(path, "rb").read()
    offs = D.shard_offsets(path, world, [EOT])
    lo, hi = offs[rank], offs[rank + 1]
    mine = c_oracle.count(data[lo:hi], [EOT])
    merged: Counter = Counter()
    gathered, stride, sizes = D.all_gather_tables(*table_tensors(mine))
    assert len(sizes) == world and sizes[rank][0] == len(mine) and gathered.numel() == stride * world and stride % 16 == 0
    blob, off, ln, cnt = D.unpack_tables(gathered, stride, sizes)  # what ONE bpe_unique_import call is given
    b = blob.numpy().tobytes()
    for o, l, c in zip(off.tolist(), ln.tolist(), cnt.tolist()):
        merged[tuple(b[o:o + l])] += c
    whole = c_oracle.count(data, [EOT])
    line_offs = D.line_shard_offsets(path, world)
    ok_lines = all(o == 0 or o == len(data) or data[o - 1:o] == b"\n" for o in line_offs) and line_offs == sorted(line_offs)
    res = {"rank": rank, "offs": offs, "equal": merged == whole, "n": len(merged), "line_offs": line_offs, "ok_lines": ok_lines,
           "starts_at_special": all(o in (0, len(data)) or data[o:o + len(EOT)] == EOT.encode() for o in offs)}
    with open(f"{out}.{rank}", "w") as fh:
        json.dump(res, fh)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
