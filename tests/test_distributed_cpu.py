"""The N > 1 host logic on CPU: world_size 2 and 3 over gloo (no GPU).  Shards are cut where the reference cuts
its worker chunks, the per-rank tables are all-gathered as flat tensors, and their sum equals the whole file's
Counter (the reference's fold, pretokenization.py:109)."""
from __future__ import annotations

import json
import os
import pathlib
import socket
import subprocess
import sys

import pytest

from conftest import synth_corpus

ROOT = pathlib.Path(__file__).resolve().parent.parent


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_tables_sum_to_the_whole(world, tmp_path):
    path = synth_corpus(0, 91, 600_000, str(tmp_path / "c.txt"))
    out = str(tmp_path / "res")
    env = dict(os.environ, OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(free_port()), str(ROOT / "tests" / "_dist_worker.py"), path, out]
    subprocess.run(cmd, check=True, env=env, timeout=300, capture_output=True)
    results = [json.load(open(f"{out}.{r}")) for r in range(world)]
    assert all(r["equal"] for r in results)
    assert all(r["offs"] == results[0]["offs"] for r in results)
    assert all(r["starts_at_special"] and r["ok_lines"] for r in results)
    assert len(results[0]["offs"]) == world + 1


def test_shard_offsets_need_a_special_token(tmp_path):
    from bpe_transformer_b200 import distributed as D

    p = tmp_path / "c.txt"
    p.write_bytes(b"a\nb\n")
    with pytest.raises(ValueError):
        D.shard_offsets(str(p), 2, [])
    assert D.line_shard_offsets(str(p), 2) == [0, 2, 4] or D.line_shard_offsets(str(p), 2) == [0, 4, 4]
