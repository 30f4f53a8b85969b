"""Developer tool (torchrun, N >= 2): where the time of the table exchange of sharded training goes."""
import os, sys, pathlib, time
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
os.environ.setdefault("NCCL_DEBUG", "WARN")
import torch
import torch.distributed as dist
import bench
from bpe_transformer_b200 import _lib
from bpe_transformer_b200 import distributed as D

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
rank, world = dist.get_rank(), dist.get_world_size()
nbytes = int(sys.argv[1]) if len(sys.argv) > 1 else 2_100_000_000
path = bench.corpus_path(nbytes)
if rank == 0:
    bench.make_corpus(path, nbytes)
dist.barrier()
eng = _lib.get_engine(local)
eng.set_special_tokens(_lib.sort_specials([bench.EOT]))
offs = D.shard_offsets(path, world, [bench.EOT])
d_text, n = eng.load_file(path, offs[rank], offs[rank + 1])

def T():
    torch.cuda.synchronize()
    return time.perf_counter()

for rep in range(4):
    eng.unique_reset()
    eng.pretokenize_count(d_text, n)
    dist.barrier()
    t0 = T()
    blob, off, ln, cnt = eng.unique_export()
    t1 = T()
    gathered, stride, sizes = D.all_gather_tables(blob, off, ln, cnt)
    t2 = T()
    eng.unique_reset()
    t3 = T()
    tabs = D.unpack_tables(gathered, stride, sizes)
    t4 = T()
    eng.unique_import(*tabs)
    t5 = T()
    if rank == 0:
        print(f"rep {rep}: export {1e3 * (t1 - t0):.2f} ms, all-gather {1e3 * (t2 - t1):.2f} (stride {stride} B), reset {1e3 * (t3 - t2):.2f}, "
              f"unpack {1e3 * (t4 - t3):.2f}, import {1e3 * (t5 - t4):.2f}; total {1e3 * (t5 - t0):.2f} ms; distinct {eng.unique_stats()[0]}", flush=True)
dist.destroy_process_group()
