"""Developer tool (torchrun, N >= 2): where the time of the table exchange between GPUs goes."""
import os, sys, pathlib, time
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch, torch.distributed as dist
import bench
from bpe_transformer_b200 import _lib, distributed as D

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
path = bench.corpus_path(2_100_000_000)
if rank == 0:
    bench.make_corpus(path, 2_100_000_000)
dist.barrier()
eng = _lib.get_engine(local)
offs = D.shard_offsets(path, world, [bench.EOT])
eng.set_special_tokens(_lib.sort_specials([bench.EOT]))
d_text, n = eng.load_file(path, offs[rank], offs[rank + 1])
def sync():
    torch.cuda.synchronize()
    return time.perf_counter()
for rep in range(4):
    dist.barrier(); t0 = sync()
    eng.unique_reset(); mask = eng.scan(d_text, n); eng.count(d_text, n, mask)
    t1 = sync()
    blob, off, ln, cnt = eng.unique_export()
    t2 = sync()
    parts = D.all_gather_tables(blob, off, ln, cnt)
    t3 = sync()
    for r, (b, o, l, c) in enumerate(parts):
        if r != rank and o.numel():
            eng.unique_import(b.contiguous(), o.contiguous(), l.contiguous(), c.contiguous())
    t4 = sync()
    if rank == 0:
        print(f"rep {rep}: scan+count {1e3*(t1-t0):.2f} ms, export {1e3*(t2-t1):.2f}, all_gather {1e3*(t3-t2):.2f}, import {1e3*(t4-t3):.2f}; "
              f"{off.numel()} local entries, {blob.numel()} B", flush=True)
dist.destroy_process_group()
