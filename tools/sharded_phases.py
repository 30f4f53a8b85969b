"""Developer tool (torchrun, N >= 2): where train_bpe_sharded spends its time, phase by phase (max over ranks)."""
import os, sys, pathlib, time
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
os.environ.setdefault("NCCL_DEBUG", "WARN")
import torch
import torch.distributed as dist
import bench
from bpe_transformer_b200 import _lib
from bpe_transformer_b200 import distributed as D
from bpe_transformer_b200.tokenization.bpe_trainer import BPETrainer

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
rank, world = dist.get_rank(), dist.get_world_size()
nbytes = int(sys.argv[1]) if len(sys.argv) > 1 else 2_100_000_000
path = bench.corpus_path(nbytes)
if rank == 0:
    bench.make_corpus(path, nbytes)
dist.barrier()
eng = _lib.get_engine()
specials = [bench.EOT]


def T():
    torch.cuda.synchronize()
    return time.perf_counter()


for rep in range(5):
    dist.barrier()
    t = [T()]
    offs = D.shard_offsets(path, world, specials); t.append(T())
    D.count_shard(eng, path, offs[rank], offs[rank + 1], specials); t.append(T())
    D.exchange_unique(eng, None); t.append(T())
    tr = BPETrainer(vocab_size=10000, special_tokens=specials); t.append(T())
    tr._merge_tokens(eng); t.append(T())
    d = torch.tensor([b - a for a, b in zip(t[:-1], t[1:])] + [t[-1] - t[0]], device="cuda", dtype=torch.float64)
    dist.all_reduce(d, op=dist.ReduceOp.MAX)
    if rank == 0:
        names = ["shard offsets", "load + count", "exchange", "trainer object", "merge (host + loop)", "total"]
        print(f"rep {rep}: " + ", ".join(f"{n} {1e3 * v:.2f}" for n, v in zip(names, d.tolist())) + f" ms; loop {tr.last_timing['merge_loop_ms']:.2f} ms", flush=True)
dist.destroy_process_group()
