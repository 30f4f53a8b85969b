"""Developer tool (torchrun, N >= 2): train with the replicated and with the shared merge loop on a small corpus, compare."""
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
rank = dist.get_rank()
nbytes = int(sys.argv[1]) if len(sys.argv) > 1 else 64_000_000
vocab = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
style = int(sys.argv[3]) if len(sys.argv) > 3 else 0
path = f"/dev/shm/bpe_shared_check_{style}_{nbytes}.txt"
if rank == 0 and not os.path.exists(path):
    import ctypes
    lib = ctypes.CDLL(str(ROOT / "tools" / "libgencorpus.so")); lib.gen_corpus.restype = ctypes.c_uint64
    lib.gen_corpus(ctypes.c_int(style), ctypes.c_uint64(7), ctypes.c_uint64(nbytes), path.encode(), ctypes.c_int(16))
dist.barrier()
res = {}
for design in ("replicated", "shared", "shared", "replicated"):
    torch.cuda.synchronize(); dist.barrier(); t0 = time.perf_counter()
    v, m = D.train_bpe_sharded(path, vocab, [bench.EOT], merge_loop=design)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    t = _lib.get_engine().train_timing()
    res.setdefault(design, m)
    if rank == 0:
        print(f"{design}: {dt * 1e3:.1f} ms, merge loop {t[1]:.1f} ms, {len(m)} merges, equal to first design: {m == res['replicated']}", flush=True)
dist.barrier()
dist.destroy_process_group()
