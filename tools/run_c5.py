"""Developer tool: BASELINE configs[4] — OpenWebText-shaped synthetic corpus (tools/gen_corpus.c style 1), train_bpe vocab_size=32000,
then encode the whole file; on ONE GPU or, under torchrun, on the N GPUs of one box.

    python tools/run_c5.py [bytes=11000000000] [vocab=32000]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port P tools/run_c5.py [bytes] [vocab]

No CPU oracle finishes at 11 GB (the 128 MB version of this configuration is pinned by tests/golden/c5_owt_sha256.json.gz): what is
checked here is size independent — the N-GPU merges equal the 1-GPU merges (sha256), both designs of the merge loop agree, ids are in
range, GPU decode(encode(x)) == x on every rank.
"""
import ctypes, faulthandler, hashlib, os, pathlib, sys, time
faulthandler.dump_traceback_later(120, repeat=True)  # where the host is, should something take minutes
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
os.environ.setdefault("NCCL_DEBUG", "WARN")
import numpy as np
import torch
import bench
from bpe_transformer_b200 import _lib, train_bpe
from bpe_transformer_b200 import distributed as D
from bpe_transformer_b200.tokenization import BPETokenizer

nbytes = int(sys.argv[1]) if len(sys.argv) > 1 else 11_000_000_000
vocab_size = int(sys.argv[2]) if len(sys.argv) > 2 else 32000
world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))


def say(*a):
    if rank == 0:
        print(*a, flush=True)


def barrier():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()


path = f"/dev/shm/bpe_b200_owt_seed5_{nbytes}.txt"
if rank == 0 and not os.path.exists(path):
    lib = ctypes.CDLL(str(ROOT / "tools" / "libgencorpus.so"))
    lib.gen_corpus.restype = ctypes.c_uint64
    t0 = time.perf_counter()
    n = lib.gen_corpus(ctypes.c_int(1), ctypes.c_uint64(5), ctypes.c_uint64(nbytes), path.encode(), ctypes.c_int(min(32, os.cpu_count() or 1)))
    say(f"corpus {path}: {n} B in {time.perf_counter() - t0:.1f} s")
barrier()
size = os.path.getsize(path)
eng = _lib.get_engine(local)
specials = [bench.EOT]
shas = {}
for design in (["replicated", "shared", "replicated", "shared"] if world > 1 else ["one GPU", "one GPU"]):
    barrier(); t0 = time.perf_counter()
    if world > 1:
        vocab, merges = D.train_bpe_sharded(path, vocab_size, specials, merge_loop=design)
    else:
        vocab, merges = train_bpe(path, vocab_size, specials)
    barrier(); dt = time.perf_counter() - t0
    n_unique, n_bytes, n_tokens = eng.unique_stats()
    sha = bench.merges_sha256(merges)
    shas.setdefault(design, sha)
    assert shas[design] == sha, "two trainings disagree"
    say(f"train [{design}, {world} GPU(s)]: {dt:.3f} s, {len(merges)} merges (sha256 {sha[:16]}), {n_unique} distinct pretokens ({n_bytes} B), "
        f"{n_tokens} pretokens in the table, device timing (pair init ms, merge loop ms, iterations) {eng.train_timing()}, "
        f"GPU memory in use {torch.cuda.memory_allocated(local) / 2**30:.1f} GiB (torch) / {(torch.cuda.mem_get_info(local)[1] - torch.cuda.mem_get_info(local)[0]) / 2**30:.1f} GiB (device)")
assert len(set(shas.values())) == 1, f"designs disagree: {shas}"
assert len(vocab) == vocab_size and len(merges) == vocab_size - 257
if world > 1:  # the same training on one GPU (rank 0; the others wait): must give the same merges
    if rank == 0:
        t0 = time.perf_counter()
        v1, m1 = train_bpe(path, vocab_size, specials)
        sha1 = bench.merges_sha256(m1)
        say(f"train [one GPU, for comparison]: {time.perf_counter() - t0:.3f} s, sha256 {sha1[:16]}, equal to the {world}-GPU merges: {m1 == merges}; device timing {eng.train_timing()}")
        assert m1 == merges
    barrier()
tok = BPETokenizer(vocab, merges, specials)
for rep in range(2):
    barrier(); t0 = time.perf_counter()
    if world > 1:
        ids, first, total = D.encode_file_sharded(tok, path)
    else:
        ids = tok.encode_file_to_array(path, engine=eng); first, total = 0, len(ids)
    barrier(); dt = time.perf_counter() - t0
    say(f"encode {rep} [{world} GPU(s)]: {dt:.3f} s = {size / dt / 1e9:.2f} GB/s end to end, {total} ids ({size / total:.2f} B per id), dtype {ids.dtype}, max id on rank 0 {int(ids.max())}")
# roundtrip (GPU decode) of this rank's ids against this rank's bytes, in pieces of 512M ids
offs = D.line_shard_offsets(path, world) if world > 1 else [0, size]
t0 = time.perf_counter()
pos = offs[rank]
with open(path, "rb") as fh:
    fh.seek(pos)
    for a in range(0, len(ids), 1 << 29):
        back = tok.decode_array(ids[a:a + (1 << 29)], engine=eng)
        want = fh.read(len(back))
        assert back == want, f"roundtrip differs in the piece at id {a} of rank {rank}"
        pos += len(back)
assert pos == offs[rank + 1]
barrier()
say(f"roundtrip ok on every rank ({time.perf_counter() - t0:.1f} s)")
if world > 1:
    dist.destroy_process_group()
