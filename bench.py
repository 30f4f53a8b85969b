#!/usr/bin/env python3
"""bench.py — BASELINE.json's metric on B200: train_bpe seconds (10k vocab, 2.1 GB synthetic TinyStories-shaped
corpus) and encode GB/s, next to the reference's CPU path timed on the same box.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--bytes B] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

One JSON line on stdout (rank 0).  A "step" is one full pass of the training hot path over the corpus:
split + hash + count of the pretokens (one kernel) -> (N > 1: all-gather of the tables) -> pair histogram -> merge loop.
  value      seconds per step with the corpus already resident in HBM (CUDA events, max over ranks)
  e2e        seconds per call of the public API train_bpe(path, 10000, ["<|endoftext|>"]) — file -> pinned host
             memory -> HBM -> kernels -> merges back on the host, all inside the timed region
  encode     the second half of the metric: GB/s of corpus through BPETokenizer (resident and end to end)
  roofline   the slower of the two HBM-bound kernels of a step (scan / count): algorithmic bytes = corpus bytes
  cpu_baseline  oracle/bpe_ref_port.py (the reference's algorithm, regex + Counter + heap, all host cores) over the
             WHOLE corpus, one measured run; its pretoken Counter, vocab and merges are compared with the GPU's
  parity     the gate: merges / vocab / Counter equal the CPU run's, ids equal the C oracle's over the whole corpus and
             the port's over a 64 MB prefix; sha256 of merges and ids at every N.  Any mismatch: exit code 3
The corpus (2.1 GB) is far larger than L2 (126 MB), so no L2 flush is needed between iterations.
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import pathlib
import subprocess
import sys
import threading
import time

ROOT = pathlib.Path(__file__).resolve().parent
for p in (str(ROOT), str(ROOT / "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

EOT = "<|endoftext|>"
VOCAB_SIZE = 10_000
METRIC = "train_bpe_seconds_10k_vocab_2.1GB"


def log(*a):
    if int(os.environ.get("RANK", "0")) == 0:  # one copy of every progress line under torchrun
        print(*a, file=sys.stderr, flush=True)


def corpus_path(nbytes: int, seed: int = 0) -> str:
    base = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else "/tmp"
    return f"{base}/bpe_b200_ts_seed{seed}_{nbytes}.txt"


def make_corpus(path: str, nbytes: int, seed: int = 0) -> None:
    if os.path.exists(path) and os.path.getsize(path) >= nbytes:
        return
    so = ROOT / "tools" / "libgencorpus.so"
    if not so.exists():
        subprocess.check_call(["gcc", "-O2", "-shared", "-fPIC", "-pthread", "-o", str(so), str(ROOT / "tools" / "gen_corpus.c")])
    lib = ctypes.CDLL(str(so))
    lib.gen_corpus.restype = ctypes.c_uint64
    tmp = path + ".tmp"
    n = lib.gen_corpus(ctypes.c_int(0), ctypes.c_uint64(seed), ctypes.c_uint64(nbytes), tmp.encode(), ctypes.c_int(min(32, os.cpu_count() or 1)))
    if not n:
        raise RuntimeError("corpus generation failed")
    os.replace(tmp, path)


def cut_sample(path: str, nbytes: int, out: str) -> int:
    """First ~nbytes of the corpus, cut at a document end."""
    with open(path, "rb") as fh:
        data = fh.read(nbytes + 65536)
    end = data.find(EOT.encode(), nbytes)
    end = len(data) if end < 0 else end + len(EOT) + 1
    with open(out, "wb") as fh:
        fh.write(data[:end])
    return end


# ---------------------------------------------------------------------------------------------------------
# CPU side: the reference's algorithm (port), all host cores
# ---------------------------------------------------------------------------------------------------------
def merges_sha256(merges) -> str:
    import hashlib

    h = hashlib.sha256()
    for a, b in merges:
        h.update(a.hex().encode() + b" " + b.hex().encode() + b"\n")
    return h.hexdigest()


def cpu_train_full(path: str) -> dict:
    """The reference's algorithm (oracle/bpe_ref_port.py: regex + Counter + Pool + heapq) over the WHOLE corpus, all host
    cores, timed once.  Returns the measured seconds and the results (Counter of pretokens, vocab, merges) for parity."""
    import bpe_ref_port as port

    cores = os.cpu_count() or 1
    total = os.path.getsize(path)
    t0 = time.perf_counter()
    words = port.count_pretokens(path, cores, [EOT])
    t_pre = time.perf_counter() - t0
    counter = words.copy()  # train_from_counts consumes its argument
    t0 = time.perf_counter()
    vocab, merges = port.train_from_counts(words, VOCAB_SIZE, [EOT])
    t_merge = time.perf_counter() - t0
    return {"value": t_pre + t_merge, "unit": "s", "cores": cores, "kind": "port",
            "sample": f"full corpus, 1 measured run: {total} B, pretokenize+count on {cores} processes {t_pre:.2f} s, "
                      f"pair histogram + merge loop to {len(merges)} merges {t_merge:.2f} s",
            "pretokenize_s": t_pre, "merge_s": t_merge, "_counter": counter, "_vocab": vocab, "_merges": merges}


def cpu_train_sample(path: str, total_bytes: int, sample_bytes: int) -> dict:
    """Dispersion only (never the reported value): pretokenize+count of a prefix, scaled to the corpus."""
    import bpe_ref_port as port

    cores = os.cpu_count() or 1
    sample = path + f".sample{sample_bytes}"
    n_sample = cut_sample(path, sample_bytes, sample)
    t0 = time.perf_counter()
    port.count_pretokens(sample, cores, [EOT])
    t_pre = time.perf_counter() - t0
    os.unlink(sample)
    return {"pretokenize_s_scaled": t_pre * total_bytes / n_sample, "sample_bytes": n_sample}


def cpu_encode_sample(path: str, vocab, merges, serial_bytes: int, parallel_bytes: int) -> dict:
    """BPETokenizer.encode_iterable of the reference (port) over prefixes of the corpus: the serial variant and the
    n_workers=cpu_count() variant (bpe_tokenizer.py:381-423), the faster one is the baseline.  Returns the serial ids too."""
    import numpy as np

    import bpe_ref_port as port

    cores = os.cpu_count() or 1
    out = {}
    ids_serial = None
    for name, nbytes, workers in (("serial", serial_bytes, None), ("parallel", parallel_bytes, cores)):
        sample = path + f".encsample{nbytes}"
        n_sample = cut_sample(path, nbytes, sample)
        tok = port.PortTokenizer(vocab, merges, [EOT])
        t0 = time.perf_counter()
        with open(sample, encoding="utf-8") as fh:
            ids = np.fromiter(tok.encode_iterable(fh, n_workers=workers), dtype=np.uint32)
        dt = time.perf_counter() - t0
        os.unlink(sample)
        out[name] = {"GBps": n_sample / dt / 1e9, "sample_bytes": n_sample, "seconds": dt, "ids": int(ids.size), "n_workers": workers or 1}
        if name == "serial":
            ids_serial = ids
    best = max(("serial", "parallel"), key=lambda k: out[k]["GBps"])
    return {"value": out[best]["GBps"], "unit": "GB/s", "cores": out[best]["n_workers"], "kind": "port", "variant": best,
            "sample": f"encode_iterable over a prefix cut at a document end: serial {out['serial']['sample_bytes']} B in {out['serial']['seconds']:.2f} s, "
                      f"n_workers={cores} {out['parallel']['sample_bytes']} B in {out['parallel']['seconds']:.2f} s; the faster ({best}) is the baseline",
            "serial": out["serial"], "parallel": out["parallel"], "_ids_serial": ids_serial, "_serial_bytes": out["serial"]["sample_bytes"]}


def c_oracle_ids_sha256(path: str, vocab, merges) -> tuple[str, int]:
    """sha256 of the uint16 ids the C oracle (oracle/bpe_oracle.c: sequential scanner + literal one-merge-at-a-time
    encoder) gives for encode_iterable(open(path)) over the whole file.  Lines are independent segments there, so the
    file is cut after newlines into ranges that host threads encode side by side (ctypes releases the GIL)."""
    import hashlib
    from concurrent.futures import ThreadPoolExecutor

    import numpy as np

    import c_oracle

    data = np.fromfile(path, dtype=np.uint8)
    n = int(data.size)
    cores = os.cpu_count() or 1
    n_ranges = max(1, min(cores * 4, n // (4 << 20)))
    cuts = [0]
    for k in range(1, n_ranges):
        at = max(cuts[-1], k * n // n_ranges)
        nl = np.flatnonzero(data[at:at + (1 << 20)] == 10)
        cuts.append(at + int(nl[0]) + 1 if nl.size else cuts[-1])
    cuts.append(n)
    ct = c_oracle.COracleTokenizer(vocab, merges, [EOT])

    def work(k):
        lo, hi = cuts[k], cuts[k + 1]
        if hi <= lo:
            return np.empty(0, dtype=np.uint16)
        return ct.encode_array(data[lo:hi], flags=c_oracle.CUT_AFTER_NEWLINE).astype(np.uint16)

    h = hashlib.sha256()
    total = 0
    with ThreadPoolExecutor(max_workers=cores) as pool:
        for part in pool.map(work, range(len(cuts) - 1)):
            h.update(part.tobytes())
            total += int(part.size)
    return h.hexdigest(), total


def run_reference(args, rank: int, world: int) -> None:
    """--impl reference: the reference's own CPU implementation of the path (the port: the reference is pure
    Python and /root/reference does not travel), all host cores, same metric / unit / config.  The value is ONE
    measured run over the whole corpus (~1 min on 16-32 cores), whatever --steps says: repeating it K times would take
    the run far beyond "a few minutes".  Further steps time the pretokenisation of a 64 MB prefix, for dispersion only."""
    if rank != 0:
        return
    path = corpus_path(args.bytes)
    make_corpus(path, args.bytes)
    total = os.path.getsize(path)
    full = cpu_train_full(path)
    log(f"[reference] full corpus: {full['value']:.1f} s ({full['pretokenize_s']:.1f} + {full['merge_s']:.1f})")
    extra = []
    for _ in range(min(2, max(0, args.steps - 1))):
        extra.append(cpu_train_sample(path, total, min(total, args.cpu_sample_bytes)))
    value = full["value"]
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": value * 1e3, "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(total, world),
        "cpu_baseline": {"value": value, "unit": "s", "cores": full["cores"], "kind": "port", "sample": full["sample"]},
        "e2e": {"value": value, "unit": "s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "timed_runs": 1, "merges": len(full["_merges"]), "merges_sha256": merges_sha256(full["_merges"]),
        "dispersion_pretokenize_64MB_scaled_s": [round(e["pretokenize_s_scaled"], 2) for e in extra],
    }
    if not args.no_cpu_encode:
        enc = cpu_encode_sample(path, full["_vocab"], full["_merges"], min(total, 16_000_000), min(total, 2_000_000))
        line["encode"] = {"metric": "encode_GBps_corpus_to_uint16", "value": enc["value"], "unit": "GB/s", "variant": enc["variant"], "cores": enc["cores"],
                          "sample": enc["sample"]}
    emit_line(line)


def workload_config(total: int, world: int) -> dict:
    return {"workload": f"BASELINE configs[1]{'/[2]' if world > 1 else ''}: train_bpe on synthetic TinyStories-shaped corpus "
                        f"(tools/gen_corpus.c style 0 seed 0, {total} B), vocab_size=10000, special_tokens=['<|endoftext|>']; "
                        f"encode = configs[3] over the same corpus to uint16",
            "corpus_bytes": total, "vocab_size": VOCAB_SIZE, "sharding": f"{world} shard(s) cut at <|endoftext|> starts",
            "l2": "inputs (2.1 GB) larger than L2 (126 MB); no flush between iterations"}


# ---------------------------------------------------------------------------------------------------------
# clocks
# ---------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        self.index = index
        self.rows: list[list[str]] = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(self.index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
        sm, mx, reasons = [], 0, set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx = max(mx, float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, cell in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if cell.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------
# GPU side
# ---------------------------------------------------------------------------------------------------------
def run_gpu(args, rank: int, world: int, local_rank: int) -> None:
    import numpy as np
    import torch
    import torch.distributed as dist

    import __graft_entry__ as entry

    if rank == 0:
        entry.build()
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.barrier()

    from bpe_transformer_b200 import _lib, train_bpe
    from bpe_transformer_b200 import distributed as D
    from bpe_transformer_b200.tokenization import BPETokenizer, BPETrainer

    path = corpus_path(args.bytes)
    if rank == 0:
        t0 = time.perf_counter()
        make_corpus(path, args.bytes)
        log(f"[bench] corpus {path}: {os.path.getsize(path)} B ({time.perf_counter() - t0:.1f} s)")
    if world > 1:
        dist.barrier()
    total = os.path.getsize(path)
    eng = _lib.get_engine(local_rank)
    dev = f"cuda:{local_rank}"
    specials = [EOT]

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- this rank's shard, resident in HBM ------------------------------------------------------------------
    offs = D.shard_offsets(path, world, specials) if world > 1 else [0, total]
    lo, hi = offs[rank], offs[rank + 1]
    d_text, n_local = eng.load_file(path, lo, hi)
    torch.cuda.synchronize()
    eng.set_special_tokens(_lib.sort_specials(specials))

    result = {}

    def train_step_resident(design=None):
        design = design or args.merge_loop
        eng.unique_reset()
        eng.pretokenize_count(d_text, n_local)
        if world > 1:
            D.exchange_unique(eng)
        tr = BPETrainer(vocab_size=VOCAB_SIZE, special_tokens=specials)
        if world > 1 and design == "shared":
            D.merge_tokens_together(tr, eng)
        else:
            tr._merge_tokens(eng)
        result["trainer"] = tr

    for _ in range(args.warmup):
        train_step_resident()
    eng.set_profiling(False)  # restart the launch counters
    sampler = ClockSampler(local_rank)
    sync_all()
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        train_step_resident()
    ev1.record()
    sync_all()
    step_ms = max_over_ranks(ev0.elapsed_time(ev1) / args.steps)
    launches = sum(v[1] for v in eng.kernel_times().values())
    tr = result["trainer"]
    timing = tr.last_timing
    n_unique, n_unique_bytes, n_tokens = eng.unique_stats()
    log(f"[bench] train resident: {step_ms:.1f} ms/step, {len(tr.merges)} merges, {n_unique} distinct pretokens, "
        f"merge loop {timing['merge_loop_ms']:.1f} ms / {timing['iterations']} iterations")

    # ---- N > 1: the other design of the merge loop, measured the same way (DESIGN.md section 5 keeps both numbers) ----------
    designs = None
    if world > 1:
        other = "shared" if args.merge_loop == "replicated" else "replicated"
        train_step_resident(other)
        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(2):
            train_step_resident(other)
        e1.record()
        sync_all()
        other_ms = max_over_ranks(e0.elapsed_time(e1) / 2)
        tr_other = result["trainer"]
        designs = {args.merge_loop: {"step_ms": step_ms, "merge_loop_ms": timing["merge_loop_ms"]},
                   other: {"step_ms": other_ms, "merge_loop_ms": tr_other.last_timing["merge_loop_ms"]},
                   "timed": args.merge_loop, "same_merges": tr_other.merges == tr.merges}
        log(f"[bench] merge loop designs at N={world}: {args.merge_loop} {step_ms:.1f} ms/step (loop {timing['merge_loop_ms']:.1f}), "
            f"{other} {other_ms:.1f} ms/step (loop {tr_other.last_timing['merge_loop_ms']:.1f}), same merges {designs['same_merges']}")

    # ---- end to end through the public API ---------------------------------------------------------------------
    def train_e2e():
        if world == 1:
            return train_bpe(path, VOCAB_SIZE, specials)
        return D.train_bpe_sharded(path, VOCAB_SIZE, specials, merge_loop=args.merge_loop)

    for _ in range(2):  # the first call of a process allocates the pinned staging lanes and the piece buffer
        train_e2e()
    sync_all()
    t0 = time.perf_counter()
    calls = []
    for _ in range(args.steps):
        t1 = time.perf_counter()
        vocab, merges = train_e2e()
        calls.append(time.perf_counter() - t1)
    sync_all()
    e2e_s = max_over_ranks((time.perf_counter() - t0) / args.steps)
    log(f"[bench] train end to end: {e2e_s:.3f} s/call (" + ", ".join(f"{c * 1e3:.0f}" for c in calls) + " ms)")

    # ---- encode: configs[3] ------------------------------------------------------------------------------------
    tok = BPETokenizer(vocab, merges, specials)
    line_offs = D.line_shard_offsets(path, world)
    if world > 1:
        # encode shards are cut at line ends, not at document ends: reload this rank's lines
        llo, lhi = line_offs[rank], line_offs[rank + 1]
        d_enc, n_enc = eng.load_file(path, llo, lhi)
    else:
        d_enc, n_enc = d_text, n_local
    torch.cuda.synchronize()
    ids = None
    for _ in range(args.warmup):
        ids = tok.encode_bytes_to_device(None, flags=_lib.CUT_AFTER_NEWLINE, engine=eng, on_device=(d_enc, n_enc))
    sync_all()
    ev0.record()
    t_calls = [time.perf_counter()]
    for _ in range(args.steps):
        ids = tok.encode_bytes_to_device(None, flags=_lib.CUT_AFTER_NEWLINE, engine=eng, on_device=(d_enc, n_enc))
        t_calls.append(time.perf_counter())  # bpe_encode returns after its stream work: host time per call = device time
    ev1.record()
    sync_all()
    enc_ms = max_over_ranks(ev0.elapsed_time(ev1) / args.steps)
    log("[bench] encode resident, ms per call: " + ", ".join(f"{(b - a) * 1e3:.1f}" for a, b in zip(t_calls[:-1], t_calls[1:])))
    n_ids_local = int(ids.numel())
    # roundtrip at full size: GPU decode of the ids must give back the bytes
    back = tok.decode_array(ids, engine=eng, to_host=False) if n_enc else d_enc[:0]
    roundtrip_ok = bool(back.numel() == n_enc and torch.equal(back, d_enc[:n_enc]))
    del back

    # End to end twice: into a page-locked uint16 array that is allocated once and reused (ids arrive by DMA; what a
    # dataset-building loop would do — the headline), and into a fresh pageable NumPy array per call (the default).
    pinned_out = _lib.pinned_array(int(n_ids_local * 1.02) + 4096, np.uint16)

    def encode_e2e(out=None):
        if world == 1:
            return tok.encode_file_to_array(path, engine=eng, out=out)
        return D.encode_file_sharded(tok, path, out=out)[0]

    def time_e2e(out):
        encode_e2e(out)
        sync_all()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            arr = encode_e2e(out)
        sync_all()
        assert int(arr.size) == n_ids_local
        return max_over_ranks((time.perf_counter() - t0) / args.steps)

    enc_e2e_pageable_s = time_e2e(None)
    enc_e2e_s = time_e2e(pinned_out)
    n_ids_total = n_ids_local
    if world > 1:
        t = torch.tensor([n_ids_local], dtype=torch.int64, device=dev)
        dist.all_reduce(t)
        n_ids_total = int(t.item())
    log(f"[bench] encode: resident {total / enc_ms / 1e6:.2f} GB/s, end to end {total / enc_e2e_s / 1e9:.2f} GB/s into a reused page-locked array "
        f"({total / enc_e2e_pageable_s / 1e9:.2f} GB/s into a fresh pageable one), {n_ids_total} ids, roundtrip {roundtrip_ok}")

    # ---- per-kernel device time (separate pass, CUDA events around every launch) ---------------------------------
    eng.set_profiling(True)
    train_step_resident()
    k_train = eng.kernel_times()
    eng.set_profiling(True)  # the training above emptied the pretoken table: this encode is a first one (split + count in one pass, new pretokens encoded) ...
    tok.encode_bytes_to_device(None, flags=_lib.CUT_AFTER_NEWLINE, engine=eng, on_device=(d_enc, n_enc))
    k_enc_first = eng.kernel_times()
    eng.set_profiling(True)  # ... and this one runs with the cache warm, like the timed calls
    tok.encode_bytes_to_device(None, flags=_lib.CUT_AFTER_NEWLINE, engine=eng, on_device=(d_enc, n_enc))
    k_enc = eng.kernel_times()
    eng.set_profiling(False)
    clocks = sampler.stop() if rank == 0 else {}

    peaks_file = ROOT / "MEASURED_PEAKS.json"
    if peaks_file.exists():
        peak, peak_src = float(json.loads(peaks_file.read_text())["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    else:
        peak, peak_src = 6650.0, "B200_PROFILING.md fallback (of fallback)"

    def gbps(nbytes, ms):
        return nbytes / ms / 1e6 if ms else None

    kernels = {}
    for name, (ms, n) in k_train.items():
        kernels[name] = {"ms": ms, "launches": n}
    for name, (ms, n) in k_enc.items():
        kernels["encode:" + name] = {"ms": ms, "launches": n}
    for name, (ms, n) in k_enc_first.items():
        kernels["encode_first_call:" + name] = {"ms": ms, "launches": n}
    hbm_kernels = {k: gbps(n_local, k_train[k][0]) for k in ("fused_count_kernel", "scan_kernel", "count_kernel") if k in k_train and k_train[k][0]}
    dominant = min(hbm_kernels, key=hbm_kernels.get) if hbm_kernels else None
    # DRAM bytes per launch from the committed `ncu --set full` capture (tools/ncu_summary.py); only quoted for the
    # corpus size it was taken on, one GPU.
    traffic_file = ROOT / "profiles" / "r02_ncu_dram_bytes_per_launch_2.1GB.json"
    ncu_traffic = {}
    if traffic_file.exists() and world == 1 and abs(total - 2_100_000_403) < 1_000_000:
        ncu_traffic = json.loads(traffic_file.read_text())
    ncu_name = {"scan_kernel": "scan_kernel<0>", "count_kernel": "count_kernel", "emit_fused_kernel": "emit_fused_kernel<unsigned short>",
                "fused_count_kernel": "fused_count_kernel<0, 0>", "fused_emit_kernel": "fused_emit_kernel"}
    roofline = None
    if dominant:
        others = {k: {"achieved": v, "frac": v / peak, "traffic": ncu_traffic.get(ncu_name[k])} for k, v in hbm_kernels.items() if k != dominant}
        if k_enc.get("emit_fused_kernel", (0, 0))[0]:
            emit_bytes = int(n_enc + ids.numel() * ids.element_size())  # N read + 2 T written
            v = gbps(emit_bytes, k_enc["emit_fused_kernel"][0])
            others["encode:emit_fused_kernel"] = {"achieved": v, "frac": v / peak, "traffic": ncu_traffic.get(ncu_name["emit_fused_kernel"]),
                                                  "algorithmic_bytes_per_launch": emit_bytes}
        if k_enc.get("scan_kernel", (0, 0))[0]:
            scan_bytes = int(n_enc + n_enc // 8)  # N read + N/8 of mask written
            v = gbps(scan_bytes, k_enc["scan_kernel"][0])
            others["encode:scan_kernel"] = {"achieved": v, "frac": v / peak, "traffic": ncu_traffic.get("scan_kernel<1>"),
                                            "algorithmic_bytes_per_launch": scan_bytes}
        roofline = {"kernel": dominant, "bound": "hbm", "achieved": hbm_kernels[dominant], "peak": peak, "unit": "GB/s",
                    "frac": hbm_kernels[dominant] / peak, "traffic": ncu_traffic.get(ncu_name[dominant]), "peak_source": peak_src,
                    "traffic_source": "profiles/r02_ncu_dram_bytes_per_launch_2.1GB.json (ncu --set full, dram read + write, per launch)" if ncu_traffic else None,
                    "algorithmic_bytes_per_launch": n_local,
                    "note": "scan/count/emit are instruction-issue bound, not HBM bound (DESIGN.md section 4); the merge loop is latency bound and reported as us_per_merge",
                    "others": others}

    line = {
        "metric": METRIC, "value": step_ms / 1e3, "unit": "s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": step_ms, "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(total, world),
        "e2e": {"value": e2e_s, "unit": "s", "h2d_bytes_per_step": int(n_local), "d2h_bytes_per_step": int(8 * len(merges) + 64),
                "api": "bpe_transformer_b200.train_bpe(path, 10000, ['<|endoftext|>'])" if world == 1 else "bpe_transformer_b200.distributed.train_bpe_sharded"},
        "gpu_launches": int(launches),
        "roofline": roofline,
        "merge_loop": {"ms": timing["merge_loop_ms"], "iterations": timing["iterations"], "merges": timing["merges"],
                       "us_per_merge": timing["merge_loop_ms"] * 1e3 / max(1, timing["merges"]), "pair_init_ms": timing["pair_init_ms"]},
        "encode": {"metric": "encode_GBps_corpus_to_uint16", "value": total / enc_ms / 1e6, "unit": "GB/s",
                   "e2e": {"value": total / enc_e2e_s / 1e9, "unit": "GB/s", "h2d_bytes_per_step": int(n_enc), "d2h_bytes_per_step": int(2 * n_ids_local),
                           "api": "BPETokenizer.encode_file_to_array(path, out=<page-locked uint16 array, allocated once>) (= encode_iterable(open(path)) as an array)",
                           "fresh_pageable_array_per_call": total / enc_e2e_pageable_s / 1e9},
                   "ids": n_ids_total, "algorithmic_bytes": int(total + 2 * n_ids_total), "roundtrip_decode_equals_input": roundtrip_ok},
        "merge_loop_designs": designs,
        "distinct_pretokens": int(n_unique), "pretokens": int(n_tokens), "merges": len(merges),
        "kernels": kernels, "clocks": clocks,
    }
    # ---- the results themselves: hashes at every N (the N-GPU merges and ids must equal the 1-GPU ones) --------------
    import hashlib

    ids_host = _lib.host_array(n_ids_local, np.uint16)
    eng.store_into(ids, ids_host)
    ids_sha = None
    if world == 1:
        ids_sha = hashlib.sha256(ids_host.tobytes()).hexdigest()
    else:  # the ranks' ids in rank order, through files on this box (one node): rank 0 hashes the concatenation
        part = f"{path}.ids.rank{rank}"
        ids_host.tofile(part)
        dist.barrier()
        if rank == 0:
            h = hashlib.sha256()
            for r in range(world):
                with open(f"{path}.ids.rank{r}", "rb") as fh:
                    while True:
                        block = fh.read(64 << 20)
                        if not block:
                            break
                        h.update(block)
            ids_sha = h.hexdigest()
        dist.barrier()
        os.unlink(part)
    parity = {"merges_sha256": merges_sha256(merges), "ids_sha256": ids_sha, "resident_merges_equal_e2e_merges": merges == tr.merges,
              "roundtrip_decode_equals_input": roundtrip_ok}
    line["parity"] = parity
    failed = [k for k in ("resident_merges_equal_e2e_merges", "roundtrip_decode_equals_input") if not parity[k]]

    # ---- CPU side, same run, same box: the reference's algorithm over the WHOLE corpus, and the parity gate -------------
    if rank == 0 and world == 1 and not args.no_cpu:
        from collections import Counter

        from bpe_transformer_b200.tokenization.preprocessing.pretokenization import counter_from_engine

        t0 = time.perf_counter()
        full = cpu_train_full(path)
        line["cpu_baseline"] = {k: v for k, v in full.items() if not k.startswith("_")}
        log(f"[bench] cpu baseline (full corpus, measured): train {full['value']:.1f} s")
        parity["oracle"] = "oracle/bpe_ref_port.py (the reference's algorithm) over the whole corpus; ids: oracle/bpe_oracle.c over the whole corpus + the port on a prefix"
        parity["merges_equal"] = merges == full["_merges"]
        parity["vocab_equal"] = vocab == full["_vocab"]
        parity["cpu_merges_sha256"] = merges_sha256(full["_merges"])
        eng.unique_reset()
        eng.pretokenize_count(d_text, n_local)
        parity["pretoken_counter_equal"] = counter_from_engine(eng) == Counter(full["_counter"])
        # ids: the C oracle over the whole corpus (threads over line ranges), the port over a prefix (also the timed baseline)
        c_sha, c_n = c_oracle_ids_sha256(path, vocab, merges)
        parity["ids_equal_full_corpus"] = (c_sha == ids_sha and c_n == n_ids_local)
        parity["cpu_ids_sha256"] = c_sha
        enc_cpu = cpu_encode_sample(path, vocab, merges, min(total, args.cpu_sample_bytes), min(total, 2_000_000))
        ref_ids = enc_cpu.pop("_ids_serial")
        enc_cpu.pop("_serial_bytes")
        parity["ids_equal_first_64MB"] = bool(ref_ids.size <= ids_host.size and np.array_equal(ids_host[:ref_ids.size].astype(np.uint32), ref_ids))
        line["encode"]["cpu_baseline"] = enc_cpu
        line["encode_vs_cpu"] = {"gpu_e2e_GBps": line["encode"]["e2e"]["value"], "cpu_GBps": enc_cpu["value"], "cpu_variant": enc_cpu["variant"],
                                 "ratio": line["encode"]["e2e"]["value"] / enc_cpu["value"]}
        failed += [k for k in ("merges_equal", "vocab_equal", "pretoken_counter_equal", "ids_equal_full_corpus", "ids_equal_first_64MB") if not parity[k]]
        log(f"[bench] parity vs the CPU run: " + ", ".join(f"{k}={parity[k]}" for k in ("merges_equal", "vocab_equal", "pretoken_counter_equal",
                                                                                          "ids_equal_full_corpus", "ids_equal_first_64MB"))
            + f"; cpu encode {enc_cpu['value'] * 1e3:.2f} MB/s ({enc_cpu['variant']}); CPU side took {time.perf_counter() - t0:.1f} s")
    if rank == 0:
        emit_line(line)
    if failed:
        log(f"[bench] PARITY FAILED: {failed}")
        raise SystemExit(3)


_json_fd = None


def guard_stdout():
    """stdout carries ONE JSON line: whatever a library prints to file descriptor 1 (NCCL's version banner does) goes to
    stderr instead, and the line is written to the real stdout by emit_line()."""
    global _json_fd
    if _json_fd is None:
        sys.stdout.flush()
        _json_fd = os.dup(1)
        os.dup2(2, 1)


def emit_line(line: dict):
    data = (json.dumps(line) + "\n").encode()
    if _json_fd is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_json_fd, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--bytes", type=int, default=int(os.environ.get("BPE_BENCH_BYTES", 2_100_000_000)))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-sample-bytes", type=int, default=64_000_000)
    ap.add_argument("--merge-loop", default="replicated", choices=["replicated", "shared"],
                    help="N > 1: the design of the merge loop that is timed (the other one is measured beside it)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline and the parity gate against it")
    ap.add_argument("--no-cpu-encode", action="store_true", help="--impl reference: skip the CPU encode timings")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    guard_stdout()
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG", "WARN")  # NCCL's version banner goes to stdout otherwise: stdout carries ONE JSON line
        import torch
        import torch.distributed as dist

        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))
    try:
        run_gpu(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist

            dist.destroy_process_group()


if __name__ == "__main__":
    main()
