"""Developer tool: what the box's PCIe link gives with page-locked buffers — host->device alone, device->host alone, both at once.
The ceiling of any end-to-end number that moves the corpus in and the ids out.

    python tools/pcie_duplex.py [MB per direction = 1024]
"""
import sys, time
import torch

mb = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
n = mb << 20
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.zeros(n, dtype=torch.uint8, device="cuda")
s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()


def run(h2d: bool, d2h: bool) -> float:
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    if h2d:
        with torch.cuda.stream(s_in):
            d_in.copy_(h_in, non_blocking=True)
    if d2h:
        with torch.cuda.stream(s_out):
            h_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize()
    return time.perf_counter() - t0


for name, a, b in (("host->device alone", True, False), ("device->host alone", False, True), ("both at once", True, True)):
    run(a, b)
    dt = min(run(a, b) for _ in range(5))
    each = n / dt / 1e9
    print(f"{name}: {mb} MB per direction in {dt * 1e3:.2f} ms = {each:.1f} GB/s per direction" + (f", {2 * each:.1f} GB/s in total" if a and b else ""), flush=True)
