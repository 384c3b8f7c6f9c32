"""Development probe: per-step torch-event time around the device-resident call vs the library's own kernel_ms."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
import biolib_b200
from biolib_b200 import synth
dev = torch.device("cuda", 0)
ctx = biolib_b200.Context(0)
n = 100_000_000
s = synth.generate(42, n)
cap = int(n * 2.0 / 14 * 1.3) + 65536
d_bases = torch.from_numpy(s).to(dev)
d_value = torch.empty(cap, dtype=torch.int64, device=dev); d_mm = torch.empty(cap, dtype=torch.int64, device=dev)
d_size = torch.empty(cap, dtype=torch.uint8, device=dev); d_pos = torch.empty(cap, dtype=torch.uint8, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
dummy = torch.zeros(1024, device=dev)
torch.cuda.set_stream(torch.cuda.Stream())
stream = torch.cuda.current_stream()
ctx.set_stream(stream.cuda_stream)
for mode in ("plain", "dummy_after_flush", "sync_after_flush", "noflush"):
    rows = []
    for i in range(8):
        if mode != "noflush":
            flush.zero_()
        if mode == "dummy_after_flush":
            dummy.add_(1.0)
        if mode == "sync_after_flush":
            torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        ctx.super_kmers_device(d_bases.data_ptr(), n, 31, 19, 42, True, d_value.data_ptr(), d_mm.data_ptr(), d_size.data_ptr(), d_pos.data_ptr(), cap)
        b.record(stream)
        torch.cuda.synchronize()
        rows.append((a.elapsed_time(b), ctx.stats()["kernel_ms"]))
    print(mode, " ".join(f"{x:.3f}/{y:.3f}" for x, y in rows[2:]), flush=True)
ctx.close()
