#!/usr/bin/env python
"""Development A/B tool: kernel time of blg_super_kmers (device-resident) for the library at BLG_LIB_PATH.

    BLG_LIB_PATH=biolib_b200/variants/libX.so python tools/quickbench.py [--k 31 --m 19 --n 100000000 --steps 10]

Prints one line: variant, kernel ms (median of steps, CUDA events inside the library), Gbases/s and an
order-dependent checksum of the records so two builds can be compared for identical output without the oracle.
Not part of the product or of bench.py.
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--k", type=int, default=31)
    ap.add_argument("--m", type=int, default=19)
    ap.add_argument("--n", type=int, default=100_000_000)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--canon", type=int, default=1)
    ap.add_argument("--breaks", type=float, default=0.0, help="fraction of bases replaced by N")
    ap.add_argument("--reads", type=int, default=0, help="split into sequences of this length (0 = one sequence)")
    args = ap.parse_args()
    import torch
    import biolib_b200
    from biolib_b200 import synth

    dev = torch.device("cuda", 0)
    ctx = biolib_b200.Context(0)
    s = synth.generate(42, args.n)
    if args.breaks > 0:
        rng = np.random.default_rng(7)
        s = s.copy()
        s[rng.random(args.n) < args.breaks] = ord("N")
    offs_d = None
    n_seq = 0
    if args.reads:
        n_seq = args.n // args.reads
        offs = (np.arange(n_seq + 1, dtype=np.uint64) * np.uint64(args.reads))
        offs[-1] = args.n
        offs_d = torch.from_numpy(offs.view(np.int64)).to(dev)
    w = args.k - args.m + 1
    cap = int(args.n * (2.0 / (w + 1) * 1.3 + (2.0 / args.reads if args.reads else 0) + 4 * args.breaks)) + 65536
    d_bases = torch.from_numpy(s).to(dev)
    d_value = torch.empty(cap, dtype=torch.int64, device=dev)
    d_mm = torch.empty(cap, dtype=torch.int64, device=dev)
    d_size = torch.empty(cap, dtype=torch.uint8, device=dev)
    d_pos = torch.empty(cap, dtype=torch.uint8, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    torch.cuda.set_stream(torch.cuda.Stream())   # not the legacy stream 0 (blg_set_stream reads 0 as "the context's own stream")
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    ms = []
    count = 0
    for i in range(3 + args.steps):
        flush.zero_()
        torch.cuda.synchronize()      # an event recorded while the flush still runs time-stamps early on this driver (bench.py DeviceRun)
        count = ctx.super_kmers_device(d_bases.data_ptr(), args.n, args.k, args.m, 42, bool(args.canon), d_value.data_ptr(),
                                       d_mm.data_ptr(), d_size.data_ptr(), d_pos.data_ptr(), cap,
                                       d_offsets=offs_d.data_ptr() if offs_d is not None else 0, n_seq=n_seq)
        if i >= 3:
            ms.append(ctx.stats()["kernel_ms"])
    torch.cuda.synchronize()
    idx = torch.arange(1, count + 1, dtype=torch.int64, device=dev)
    chk = int((d_value[:count] * idx).sum().item()) ^ int((d_mm[:count] * idx).sum().item()) ^ \
        int((d_size[:count].to(torch.int64) * idx).sum().item() * 257 + (d_pos[:count].to(torch.int64) * idx).sum().item())
    med = float(np.median(ms))
    print(f"{os.environ.get('BLG_LIB_PATH', 'default'):48s} k={args.k} m={args.m} n={args.n} reads={args.reads} breaks={args.breaks} "
          f"kernel_ms={med:.4f} min={min(ms):.4f} Gbases/s={args.n / med / 1e6:.1f} records={count} chk={chk & (2**64 - 1):016x}",
          flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
