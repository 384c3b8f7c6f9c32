#!/usr/bin/env python
"""Development tool: device-resident timing of the two small entry points of the path,
blg_hash64 (hash::hash64) and blg_minimizer_positions (minimizer_position_extractor, row X1).

    python tools/auxbench.py [--n 100000000]

CUDA events on the context stream around each call (the calls only enqueue when both pointers are device pointers),
L2 flushed between runs, median of 10. Not part of the product or of bench.py.
"""
import argparse
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=100_000_000)
    args = ap.parse_args()
    import torch
    import biolib_b200
    from biolib_b200.context import F_DEVICE_INPUT, F_DEVICE_OUTPUT

    dev = torch.device("cuda", 0)
    ctx = biolib_b200.Context(0)
    torch.cuda.set_stream(torch.cuda.Stream())   # not the legacy stream 0 (blg_set_stream reads 0 as "the context's own stream")
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    n = args.n
    vals = torch.randint(0, 2**62, (n,), dtype=torch.int64, device=dev)
    out64 = torch.empty(n, dtype=torch.int64, device=dev)
    out8 = torch.empty(n, dtype=torch.uint8, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    fl = F_DEVICE_INPUT | F_DEVICE_OUTPUT

    def timed(fn):
        ms = []
        for i in range(13):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            if i >= 3:
                ms.append(e0.elapsed_time(e1))
        return float(np.median(ms))

    t = timed(lambda: ctx._check(ctx.L.blg_hash64(ctx.h, C.c_void_p(vals.data_ptr()), n, 42, fl, C.c_void_p(out64.data_ptr()))))
    print(f"hash64: n={n} ms={t:.4f} Gvalues/s={n / t / 1e6:.1f} GB/s={16 * n / t / 1e6:.0f} (16 B per value)")
    for k, m in ((31, 19), (21, 11), (32, 16)):
        kv = vals & ((1 << (2 * k)) - 1) if k < 32 else vals
        t = timed(lambda: ctx._check(ctx.L.blg_minimizer_positions(ctx.h, C.c_void_p(kv.data_ptr()), n, k, m, fl, C.c_void_p(out8.data_ptr()))))
        w = k - m + 1
        print(f"minimizer_positions k={k} m={m}: n={n} ms={t:.4f} Gkmers/s={n / t / 1e6:.1f} Ghashes/s={n * w / t / 1e6:.0f} GB/s={9 * n / t / 1e6:.0f} (9 B per k-mer)")
    ctx.close()


if __name__ == "__main__":
    main()
