#!/usr/bin/env python
"""One blg_kmer_jaccard call over two device-resident inputs (for an ncu launch list of the sort / unique / lookup kernels).

    python tools/jaccard_probe.py [--bases 50000000] [--reps 2]
"""
import argparse
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bases", type=int, default=50_000_000)
    ap.add_argument("--reps", type=int, default=2)
    a = ap.parse_args()
    import torch
    import biolib_b200
    from biolib_b200 import synth, _capi
    ctx = biolib_b200.Context(0)
    s1, s2 = synth.related_genomes_slice(0, a.bases)
    d1, d2 = torch.from_numpy(s1).cuda(), torch.from_numpy(s2).cuda()
    p = biolib_b200.context.Params(k=31, m=0, canonical=1, flags=_capi.F_DEVICE_INPUT, seed=0)
    for _ in range(a.reps):
        o = _capi.JaccardOut()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        st = ctx.L.blg_kmer_jaccard(ctx.h, C.c_void_p(d1.data_ptr()), a.bases, None, 0, C.c_void_p(d2.data_ptr()), a.bases, None, 0, C.byref(p), C.byref(o))
        torch.cuda.synchronize()
        assert st == 0
        print(f"bases_each={a.bases} ms={(time.perf_counter() - t0) * 1e3:.2f} kernel_ms={ctx.stats()['kernel_ms']:.2f} result={(o.intersection, o.union_size, o.size1, o.size2)}")
    ctx.close()


if __name__ == "__main__":
    main()
