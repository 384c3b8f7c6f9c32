#!/usr/bin/env python
"""Development tool: end-to-end (pinned host in, pinned host out) time of blg_super_kmers for the bench workload, over the
shapes of the chunked host path (blg_set_pipeline) and the four record layouts.

    python tools/e2e.py [--layouts tight,runs,pos32,full] [--chunks 4,6,8,12,16] [--divs 2,4,8]
"""
import argparse
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import biolib_b200
from biolib_b200 import _capi, synth


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layouts", default="runs,pos32,full")
    ap.add_argument("--chunks", default="4,6,8,12,16")
    ap.add_argument("--divs", default="2,4,8")
    ap.add_argument("--reps", type=int, default=8)
    a = ap.parse_args()
    n = 100_000_000
    ctx = biolib_b200.Context(0)
    s = synth.generate(42, n)
    cap = int(n * 2.0 / 14 * 1.3) + 65536
    run_cap = 1 << 16

    def npp(ptr, nb, dt):
        return np.frombuffer((C.c_uint8 * nb).from_address(ptr), dtype=dt)
    hin = ctx.host_alloc(n)
    npp(hin, n, np.uint8)[:] = s
    hp = [ctx.host_alloc(cap * 8), ctx.host_alloc(cap * 8), ctx.host_alloc(cap), ctx.host_alloc(cap)]
    hr = [ctx.host_alloc(run_cap * 8), ctx.host_alloc(run_cap * 8)]
    flags = {"full": 0, "pos32": _capi.F_POS32, "runs": _capi.F_RUNS, "tight": _capi.F_RUNS | _capi.F_TIGHT}
    for lay in a.layouts.split(","):
        f = flags[lay]

        def call():
            return ctx.super_kmers_raw(hin, n, 31, 19, 42, True, hp[0], hp[1], hp[2], hp[3], cap, flags=f,
                                       run_ptrs=hr if f & _capi.F_RUNS else None, run_capacity=run_cap if f & _capi.F_RUNS else 0,
                                       tight_ptrs=(hp[0], hp[2], hp[3]) if f & _capi.F_TIGHT else None)
        for ch in (int(x) for x in a.chunks.split(",")):
            for dv in (int(x) for x in a.divs.split(",")):
                ctx._check(ctx.L.blg_set_pipeline(ctx.h, ch, dv))
                for _ in range(2):
                    cnt = call()
                ts = []
                for _ in range(a.reps):
                    t0 = time.perf_counter()
                    call()
                    ts.append(time.perf_counter() - t0)
                st = ctx.stats()
                print(f"layout={lay} chunk_mi={ch} first_div={dv} e2e_ms median={np.median(ts) * 1e3:.3f} min={min(ts) * 1e3:.3f} "
                      f"Gbases/s={n / np.median(ts) / 1e9:.2f} launches={st['kernel_launches']} d2h={st['bytes_d2h']} records={cnt}", flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
