#!/usr/bin/env python
"""Development tool: end-to-end (pinned host in, pinned host out) time of blg_super_kmers for the bench workload."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import biolib_b200
from biolib_b200 import synth
import ctypes as C
n = 100_000_000
ctx = biolib_b200.Context(0)
s = synth.generate(42, n)
cap = int(n * 2.0 / 14 * 1.3) + 65536
def npp(ptr, nb, dt): return np.frombuffer((C.c_uint8 * nb).from_address(ptr), dtype=dt)
hin = ctx.host_alloc(n); npp(hin, n, np.uint8)[:] = s
hp = [ctx.host_alloc(cap * 8), ctx.host_alloc(cap * 8), ctx.host_alloc(cap), ctx.host_alloc(cap)]
for _ in range(2): cnt = ctx.super_kmers_raw(hin, n, 31, 19, 42, True, hp[0], hp[1], hp[2], hp[3], cap)
ts = []
for _ in range(8):
    t0 = time.perf_counter(); ctx.super_kmers_raw(hin, n, 31, 19, 42, True, hp[0], hp[1], hp[2], hp[3], cap); ts.append(time.perf_counter() - t0)
st = ctx.stats()
print(f"chunk_m={os.environ.get('BLG_PIPE_CHUNK_M','-')} nopipe={os.environ.get('BLG_NO_PIPELINE','0')} e2e_ms median={np.median(ts)*1e3:.3f} min={min(ts)*1e3:.3f} "
      f"Gbases/s={n/np.median(ts)/1e9:.2f} launches={st['kernel_launches']} kernel_ms={st['kernel_ms']:.3f} records={cnt}")
