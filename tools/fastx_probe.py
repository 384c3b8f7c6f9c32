#!/usr/bin/env python
"""Development tool: wall time of blg_parse_fastx (device text in, device bases + offsets out) on FASTA and FASTQ text.

    python tools/fastx_probe.py [--n 100000000]     (under ncu: -k regex:fx_ lists the three kernels)
"""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=100_000_000)
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    import torch
    import biolib_b200
    from biolib_b200 import synth
    ctx = biolib_b200.Context(0)
    s = synth.generate(42, a.n)
    # FASTA: reads of 10 kbp, 80-column lines; FASTQ: reads of 150 bp
    L, wrap = 10_000, 80
    nrec = a.n // L
    body = np.concatenate([s[: nrec * L].reshape(nrec, L // wrap, wrap), np.full((nrec, L // wrap, 1), 10, np.uint8)], axis=2).reshape(nrec, -1)
    fasta = np.concatenate([np.tile(np.frombuffer(b">read_0000000000\n", np.uint8), (nrec, 1)), body], axis=1).reshape(-1)
    R = 150
    nq = a.n // R
    seqs = s[: nq * R].reshape(nq, R)
    nl = np.full((nq, 1), 10, np.uint8)
    fastq = np.concatenate([np.tile(np.frombuffer(b"@r0000000/1\n", np.uint8), (nq, 1)), seqs, nl, np.tile(np.frombuffer(b"+\n", np.uint8), (nq, 1)),
                            np.full((nq, R), ord("I"), np.uint8), nl], axis=1).reshape(-1)
    for name, text, fmt, want_b, want_s in (("fasta", fasta, 0, nrec * L, nrec), ("fastq", fastq, 1, nq * R, nq)):
        d = torch.from_numpy(text).cuda()
        best = None
        for _ in range(a.reps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            fx = ctx.parse_fastx_device(d.data_ptr(), text.size, fmt)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            assert fx.n_bases == want_b and fx.n_seq == want_s, (fx.n_bases, fx.n_seq)
            ctx.free_fastx(fx)
            best = dt if best is None else min(best, dt)
        print(f"{name}: text {text.size} B, {want_s} records, {best * 1e3:.3f} ms, {text.size / best / 1e9:.1f} GB/s of text")
    ctx.close()


if __name__ == "__main__":
    main()
