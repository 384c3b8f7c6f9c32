#!/usr/bin/env python
"""Development tool: single-thread time of the reference's own sampler::syncmer_sampler over wrapper::kmer_view
(oracle/_ref, reference include/syncmer_sampler.hpp:130-137) on a bounded sample, to sit beside tools/kmerbench.py --syncmer.

    python tools/syncmer_cpu.py [--n 8000000]
"""
import argparse
import os
import sys
import time

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=8_000_000)
    args = ap.parse_args()
    import pyoracle
    from biolib_b200 import synth
    ref = pyoracle.Ref()
    s = synth.generate(42, args.n)
    for k, m in ((31, 19), (21, 11)):
        for canon in (1, 0):
            t = time.perf_counter()
            r = ref.syncmer_stream(s, k, m, 0, k - m, canon)
            dt = time.perf_counter() - t
            print(f"reference syncmer_sampler over kmer_view, 1 thread: k={k} m={m} canon={canon} {args.n} bases in {dt:.3f} s = "
                  f"{args.n / dt / 1e9:.5f} Gbases/s, {len(r['lo'])} syncmers", flush=True)


if __name__ == "__main__":
    main()
