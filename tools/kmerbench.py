#!/usr/bin/env python
"""Development tool: kernel time of blg_kmers (kmer_view) with device-resident input and output.

    python tools/kmerbench.py [--k 31 --n 100000000 --canon 1 --steps 10 --pos 1]

One line: kernel ms (median, CUDA events inside the library, L2 flushed between runs), Gbases/s, algorithmic GB/s
(1 B per base + 8 B per value word + 8 B per position) and its share of the measured HBM copy peak
(MEASURED_PEAKS.json when present). Not part of the product or of bench.py.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--k", type=int, default=31)
    ap.add_argument("--n", type=int, default=100_000_000)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--canon", type=int, default=1)
    ap.add_argument("--pos", type=int, default=1)
    ap.add_argument("--breaks", type=float, default=0.0)
    ap.add_argument("--syncmer", type=int, default=0, help="m: closed-syncmer filter (blg_syncmers, offsets 0 and k - m)")
    ap.add_argument("--hash-sample", type=float, default=-1.0, help="rate: hash_sampler filter (blg_hash_sampled_kmers, seed 42)")
    ap.add_argument("--reads", type=int, default=0, help="split into sequences of this length (0 = one sequence)")
    args = ap.parse_args()
    import torch
    import biolib_b200
    from biolib_b200 import synth

    peak = 6546.2
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    dev = torch.device("cuda", 0)
    ctx = biolib_b200.Context(0)
    s = synth.generate(42, args.n)
    if args.breaks > 0:
        rng = np.random.default_rng(7)
        s = s.copy()
        s[rng.random(args.n) < args.breaks] = ord("N")
    d_bases = torch.from_numpy(s).to(dev)
    offs_d, n_seq = None, 0
    if args.reads:
        n_seq = args.n // args.reads
        offs = (np.arange(n_seq + 1, dtype=np.uint64) * np.uint64(args.reads))
        offs[-1] = args.n
        offs_d = torch.from_numpy(offs.view(np.int64)).to(dev)
    cap = args.n
    wide = args.k > 32
    d_lo = torch.empty(cap, dtype=torch.int64, device=dev)
    d_hi = torch.empty(cap if wide else 1, dtype=torch.int64, device=dev)
    d_pos = torch.empty(cap if args.pos else 1, dtype=torch.int64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    torch.cuda.set_stream(torch.cuda.Stream())   # not the legacy stream 0 (blg_set_stream reads 0 as "the context's own stream")
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    ms, count = [], 0
    for i in range(3 + args.steps):
        flush.zero_()
        count = ctx.kmers_device(d_bases.data_ptr(), args.n, args.k, bool(args.canon), d_lo.data_ptr(), d_hi.data_ptr() if wide else 0,
                                 d_pos.data_ptr() if args.pos else 0, cap, d_offsets=offs_d.data_ptr() if offs_d is not None else 0, n_seq=n_seq,
                                 syncmer=(args.syncmer, 0, args.k - args.syncmer) if args.syncmer else None,
                                 hash_sample=(42, ctx.hash_sampler_threshold(args.hash_sample)) if args.hash_sample >= 0 else None)
        if i >= 3:
            ms.append(ctx.stats()["kernel_ms"])
    torch.cuda.synchronize()
    med = float(np.median(ms))
    per = 8 * (2 if wide else 1) + (8 if args.pos else 0)
    gb = (args.n + per * count) / 1e9
    idx = torch.arange(1, count + 1, dtype=torch.int64, device=dev)
    chk = int((d_lo[:count] * idx).sum().item())
    if args.pos:
        chk ^= int((d_pos[:count] * idx).sum().item())
    if wide:
        chk ^= int((d_hi[:count] * (idx + 1)).sum().item())
    print(f"k={args.k} syncmer_m={args.syncmer} hash_sample={args.hash_sample} canon={args.canon} pos={args.pos} reads={args.reads} breaks={args.breaks} kmers={count} kernel_ms={med:.4f} min={min(ms):.4f} "
          f"Gbases/s={args.n / med / 1e6:.1f} GB/s={gb / med * 1e3:.0f} frac_of_{peak:.0f}={gb / med * 1e3 / peak:.3f} chk={chk & (2**64 - 1):016x}", flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
