#!/usr/bin/env python
"""Sort / unique / Jaccard of two related genomes across N GPUs (biolib_b200/multi_gpu.py), checked against the one-GPU call.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        tools/multi_gpu_jaccard.py [--bases 200000000] [--k 31]

Two inputs of --bases bases (the second = the first with 2 % substitutions), each cut into N chunks of k-mer starts (+ k - 1 bases);
rank r holds chunk r of both. Timed (wall clock between barriers + device synchronise, max over ranks): local extraction + sort +
unique, the agreement on key ranges, the two all-to-alls, the merge, the per-range Jaccard and the final all-reduce (both strategies of
biolib_b200/multi_gpu.py).
Rank 0 then runs blg_kmer_jaccard over the whole inputs on its GPU alone and compares the four counters.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bases", type=int, default=200_000_000)
    ap.add_argument("--k", type=int, default=31)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--strategy", default="partition", help="partition (exchange, then sort once) or sorted (sort, exchange distinct keys, merge)")
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    import biolib_b200
    from biolib_b200 import multi_gpu as mg, synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ctx = biolib_b200.Context(local)
    ops = mg.CudaKmerOps(ctx, torch, dev)

    n, k = a.bases, a.k
    make_inputs = synth.related_genomes_slice

    lo, hi = mg.kmer_chunk(n, k, world, rank)
    s1, s2 = make_inputs(lo, hi)
    d1, d2 = torch.from_numpy(s1).to(dev), torch.from_numpy(s2).to(dev)

    def barrier():
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()

    res, times = None, []
    for _ in range(a.reps + 1):
        barrier()
        t0 = time.perf_counter()
        res = mg.distributed_jaccard(ops, dist, d1, d2, k, True, strategy=a.strategy)
        barrier()
        times.append(time.perf_counter() - t0)
    t = torch.tensor([min(times[1:])], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0]) * 1e3

    line = {"n_gpus": world, "k": k, "bases_each": n, "ms": ms, "kmers_per_s": 2 * (n - k + 1) / (ms * 1e-3), **res}
    # the one-GPU call over the whole inputs, on rank 0
    if rank == 0:
        w1, w2 = (s1, s2) if world == 1 else make_inputs(0, n)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        one = ctx.kmer_jaccard(w1, w2, k, True)
        line["one_gpu_host_input_ms"] = (time.perf_counter() - t0) * 1e3
        line["one_gpu_result"] = [int(x) for x in one]
        line["equal_to_one_gpu"] = tuple(int(x) for x in one) == (res["intersection"], res["union"], res["size1"], res["size2"])
        print(json.dumps(line))
        assert line["equal_to_one_gpu"], line
    dist.barrier()
    dist.destroy_process_group()
    ctx.close()


if __name__ == "__main__":
    main()
