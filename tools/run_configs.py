#!/usr/bin/env python
"""BASELINE.json configurations 2-5 at full size on N GPUs of one box (not the bench contract: bench.py measures
configs[1]; this records the others beside it, one JSON line per configuration, rank 0 prints).

    python tools/run_configs.py                       # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/run_configs.py

Device-resident timing (CUDA events inside the library around each kernel, L2 flushed between runs, max over ranks);
sharding by biolib_b200.shard (chunks with halos for one long sequence, read ranges for batches); the only collective
is the sum of the counts. The CPU arm beside each line runs the reference's read_base()+hash64 verbatim (oracle/_ref,
when present) + the restated windowing on a bounded sample with all host threads, rank 0 only.
"""
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def cpu_arm(chunks, k, m):
    import pyoracle
    if os.path.exists(pyoracle.REF_SO):
        eng = pyoracle.Ref()
        fn = lambda s: eng.super_kmers_sums(s, k, m, 42, True)[0]
        kind = "reference"
    else:
        eng = pyoracle.Oracle()
        fn = lambda s: eng.super_kmers(s, k, m, 42, True, want=())["n"]
        kind = "port"
    ths = [threading.Thread(target=fn, args=(c,)) for c in chunks]
    t0 = time.perf_counter()
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    dt = time.perf_counter() - t0
    return sum(c.size for c in chunks) / dt / 1e9, kind, len(chunks)


def main():
    import torch
    import torch.distributed as dist
    import biolib_b200
    from biolib_b200 import shard, synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    ctx = biolib_b200.Context(local)
    torch.cuda.set_stream(torch.cuda.Stream())   # not the legacy stream 0 (blg_set_stream reads 0 as "the context's own stream")
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    only = set(sys.argv[1:])

    def timed(d_bases, n, k, m, offs=None, win=None, steps=5):
        w = k - m + 1
        n_seq = 0 if offs is None else offs.numel() - 1
        nwin = n if win is None else win[1] - win[0]
        cap = int(nwin * 2.0 / (w + 1) * 1.35) + 3 * n_seq + 65536
        v = torch.empty(cap, dtype=torch.int64, device=dev); mm = torch.empty_like(v)
        sz = torch.empty(cap, dtype=torch.uint8, device=dev); ps = torch.empty_like(sz)
        ms, cnt = [], 0
        for i in range(2 + steps):
            flush.zero_()
            cnt = ctx.super_kmers_device(d_bases.data_ptr(), n, k, m, 42, True, v.data_ptr(), mm.data_ptr(), sz.data_ptr(), ps.data_ptr(),
                                         cap, d_offsets=offs.data_ptr() if offs is not None else 0, n_seq=n_seq, win_range=win)
            if i >= 2:
                ms.append(ctx.stats()["kernel_ms"])
        torch.cuda.synchronize()
        del v, mm, sz, ps
        return float(np.median(ms)), cnt

    def report(name, k, m, bases_rank, ms, records, extra, sample):
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        c = torch.tensor([records, bases_rank], dtype=torch.int64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(c, op=dist.ReduceOp.SUM)
        if rank == 0:
            total_bases, total_rec = int(c[1]), int(c[0])
            ms_max = float(t[0])
            line = {"config": name, "k": k, "m": m, "n_gpus": world, "bases": total_bases, "records": total_rec, "kernel_ms_max_over_ranks": ms_max,
                    "Gbases_per_s": total_bases / ms_max / 1e6, "algorithmic_bytes_per_base": (total_bases + 18 * total_rec) / total_bases,
                    "achieved_GBps_per_gpu": (total_bases + 18 * total_rec) / world / ms_max / 1e6, **extra}
            if sample is not None:
                v, kind, th = cpu_arm(sample, k, m)
                line["cpu_baseline"] = {"value": v, "unit": "Gbases/s", "cores": th, "kind": kind, "sample": f"{th} threads x {sample[0].size / 1e6:.0f} Mbp of the same input"}
            print(json.dumps(line), flush=True)

    threads = os.cpu_count() or 1

    # ---- configs[1]: k=31 m=19, one 100 Mbp sequence, chunk + halo across ranks (strong scaling here) --------------
    if not only or "2" in only:
        n, k, m = 100_000_000, 31, 19
        pl = shard.plan_chunks(n, k, m, world)[rank]
        s = synth.generate(42, n)[pl.buf_lo:pl.buf_hi]
        d = torch.from_numpy(s).to(dev)
        ms, cnt = timed(d, s.size, k, m, win=pl.local_window)
        report("2: minimizer/super-k-mer k=31 m=19, 100 Mbp, chunk+halo", k, m, pl.win_hi - pl.win_lo, ms, cnt, {"scaling": "strong"},
               [synth.generate(1042 + i, 4_000_000) for i in range(threads)] if rank == 0 else None)
        del d

    # ---- configs[2]: 150 bp reads x 100 M with N runs, batches of 4 M reads dealt round-robin to the ranks ----------
    if not only or "3" in only:
        k, m, L, per_batch, batches = 31, 19, 150, 4_000_000, 25
        offs = torch.arange(0, per_batch + 1, dtype=torch.int64, device=dev) * L
        ms_sum, rec_sum, bases_sum, sample = 0.0, 0, 0, None
        for bi in range(rank, batches, world):
            s = synth.generate_sharded_parallel(43 + 100 * bi, per_batch * L, 1 << 24)
            synth.inject_n_runs(s, L, seed=44 + bi)
            d = torch.from_numpy(s).to(dev)
            ms, cnt = timed(d, s.size, k, m, offs=offs, steps=3)
            ms_sum += ms; rec_sum += cnt; bases_sum += s.size
            if sample is None and rank == 0:
                sample = [s[i * 3_000_000:(i + 1) * 3_000_000].copy() for i in range(threads)]
            del d
        report("3: super-k-mers k=31 m=19, 100 M reads x 150 bp with N runs, 25 batches of 4 M reads", k, m, bases_sum, ms_sum, rec_sum,
               {"scaling": "strong", "note": "cpu sample ignores read boundaries (rate only)"}, sample)

    # ---- configs[3]: k=63 m=31, 3 Gbp genome, chunk + halo across ranks (strong scaling) ----------------------------
    if not only or "4" in only:
        n, k, m, shard_len = 3_000_000_000, 63, 31, 1 << 26
        pl = shard.plan_chunks(n, k, m, world)[rank]
        first = pl.buf_lo // shard_len
        length = pl.buf_hi - first * shard_len
        s = synth.generate_sharded_parallel(45 + first, length, shard_len)[pl.buf_lo - first * shard_len:]
        d = torch.from_numpy(s).to(dev)
        ms, cnt = timed(d, s.size, k, m, win=pl.local_window)
        report("4: super-k-mers k=63 m=31, 3 Gbp genome, chunk+halo", k, m, pl.win_hi - pl.win_lo, ms, cnt, {"scaling": "strong"},
               [s[i * 3_000_000:(i + 1) * 3_000_000].copy() for i in range(threads)] if rank == 0 else None)
        del d

    # ---- configs[4]: k=21 m=11, 20 000 mutated long reads (10-100 kbp), sharded by read ------------------------------
    if not only or "5" in only:
        k, m = 21, 11
        # read i = sequence_mutator(source slice i, 0.10, 0.3, 0.3) from cbegin(47 + i); every rank builds the
        # whole batch (deterministic) and takes its read range
        s_all, offs_h = synth.mutated_long_reads(20_000)
        a, b = shard.plan_sequences(offs_h, world)[rank]
        base, end = int(offs_h[a]), int(offs_h[b])
        s = s_all[base:end]
        d = torch.from_numpy(s).to(dev)
        offs = torch.from_numpy((offs_h[a:b + 1] - np.uint64(base)).view(np.int64)).to(dev)
        ms, cnt = timed(d, s.size, k, m, offs=offs)
        report("5: minimizers+super-k-mers k=21 m=11, 20 000 long reads 10-100 kbp, sequence_mutator(0.10, 0.3, 0.3), sharded by read", k, m, s.size, ms, cnt,
               {"scaling": "strong", "reads": int(b - a)}, [s[i * 3_000_000:(i + 1) * 3_000_000].copy() for i in range(threads)] if rank == 0 else None)
        del d

    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
