#!/usr/bin/env python
"""bench.py -- super-k-mer extraction throughput (k=31, m=19, canonical, seed 42), BASELINE.json's metric.

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--no-extra]

A "step" is one pass of the hot path (blg_super_kmers_range through the C ABI) over one batch of synthetic input:
per GPU 100 Mbp of the reference generator's stream (BASELINE configs[1]). With N > 1 (torchrun, one rank per
GPU) rank r owns windows [r*100M, (r+1)*100M) of one N*100 Mbp genome built from shards uniform(42 + shard), plus
the k+w-2 halo from the next shard: weak scaling, no data-path collective, one NCCL all-reduce of the counts.

  value         Gbases/s, whole job, inputs resident in HBM, CUDA events on the launching stream around each step (the K steps
                are enqueued with BLG_F_ASYNC, the record count is collected after the timed region), max over ranks
  e2e           the same call with HOST (pinned) buffers and the smallest record layout (BLG_F_RUNS | BLG_F_TIGHT, 6 B per record
                + a run list): H2D of the bases and D2H of the records inside the timing; e2e_runs10 = BLG_F_RUNS alone (10 B),
                e2e_compact14 = BLG_F_POS32 (14 B), e2e_records18 = the default 18-byte layout; e2e_packed_input(_tight) = the
                same fed 2-bit packed bases
  roofline      algorithmic bytes (1 B/base + 18 B/record) / kernel time vs the measured HBM copy peak
  roofline_int  why the HBM fraction is what it is: executed instructions per base and pipe utilisation (from the committed
                ncu capture, with its git sha), hashes per second of this run against the hash rate the machine sustains
                (measured live with the minimizer_position_extractor kernel, which does nothing but hash)
  parity        the records of the timed input are compared with the oracle on a prefix, on the same run
  cpu_baseline  reference read_base()+hash64 (verbatim, oracle/_ref) + restated windowing on the host cores, same input
  extra         the other configurations north_star names, device-resident, at this N: config 4 (k=63 m=31, 3 Gbp genome,
                chunk + halo, STRONG scaling) and config 3 (150-bp reads with N runs, 32 batches dealt to the ranks)

--impl reference times that CPU arm alone (all host threads) on the same input.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

K, M, SEED, CANON = 31, 19, 42, True
W = K - M + 1
SHARD = 100_000_000          # bases per GPU per step (configs[1])
GEN_SEED0 = 42               # shard i of the genome = sequence_generator uniform(42 + i)
HALO = K + W - 2             # bases of the next shard a rank needs: its last record can reach k + w - 2 bases past its last start
HALO_PAD = 64                # ... rounded up so that the slice keeps the 16-byte alignment the vector loads want
REC_BYTES = 18               # value u64 + mm_start u64 + size u8 + mm_pos u8
METRIC = "Gbases/s super-k-mer extraction (k=31,m=19)"
PARITY_PREFIX = 2_000_000    # bases of the timed input checked against the oracle on the timed run


def load_json(name):
    try:
        with open(os.path.join(ROOT, "profiles", name)) as f:
            return json.load(f)
    except Exception:
        return None


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def host_description():
    """CPU model / core count of the box, for the CPU baseline."""
    model = "?"
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                model = line.split(":", 1)[1].strip()
                break
    except Exception:
        pass
    return {"cpu": model, "logical_cpus": os.cpu_count() or 1,
            "build": "g++ -std=c++17 -O3 -DNDEBUG -fno-lifetime-dse (oracle/Makefile): the reference headers compiled where they lie"}


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons during the timed region (NVML, ~5 ms period)."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.sm, self.reasons, self.max_mhz = index, False, [], set(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self.stop_flag:
            try:
                self.sm.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                break
            time.sleep(0.005)

    def summary(self):
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


# ----------------------------------------------------------------------------------------------- CPU arm
def cpu_engine():
    """(callable bases -> record count, kind, description): the reference's own arithmetic when oracle/_ref travelled here."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle
    if os.path.exists(pyoracle.REF_SO):
        eng = pyoracle.Ref()
        return (lambda s: eng.super_kmers_sums(s, K, M, SEED, CANON)[0]), "reference", \
            "reference read_base()+hash64 verbatim (oracle/_ref) + restated deque windowing"
    eng = pyoracle.Oracle()
    return (lambda s: eng.super_kmers(s, K, M, SEED, CANON, want=())["n"]), "port", "oracle.c port (rolling words + Murmur + deque windowing)"


def shard_chunks(total_bases: int, threads: int):
    """The bench's own input -- the first `total_bases` of shard 0, uniform(42) -- cut into one chunk per thread with the
    k + w - 2 overlap a record can span, so that the threads together produce exactly the records of those bases."""
    from biolib_b200 import synth
    s = synth.generate(GEN_SEED0, min(total_bases + HALO, SHARD))
    per = (total_bases + threads - 1) // threads
    return [s[i * per: min((i + 1) * per + HALO, s.size)] for i in range(threads) if i * per < total_bases]


def cpu_arm(chunks, repeats: int = 1):
    """All chunks concurrently, one thread each. Returns (Gbases/s over the owned bases, kind, description, seconds)."""
    fn, kind, what = cpu_engine()
    best = None
    for _ in range(repeats):
        ths = [threading.Thread(target=fn, args=(c,)) for c in chunks]
        t0 = time.perf_counter()
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    total = sum(c.size for c in chunks) - HALO * len(chunks)
    return total / best / 1e9, kind, what, best


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    steps, warm = max(args.steps, 1), max(args.warmup, 0)
    # the GPU arm's own step: the 100 Mbp shard of rank 0 (a step of it is 0.1 - 0.2 s on 16 - 32 host threads); bounded so
    # that K + W steps stay near two minutes
    total = int(min(SHARD, max(4_000_000, 120.0 * 25e6 * threads / (steps + warm))))
    chunks = shard_chunks(total, threads)
    vals = []
    kind = what = None
    for i in range(warm + steps):
        v, kind, what, _ = cpu_arm(chunks)
        if i >= warm:
            vals.append(v)
    v = float(np.mean(vals))
    one, _, _, sec1 = cpu_arm(shard_chunks(4_000_000, 1))
    sample = f"{what}; {len(chunks)} threads over the first {total / 1e6:.0f} Mbp of the GPU arm's shard (uniform(42)), chunk + halo"
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "Gbases/s", "n_gpus": args.gpus, "steps": steps,
            "warmup": warm, "ms_per_step": total / (v * 1e9) * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64", "data": "synthetic",
            "config": {"workload": "super_kmer_view k=31 m=19 canonical seed=42 over 100 Mbp/GPU of sequence_generator "
                                   "uniform(42+shard) (BASELINE configs[1])",
                       "k": K, "m": M, "bases_per_step": total, "same_input_as_gpu_arm": True},
            "cpu_baseline": {"value": v, "unit": "Gbases/s", "cores": len(chunks), "kind": kind, "sample": sample,
                             "one_core_value": one, "host": host_description()},
            "e2e": {"value": v, "unit": "Gbases/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------- GPU arm
def np_from_ptr(ptr: int, nbytes: int, dtype) -> np.ndarray:
    import ctypes as C
    buf = (C.c_uint8 * nbytes).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype)


def build_rank_input(rank: int, world: int):
    """Rank r's slice of the world*SHARD genome: [r*SHARD - LH, (r+1)*SHARD + HALO + pad) clipped to the genome.
    Returns (bases uint8, win_lo, win_hi) with the window range in slice coordinates."""
    from biolib_b200 import synth
    LH = 64 if rank > 0 else 0
    RH = HALO + HALO_PAD if rank < world - 1 else 0
    parts = []
    if LH:
        parts.append(synth.generate(GEN_SEED0 + rank - 1, SHARD)[-LH:])
    parts.append(synth.generate(GEN_SEED0 + rank, SHARD))
    if RH:
        parts.append(synth.generate(GEN_SEED0 + rank + 1, RH))
    bases = np.concatenate(parts) if len(parts) > 1 else parts[0]
    win_lo = LH
    win_hi = LH + SHARD if rank < world - 1 else LH + SHARD - K + 1
    return bases, win_lo, win_hi


def bind_to_gpu_numa_node(local: int) -> str:
    """Run this rank (and so first-touch its pinned buffers) on the CPU cores of the NUMA node its GPU hangs off:
    with 8 ranks the host<->device copies otherwise cross the socket interconnect. Best effort, never fatal."""
    try:
        import torch
        pr = torch.cuda.get_device_properties(local)
        bdf = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        base = f"/sys/bus/pci/devices/{bdf}"
        node = int(open(base + "/numa_node").read())
        if node < 0:
            return "numa: none"
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
        return f"numa node {node}, {len(cpus)} cpus"
    except Exception as e:  # noqa: BLE001
        return f"numa: unavailable ({type(e).__name__})"


class DeviceRun:
    """Device-resident runs of blg_super_kmers_range for one (k, m): output arrays sized once, kernel time from the
    library's CUDA events with the device idle at launch, L2 flushed between runs."""

    def __init__(self, torch, ctx, dev, flush, k, m):
        self.torch, self.ctx, self.dev, self.flush, self.k, self.m = torch, ctx, dev, flush, k, m
        self.cap = 0

    def ensure(self, cap):
        if cap > self.cap:
            t = self.torch
            self.v = t.empty(cap, dtype=t.int64, device=self.dev); self.mm = t.empty(cap, dtype=t.int64, device=self.dev)
            self.sz = t.empty(cap, dtype=t.uint8, device=self.dev); self.ps = t.empty(cap, dtype=t.uint8, device=self.dev)
            self.cap = cap

    def call(self, d_bases, n, win=None, offs=None, async_call=False):
        n_seq = 0 if offs is None else offs.numel() - 1
        return self.ctx.super_kmers_device(d_bases.data_ptr(), n, self.k, self.m, SEED, CANON, self.v.data_ptr(), self.mm.data_ptr(),
                                           self.sz.data_ptr(), self.ps.data_ptr(), self.cap, win_range=win,
                                           d_offsets=offs.data_ptr() if offs is not None else 0, n_seq=n_seq, async_call=async_call)

    def kernel_ms(self, d_bases, n, win=None, offs=None, reps=5, warm=1):
        """(median kernel ms, record count)"""
        nwin = n if win is None else win[1] - win[0]
        w = self.k - self.m + 1
        self.ensure(int(nwin * 2.0 / (w + 1) * 1.35) + 3 * (0 if offs is None else offs.numel()) + 65536)
        ms, cnt = [], 0
        for i in range(warm + reps):
            self.flush.zero_()
            self.torch.cuda.synchronize()
            cnt = self.call(d_bases, n, win, offs)
            if i >= warm:
                ms.append(self.ctx.stats()["kernel_ms"])
        return float(np.median(ms)), cnt


def hash_rate_ceiling(torch, ctx, dev):
    """Hashes per second of a kernel that only hashes: hash::minimizer_position_extractor (blg_minimizer_positions) over
    2^24 k-mer values, w = 13 hashes each, device-resident."""
    import ctypes as C
    n = 1 << 24
    vals = torch.randint(0, 2**62, (n,), dtype=torch.int64, device=dev)
    out = torch.empty(n, dtype=torch.uint8, device=dev)
    best = None
    for _ in range(4):
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        st = ctx.L.blg_minimizer_positions(ctx.h, C.c_void_p(vals.data_ptr()), n, K, M, 3, C.c_void_p(out.data_ptr()))
        b.record()
        torch.cuda.synchronize()
        assert st == 0
        ms = a.elapsed_time(b)
        best = ms if best is None else min(best, ms)
    return n * W / (best * 1e-3)


def check_parity(ctx, bases_h, win_lo, d_value, d_mm, d_size, count):
    """The records of the timed run against the oracle on the first PARITY_PREFIX owned bases (rank 0 of the run)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle
    orc = pyoracle.Oracle()
    pre = bases_h[win_lo: win_lo + PARITY_PREFIX]
    o = orc.super_kmers(pre, K, M, SEED, CANON)
    keep = int(o["n"]) - 2                       # the last records of a prefix are cut short by its end
    gv = d_value[:keep].cpu().numpy().view(np.uint64)
    gm = d_mm[:keep].cpu().numpy().view(np.uint64) - np.uint64(win_lo)
    gs = d_size[:keep].cpu().numpy()
    ok = keep > 0 and keep <= count and np.array_equal(gv, o["value"][:keep]) and np.array_equal(gm, o["mm_start"][:keep]) \
        and np.array_equal(gs, o["size"][:keep])
    return bool(ok), keep


def run_extras(torch, dist, ctx, dev, flush, world, rank):
    """The other configurations north_star names, device-resident, at this N (STRONG scaling: fixed total work)."""
    from biolib_b200 import shard, synth

    def reduce(ms, records, bases):
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        c = torch.tensor([records, bases], dtype=torch.int64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(c, op=dist.ReduceOp.SUM)
        return float(t[0]), int(c[0]), int(c[1])

    out = {}
    # ---- config 4: k=63 m=31 over a 3 Gbp genome of shards uniform(45 + i) of 64 Mbp, chunk + halo ----------------------
    n, k, m, shard_len = 3_000_000_000, 63, 31, 1 << 26
    pl = shard.plan_chunks(n, k, m, world)[rank]
    first = pl.buf_lo // shard_len
    s = synth.generate_sharded_parallel(45 + first, pl.buf_hi - first * shard_len, shard_len)[pl.buf_lo - first * shard_len:]
    d = torch.from_numpy(s).to(dev)
    run = DeviceRun(torch, ctx, dev, flush, k, m)
    ms, cnt = run.kernel_ms(d, s.size, win=pl.local_window, reps=3)
    ms, rec, bases = reduce(ms, cnt, pl.win_hi - pl.win_lo)
    out["config4_k63_m31_3gbp"] = {"scaling": "strong", "n_gpus": world, "bases": bases, "records": rec, "kernel_ms_max_over_ranks": ms,
                                   "Gbases_per_s": bases / ms / 1e6, "achieved_GBps_per_gpu": (bases + REC_BYTES * rec) / world / ms / 1e6,
                                   "sharding": "chunk + k+w-2 halo, records owned by the rank of their first k-mer"}
    del d, s, run
    torch.cuda.empty_cache()
    # ---- config 3: 150-bp reads with N runs: 32 batches of 1 M reads dealt round-robin to the ranks ----------------------
    k, m, L, per_batch, batches = 31, 19, 150, 1_000_000, 32
    offs = torch.arange(0, per_batch + 1, dtype=torch.int64, device=dev) * L
    run = DeviceRun(torch, ctx, dev, flush, k, m)
    ms_sum, rec_sum, bases_sum = 0.0, 0, 0
    for bi in range(rank, batches, world):
        s = synth.generate_sharded_parallel(43 + 100 * bi, per_batch * L, 1 << 24)
        synth.inject_n_runs(s, L, seed=44 + bi)
        d = torch.from_numpy(s).to(dev)
        ms, cnt = run.kernel_ms(d, s.size, offs=offs, reps=2)
        ms_sum += ms; rec_sum += cnt; bases_sum += s.size
        del d
    ms, rec, bases = reduce(ms_sum, rec_sum, bases_sum)
    out["config3_reads150_nruns"] = {"scaling": "strong", "n_gpus": world, "bases": bases, "records": rec, "kernel_ms_max_over_ranks": ms,
                                     "Gbases_per_s": bases / ms / 1e6, "reads": batches * per_batch,
                                     "note": "32 batches of 1 M reads (4.8 Gbp) of the 100 M-read configuration, 32 / N batches per rank"}
    del run
    torch.cuda.empty_cache()
    # ---- the other views of the path on this rank's 100 Mbp shard, device-resident: kmer_view (the HBM-bound kernel) and the closed
    # syncmer filter (every rank runs its own shard: weak scaling, like the headline)
    peak, _ = load_peaks()
    bases_h, _, _ = build_rank_input(rank, world)
    nb = min(bases_h.size, SHARD)
    d = torch.from_numpy(bases_h[:nb]).to(dev)
    d_lo = torch.empty(nb, dtype=torch.int64, device=dev)
    d_pos = torch.empty(nb, dtype=torch.int64, device=dev)
    for name, kw in (("kmer_view_k31", {}), ("syncmers_k31_m19", {"syncmer": (19, 0, 12)})):
        ms, cnt = [], 0
        for i in range(4):
            flush.zero_()
            torch.cuda.synchronize()
            cnt = ctx.kmers_device(d.data_ptr(), nb, 31, True, d_lo.data_ptr(), 0, d_pos.data_ptr(), nb, **kw)
            if i:
                ms.append(ctx.stats()["kernel_ms"])
        t, rec, bases = reduce(float(np.median(ms)), cnt, nb)
        algo = (nb + 16 * cnt) / 1e9
        out[name] = {"scaling": "weak", "n_gpus": world, "bases": bases, "records": rec, "kernel_ms_max_over_ranks": t, "Gbases_per_s": bases / t / 1e6,
                     "achieved_GBps_per_gpu": algo / (float(np.median(ms)) * 1e-3), "frac_of_hbm_peak": algo / (float(np.median(ms)) * 1e-3) / peak,
                     "algorithmic_bytes": "1 B per base + 16 B per (kept) k-mer"}
    del d, d_lo, d_pos
    torch.cuda.empty_cache()
    # ---- the reference's downstream pipeline (tests/test_jaccard.cpp: k-mers -> sort -> unique -> Jaccard) over two related genomes of
    # 200 Mbp, cut over the ranks by k-mer start: local sort + unique, all-to-all by key range over NVLink, merge, per-range Jaccard
    from biolib_b200 import multi_gpu as mg
    n, k = 200_000_000, 31
    lo, hi = mg.kmer_chunk(n, k, world, rank)
    s1, s2 = synth.related_genomes_slice(lo, hi)
    d1, d2 = torch.from_numpy(s1).to(dev), torch.from_numpy(s2).to(dev)
    ops = mg.CudaKmerOps(ctx, torch, dev)
    best, res = None, None
    for _ in range(3):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        res = mg.distributed_jaccard(ops, dist if world > 1 else None, d1, d2, k, True)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) * 1e3
        best = dt if best is None else min(best, dt)
    ms, _, _ = reduce(best, 0, 0)
    out["jaccard_k31_two_200mbp_genomes"] = {"scaling": "strong", "n_gpus": world, "kmers": 2 * (n - k + 1), "ms_max_over_ranks": ms,
                                             "G_kmers_per_s": 2 * (n - k + 1) / ms / 1e6, "intersection": res["intersection"], "union": res["union"],
                                             "bytes_over_nvlink": res["bytes_exchanged"],
                                             "timing": "wall clock around the whole call (extraction, sorts, splitter agreement, all-to-all, merge, all-reduce), best of 3"}
    return out


def run_gpu(args):
    import torch
    import torch.distributed as dist
    import biolib_b200
    from biolib_b200 import _capi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and rank == 0:
        print(f"warning: --gpus {args.gpus} but WORLD_SIZE={world}", file=sys.stderr)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = bind_to_gpu_numa_node(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    ctx = biolib_b200.Context(local)
    bases_h, win_lo, win_hi = build_rank_input(rank, world)
    n = bases_h.size
    nwin = win_hi - win_lo
    cap = int(nwin * 2.0 / (W + 1) * 1.35) + 65536

    d_bases = torch.from_numpy(bases_h).to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    # a stream of our own: torch's default stream is the legacy stream 0, which blg_set_stream reads as "the context's own
    # stream" -- flush, events and kernels must share one real stream to be ordered
    torch.cuda.synchronize()
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    ctx.set_stream(stream.cuda_stream)
    main_run = DeviceRun(torch, ctx, dev, flush, K, M)
    main_run.ensure(cap)
    d_value, d_mm, d_size = main_run.v, main_run.mm, main_run.sz

    def step_device():
        return main_run.call(d_bases, n, win=(win_lo, win_hi))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    warm = max(args.warmup, 3)
    for _ in range(warm):
        flush.zero_()
        count = step_device()
    sampler = ClockSampler(local)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    exact_tiles = 0
    barrier()
    sampler.start()
    # a few milliseconds of device-side delay in front of the timed steps: the host enqueues the K steps while it runs, so that a
    # launch that arrives late (eight ranks and their clock samplers share the box's cores) does not sit between a step's two events
    try:
        with torch.cuda.stream(stream):
            torch.cuda._sleep(int(4e-3 * 1.9e9))
    except Exception:
        for _ in range(16):
            flush.zero_()
    for s, e in ev:
        flush.zero_()           # L2 flush between timed iterations (outside the event pair)
        s.record(stream)
        main_run.call(d_bases, n, win=(win_lo, win_hi), async_call=True)   # BLG_F_ASYNC: enqueue only, as a pipeline would
        e.record(stream)
    barrier()
    sampler.stop_flag = True
    sampler.join()
    count_async = ctx.async_count()                                            # the last step's record count, after the fact
    assert count_async == count, (count_async, count)
    exact_tiles = ctx.stats()["exact_tiles"] // max(args.steps, 1)
    step_ms = sum(s.elapsed_time(e) for s, e in ev) / args.steps
    kernel_ms, count_k = main_run.kernel_ms(d_bases, n, win=(win_lo, win_hi), reps=min(max(args.steps, 3), 10))
    assert count_k == count
    parity_ok, parity_records = check_parity(ctx, bases_h, win_lo, d_value, d_mm, d_size, count) if rank == 0 else (True, 0)
    hash_ceiling = hash_rate_ceiling(torch, ctx, dev) if rank == 0 else 0.0

    # ---- end to end: host (pinned) buffers in and out through the same entry point, both record layouts ----
    ctx.set_stream(None)
    e2e_steps = max(3, min(args.steps, 10))
    hp_in = ctx.host_alloc(n)
    np_from_ptr(hp_in, n, np.uint8)[:] = bases_h
    hp = [ctx.host_alloc(cap * 8), ctx.host_alloc(cap * 8), ctx.host_alloc(cap), ctx.host_alloc(cap)]
    run_cap = cap // 64 + 65536
    hp_runs = [ctx.host_alloc(run_cap * 8), ctx.host_alloc(run_cap * 8)]
    e2e = {}
    TIGHT = _capi.F_RUNS | _capi.F_TIGHT
    for name, flags in (("records18", 0), ("compact14", _capi.F_POS32), ("runs10", _capi.F_RUNS), ("tight6", TIGHT)):
        def step_host():
            return ctx.super_kmers_raw(hp_in, n, K, M, SEED, CANON, hp[0], hp[1], hp[2], hp[3], cap, flags=flags, win_range=(win_lo, win_hi),
                                       run_ptrs=hp_runs if flags & _capi.F_RUNS else None, run_capacity=run_cap if flags & _capi.F_RUNS else 0,
                                       tight_ptrs=(hp[0], hp[2], hp[3]) if flags & _capi.F_TIGHT else None)
        for _ in range(2):
            c2 = step_host()
        assert c2 == count, (c2, count)
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            step_host()
            st = ctx.stats()
        torch.cuda.synchronize()
        e2e[name] = ((time.perf_counter() - t0) * 1e3 / e2e_steps, st["bytes_h2d"], st["bytes_d2h"])
        # parity of the host path: same records as the device-resident run (checksums over what landed in host memory)
        from biolib_b200.context import expand_runs, unpack_tight
        if flags & _capi.F_TIGHT:
            hv, h_size, h_pos = unpack_tight(np_from_ptr(hp[0], cap * 4, np.uint32)[:count], np_from_ptr(hp[2], cap, np.uint8)[:count],
                                             np_from_ptr(hp[3], cap, np.uint8)[:count])
        else:
            hv = np_from_ptr(hp[0], cap * 8, np.uint64)[:count]
            h_size, h_pos = np_from_ptr(hp[2], cap, np.uint8)[:count], np_from_ptr(hp[3], cap, np.uint8)[:count]
        assert int(hv.sum(dtype=np.uint64)) == int(d_value[:count].sum().item()) & (2**64 - 1)
        if flags & _capi.F_RUNS:
            nr = ctx.last_n_runs
            pos = expand_runs(np_from_ptr(hp_runs[0], run_cap * 8, np.uint64)[:nr], np_from_ptr(hp_runs[1], run_cap * 8, np.uint64)[:nr], h_size, h_pos)
        else:
            pos = np_from_ptr(hp[1], cap * 8, np.uint32 if flags else np.uint64)[:count].astype(np.uint64) + np.uint64(win_lo if flags else 0)
        assert int(pos.sum(dtype=np.uint64)) == int(d_mm[:count].sum().item()) & (2**64 - 1)
    # the same call fed the 2-bit wire format (blg_super_kmers_packed: 0.25 B per base over PCIe, unpacked on the device piece by piece);
    # the packing itself is outside the timing: it is the format a pipeline would hold its reads in
    codes, bs, bl = ctx.pack_bases(bases_h)
    hp_codes = ctx.host_alloc(codes.size * 4)
    np_from_ptr(hp_codes, codes.size * 4, np.uint32)[:] = codes

    def step_packed():
        return ctx.super_kmers_packed_raw(hp_codes, n, bs.ctypes.data if bs.size else 0, bl.ctypes.data if bl.size else 0, bs.size, K, M, SEED, CANON,
                                          hp[0], 0, hp[2], hp[3], cap, flags=_capi.F_RUNS, win_range=(win_lo, win_hi), run_ptrs=hp_runs, run_capacity=run_cap)
    for _ in range(2):
        c2 = step_packed()
    assert c2 == count, (c2, count)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_packed()
        st = ctx.stats()
    torch.cuda.synchronize()
    e2e["packed_in"] = ((time.perf_counter() - t0) * 1e3 / e2e_steps, st["bytes_h2d"], st["bytes_d2h"])
    assert int(np_from_ptr(hp[0], cap * 8, np.uint64)[:count].sum(dtype=np.uint64)) == int(d_value[:count].sum().item()) & (2**64 - 1)

    def step_packed_tight():
        return ctx.super_kmers_packed_raw(hp_codes, n, bs.ctypes.data if bs.size else 0, bl.ctypes.data if bl.size else 0, bs.size, K, M, SEED, CANON,
                                          0, 0, 0, 0, cap, flags=TIGHT, win_range=(win_lo, win_hi), run_ptrs=hp_runs, run_capacity=run_cap,
                                          tight_ptrs=(hp[0], hp[2], hp[3]))
    for _ in range(2):
        c2 = step_packed_tight()
    assert c2 == count, (c2, count)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_packed_tight()
        st = ctx.stats()
    torch.cuda.synchronize()
    e2e["packed_tight"] = ((time.perf_counter() - t0) * 1e3 / e2e_steps, st["bytes_h2d"], st["bytes_d2h"])
    hv, _, _ = unpack_tight(np_from_ptr(hp[0], cap * 4, np.uint32)[:count], np_from_ptr(hp[2], cap, np.uint8)[:count], np_from_ptr(hp[3], cap, np.uint8)[:count])
    assert int(hv.sum(dtype=np.uint64)) == int(d_value[:count].sum().item()) & (2**64 - 1)
    for p in [hp_in, hp_codes] + hp + hp_runs:
        ctx.host_free(p)

    ctx.set_stream(stream.cuda_stream)
    extras = None if args.no_extra else run_extras(torch, dist, ctx, dev, flush, world, rank)

    t = torch.tensor([step_ms, e2e["tight6"][0], kernel_ms, e2e["records18"][0], e2e["compact14"][0], e2e["packed_in"][0], e2e["runs10"][0],
                      e2e["packed_tight"][0]], dtype=torch.float64, device=dev)
    cnt = torch.tensor([count, nwin, exact_tiles], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)   # the path's only collective: per-GPU counts
    step_ms, e2e_ms, kernel_ms, e2e18_ms, e2e14_ms, e2ep_ms, e2e10_ms, e2ept_ms = t.tolist()
    total_records, total_win, total_exact = cnt.tolist()
    total_bases = SHARD * world

    if rank == 0:
        peak, peak_kind = load_peaks()
        traffic = load_json("ncu_traffic.json") or {}
        intmix = load_json("ncu_intmix.json") or {}
        algo_bytes = nwin + REC_BYTES * count          # this rank's launch
        achieved = algo_bytes / (kernel_ms * 1e-3) / 1e9
        hashes = nwin * 64.0 / 61.0                     # sk4 hashes 64 blocks of m-mer starts per 61 blocks of new windows
        line = {
            "metric": METRIC, "value": total_bases / (step_ms * 1e-3) / 1e9, "unit": "Gbases/s", "n_gpus": world,
            "steps": args.steps, "warmup": warm, "ms_per_step": step_ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": {"workload": "super_kmer_view k=31 m=19 canonical seed=42 over 100 Mbp/GPU of sequence_generator "
                                   "uniform(42+shard) (BASELINE configs[1])",
                       "k": K, "m": M, "bases_per_gpu": SHARD, "records": total_records, "kmers": total_win,
                       "l2": "256 MiB flush write between timed steps", "sharding": f"chunk+halo x{world}",
                       "host": numa, "e2e_layout": "BLG_F_RUNS | BLG_F_TIGHT: value in 5 bytes (u32 + u8 arrays) + one byte holding size - 1 and mm_pos per "
                                     "record, plus (first record, start) per run of contiguous super-k-mers; positions follow by adding sizes, as in "
                                     "the reference's own sequential view. ASCII bases in."},
            "e2e": {"value": total_bases / (e2e_ms * 1e-3) / 1e9, "unit": "Gbases/s", "h2d_bytes_per_step": int(e2e["tight6"][1]),
                    "d2h_bytes_per_step": int(e2e["tight6"][2]), "steps": e2e_steps, "ms_per_step": e2e_ms},
            "e2e_runs10": {"value": total_bases / (e2e10_ms * 1e-3) / 1e9, "unit": "Gbases/s", "h2d_bytes_per_step": int(e2e["runs10"][1]),
                           "d2h_bytes_per_step": int(e2e["runs10"][2]), "ms_per_step": e2e10_ms,
                           "note": "BLG_F_RUNS alone: value u64 + size u8 + mm_pos u8 per record (the round's earlier headline layout)"},
            "e2e_packed_input_tight": {"value": total_bases / (e2ept_ms * 1e-3) / 1e9, "unit": "Gbases/s", "h2d_bytes_per_step": int(e2e["packed_tight"][1]),
                                       "d2h_bytes_per_step": int(e2e["packed_tight"][2]), "ms_per_step": e2ept_ms,
                                       "note": "2-bit packed host input and 6-byte records: the fewest bytes over PCIe in both directions"},
            "e2e_compact14": {"value": total_bases / (e2e14_ms * 1e-3) / 1e9, "unit": "Gbases/s", "h2d_bytes_per_step": int(e2e["compact14"][1]),
                              "d2h_bytes_per_step": int(e2e["compact14"][2]), "ms_per_step": e2e14_ms},
            "e2e_packed_input": {"value": total_bases / (e2ep_ms * 1e-3) / 1e9, "unit": "Gbases/s", "h2d_bytes_per_step": int(e2e["packed_in"][1]),
                                 "d2h_bytes_per_step": int(e2e["packed_in"][2]), "ms_per_step": e2ep_ms,
                                 "note": "blg_super_kmers_packed: 2-bit packed host input (packed outside the timing), 10-byte records + run list"},
            "e2e_records18": {"value": total_bases / (e2e18_ms * 1e-3) / 1e9, "unit": "Gbases/s", "h2d_bytes_per_step": int(e2e["records18"][1]),
                              "d2h_bytes_per_step": int(e2e["records18"][2]), "ms_per_step": e2e18_ms},
            "gpu_launches": args.steps,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "frac_of_nominal_8000": achieved / 8000.0,      # north_star's ~8 TB/s figure (report against both)
                         "traffic": traffic.get("dram_bytes_per_launch"), "traffic_capture": {"git_sha": traffic.get("git_sha"), "kernel": traffic.get("kernel")},
                         "peak_kind": peak_kind, "kernel": "sk4_kernel", "kernel_ms": kernel_ms,
                         "algorithmic_bytes_per_base": algo_bytes / nwin},
            "roofline_int": {"bound": "integer issue (ALU pipe: LOP3 / SHF / ISETP / SEL / IADD3 / VIMNMX at 16 lanes per clock per scheduler)",
                             "inst_per_base": intmix.get("inst_per_base"), "alu_ops_per_base": intmix.get("alu_ops_per_base"),
                             "imad_per_base": intmix.get("imad_per_base"), "alu_pipe_pct": intmix.get("alu_pipe_pct"),
                             "fma_pipe_pct": intmix.get("fma_pipe_pct"), "issue_slots_pct": intmix.get("issue_slots_pct"),
                             "capture": {"git_sha": intmix.get("git_sha"), "kernel": intmix.get("kernel")},
                             "hashes_per_s": hashes / (kernel_ms * 1e-3), "hash_rate_ceiling": hash_ceiling,
                             "hash_rate_frac": hashes / (kernel_ms * 1e-3) / hash_ceiling if hash_ceiling else None,
                             "exact_redo_tiles_per_step": int(total_exact)},
            "parity_checked": parity_ok, "parity": {"against": "oracle/oracle.c on the first 2 Mbp of the timed input, value + mm_start + size",
                                                    "records_compared": parity_records},
            "clocks": sampler.summary(),
        }
        if extras is not None:
            line["extra"] = extras
        if world == 1:
            threads = os.cpu_count() or 1
            chunks = shard_chunks(min(SHARD, 8_000_000 * threads), threads)
            v, kind, what, sec = cpu_arm(chunks)
            line["cpu_baseline"] = {"value": v, "unit": "Gbases/s", "cores": len(chunks), "kind": kind, "host": host_description(),
                                    "sample": f"{what}; {len(chunks)} threads over the first {(sum(c.size for c in chunks) - HALO * len(chunks)) / 1e6:.0f} Mbp of this "
                                              f"run's own shard, {sec:.2f} s"}
        if not parity_ok:
            line["error"] = "parity check against the oracle FAILED on the timed run"
        print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-extra", action="store_true", help="skip the config 3 / config 4 runs")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
