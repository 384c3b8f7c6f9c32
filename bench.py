#!/usr/bin/env python
"""bench.py -- super-k-mer extraction throughput (k=31, m=19, canonical, seed 42), BASELINE.json's metric.

    python bench.py --gpus N --steps K --warmup W [--impl reference]

A "step" is one pass of the hot path (blg_super_kmers_range through the C ABI) over one batch of synthetic input:
per GPU 100 Mbp of the reference generator's stream (BASELINE configs[1]). With N > 1 (torchrun, one rank per
GPU) rank r owns windows [r*100M, (r+1)*100M) of one N*100 Mbp genome built from shards uniform(42 + shard), plus
the k+w-2 halo from the next shard: weak scaling, no data-path collective, one NCCL all-reduce of the counts.

  value      Gbases/s, whole job, inputs resident in HBM, CUDA events on the launching stream, max over ranks
  e2e        the same call with HOST (pinned) buffers: H2D of the bases and D2H of the records inside the timing
  roofline   algorithmic bytes (1 B/base + 18 B/record) / kernel time vs the measured HBM copy peak
  cpu_baseline  reference read_base()+hash64 (verbatim, oracle/_ref) + restated windowing on the host cores

--impl reference times that CPU arm alone (all host threads) on the same config.
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
HALO = K + W - 2 + 64        # bases of the next shard a rank needs (rounded up)
REC_BYTES = 18               # value u64 + mm_start u64 + size u8 + mm_pos u8
METRIC = "Gbases/s super-k-mer extraction (k=31,m=19)"


def load_traffic():
    """DRAM bytes per launch of the dominant kernel from the last committed `ncu --set full` capture of this very
    workload (profiles/ncu_traffic.json: dram__bytes_read.sum + dram__bytes_write.sum); None when there is none."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            d = json.load(f)
        return d["kernel"], float(d["dram_bytes_per_launch"])
    except Exception:
        return "sk3_kernel", None


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


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
def cpu_arm(sample_bases_per_thread: int, threads: int, repeats: int = 1, chunks=None):
    """Reference arithmetic on the host cores: each thread runs the whole path over its own chunk
    (chunks overlap nothing: records at chunk edges are irrelevant for a rate). Returns (Gbases/s, kind, desc)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle
    from biolib_b200 import synth
    use_ref = os.path.exists(pyoracle.REF_SO)
    if use_ref:
        eng = pyoracle.Ref()
        fn = lambda s: eng.super_kmers_sums(s, K, M, SEED, CANON)[0]
        kind = "reference"
        what = "reference read_base()+hash64 verbatim (oracle/_ref) + restated deque windowing"
    else:
        eng = pyoracle.Oracle()
        fn = lambda s: eng.super_kmers(s, K, M, SEED, CANON, want=())["n"]
        kind = "port"
        what = "oracle.c port (rolling words + Murmur + deque windowing)"
    if chunks is None:
        chunks = [synth.generate(GEN_SEED0 + 1000 + t, sample_bases_per_thread) for t in range(threads)]
    best = None
    for _ in range(repeats):
        res = [0] * threads
        ths = [threading.Thread(target=lambda i=i: res.__setitem__(i, fn(chunks[i]))) for i in range(threads)]
        t0 = time.perf_counter()
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    total = sample_bases_per_thread * threads
    return total / best / 1e9, kind, f"{what}; {threads} threads x {sample_bases_per_thread/1e6:.0f} Mbp of the same generator stream, {best:.2f} s"


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    steps, warm = max(args.steps, 1), max(args.warmup, 0)
    from biolib_b200 import synth
    # bounded sample: the whole run stays near two minutes at ~30 Mbases/s per thread
    per_thread = int(min(8_000_000, max(500_000, 120.0 * 30e6 / (steps + warm))))
    chunks = [synth.generate(GEN_SEED0 + 1000 + t, per_thread) for t in range(threads)]
    vals = []
    desc = kind = None
    for i in range(warm + steps):
        v, kind, desc = cpu_arm(per_thread, threads, chunks=chunks)
        if i >= warm:
            vals.append(v)
    v = float(np.mean(vals))
    ms = per_thread * threads / (v * 1e9) * 1e3
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "Gbases/s", "n_gpus": args.gpus, "steps": steps,
            "warmup": warm, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64", "data": "synthetic",
            "config": {"workload": "super_kmer_view k=31 m=19 canonical seed=42, sequence_generator stream (configs[1])",
                       "k": K, "m": M, "bases_per_step": per_thread * threads},
            "cpu_baseline": {"value": v, "unit": "Gbases/s", "cores": threads, "kind": kind, "sample": desc},
            "e2e": {"value": v, "unit": "Gbases/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------- GPU arm
def np_from_ptr(ptr: int, nbytes: int, dtype) -> np.ndarray:
    import ctypes as C
    buf = (C.c_uint8 * nbytes).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype)


def build_rank_input(rank: int, world: int):
    """Rank r's slice of the world*SHARD genome: [r*SHARD - LH, (r+1)*SHARD + HALO) clipped to the genome.
    Returns (bases uint8, win_lo, win_hi) with the window range in slice coordinates."""
    from biolib_b200 import synth
    LH = 64 if rank > 0 else 0
    RH = HALO if rank < world - 1 else 0
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


def run_gpu(args):
    import torch
    import torch.distributed as dist
    import biolib_b200

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
    cap = int(nwin * 2.0 / (W + 1) * 1.3) + 65536

    d_bases = torch.from_numpy(bases_h).to(dev)
    d_value = torch.empty(cap, dtype=torch.int64, device=dev)
    d_mm = torch.empty(cap, dtype=torch.int64, device=dev)
    d_size = torch.empty(cap, dtype=torch.uint8, device=dev)
    d_pos = torch.empty(cap, dtype=torch.uint8, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)

    def step_device():
        return ctx.super_kmers_device(d_bases.data_ptr(), n, K, M, SEED, CANON, d_value.data_ptr(), d_mm.data_ptr(),
                                      d_size.data_ptr(), d_pos.data_ptr(), cap, win_range=(win_lo, win_hi))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        flush.zero_()
        count = step_device()
    sampler = ClockSampler(local)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    kernel_ms = 0.0
    barrier()
    sampler.start()
    for s, e in ev:
        flush.zero_()           # L2 flush between timed iterations (outside the event pair)
        s.record(stream)
        count = step_device()
        e.record(stream)
        kernel_ms += ctx.stats()["kernel_ms"]
    barrier()
    sampler.stop_flag = True
    sampler.join()
    step_ms = sum(s.elapsed_time(e) for s, e in ev) / args.steps
    kernel_ms /= args.steps

    # ---- end to end: host (pinned) buffers in and out through the same entry point ----
    ctx.set_stream(None)
    e2e_steps = max(3, min(args.steps, 10))
    hp_in = ctx.host_alloc(n)
    np_from_ptr(hp_in, n, np.uint8)[:] = bases_h
    hp = [ctx.host_alloc(cap * 8), ctx.host_alloc(cap * 8), ctx.host_alloc(cap), ctx.host_alloc(cap)]

    def step_host():
        return ctx.super_kmers_raw(hp_in, n, K, M, SEED, CANON, hp[0], hp[1], hp[2], hp[3], cap, win_range=(win_lo, win_hi))

    for _ in range(2):
        c2 = step_host()
    assert c2 == count, (c2, count)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_host()
        st = ctx.stats()
    torch.cuda.synchronize()
    e2e_ms = (time.perf_counter() - t0) * 1e3 / e2e_steps
    h2d, d2h = st["bytes_h2d"], st["bytes_d2h"]
    # parity of the two paths: same records (checksums over what landed in host memory vs HBM)
    hv = np_from_ptr(hp[0], cap * 8, np.uint64)[:count]
    assert int(hv.sum(dtype=np.uint64)) == int(d_value[:count].sum().item()) & (2**64 - 1)

    t = torch.tensor([step_ms, e2e_ms, kernel_ms], dtype=torch.float64, device=dev)
    cnt = torch.tensor([count, nwin], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)   # the path's only collective: per-GPU counts
    step_ms, e2e_ms, kernel_ms = t.tolist()
    total_records, total_win = cnt.tolist()
    total_bases = SHARD * world

    if rank == 0:
        peak, peak_kind = load_peaks()
        kname, traffic = load_traffic()
        algo_bytes = nwin + REC_BYTES * count          # this rank's launch
        achieved = algo_bytes / (kernel_ms * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": total_bases / (step_ms * 1e-3) / 1e9, "unit": "Gbases/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": step_ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": {"workload": "super_kmer_view k=31 m=19 canonical seed=42 over 100 Mbp/GPU of sequence_generator "
                                   "uniform(42+shard) (BASELINE configs[1])",
                       "k": K, "m": M, "bases_per_gpu": SHARD, "records": total_records, "kmers": total_win,
                       "l2": "256 MiB flush write between timed steps", "sharding": f"chunk+halo x{world}",
                       "host": numa},
            "e2e": {"value": total_bases / (e2e_ms * 1e-3) / 1e9, "unit": "Gbases/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h), "steps": e2e_steps, "ms_per_step": e2e_ms},
            "gpu_launches": args.steps,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "peak_kind": peak_kind, "kernel": kname, "kernel_ms": kernel_ms,
                         "algorithmic_bytes_per_base": algo_bytes / nwin},
            "clocks": sampler.summary(),
        }
        if world == 1:
            threads = os.cpu_count() or 1
            v, kind, desc = cpu_arm(8_000_000, threads)
            line["cpu_baseline"] = {"value": v, "unit": "Gbases/s", "cores": threads, "kind": kind, "sample": desc}
        print(json.dumps(line), flush=True)
    for p in [hp_in] + hp:
        ctx.host_free(p)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
