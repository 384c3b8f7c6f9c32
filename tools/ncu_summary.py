#!/usr/bin/env python
"""Summarise an `ncu --set full` report of the fused kernel into the files bench.py and the judge read.

    python tools/ncu_summary.py gpurun_out/X.ncu-rep --bases 100000000 --tag r02_x [--kernel-index 0]

Writes, under profiles/:
  <tag>_details.txt        `ncu --page details` of the capture (text)
  <tag>_instmix.txt        executed instruction mix + stall reasons from the source page
  ncu_traffic.json         DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) -> bench.py roofline.traffic
  ncu_intmix.json          instructions per base, pipe utilisations, issue slots -> bench.py roofline_int
Both JSON files carry the git sha of the tree the capture was taken from, so a stale number is visible.
"""
import argparse
import collections
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")


def ncu(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True, check=True).stdout


def to_bytes(value, unit):
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    return float(value) * scale[unit]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("--bases", type=float, default=1e8)
    ap.add_argument("--tag", required=True)
    ap.add_argument("--kernel-index", type=int, default=0)
    ap.add_argument("--sha", default=None)
    ap.add_argument("--no-json", action="store_true", help="only the text summaries (captures that are not the bench workload)")
    a = ap.parse_args()
    sha = a.sha or subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    prof = os.path.join(ROOT, "profiles")
    details = ncu(["-i", a.report, "--page", "details"])
    with open(os.path.join(prof, a.tag + "_details.txt"), "w") as f:
        f.write(details)
    rows = list(csv.reader(io.StringIO(ncu(["-i", a.report, "--page", "raw", "--csv"]))))
    hdr, units, vals = rows[0], rows[1], rows[2 + a.kernel_index]
    m = {k: (v, u) for k, u, v in zip(hdr, units, vals)}
    get = lambda k: float(m[k][0].replace(",", ""))
    kernel = m["Kernel Name"][0]
    dram = to_bytes(*m["dram__bytes_read.sum"]) + to_bytes(*m["dram__bytes_write.sum"])
    dur_us = get("gpu__time_duration.sum") * {"us": 1, "ms": 1e3, "ns": 1e-3, "s": 1e6}[m["gpu__time_duration.sum"][1]]
    warp_inst = get("smsp__inst_executed.sum")
    intmix = {
        "kernel": kernel, "git_sha": sha, "report": os.path.basename(a.report), "bases": a.bases,
        "duration_us_under_ncu": dur_us,
        "warp_instructions": warp_inst, "inst_per_base": warp_inst * 32 / a.bases,
        "alu_pipe_pct": get("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"),
        "fma_pipe_pct": get("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
        "issue_slots_pct": get("smsp__issue_active.avg.pct_of_peak_sustained_active"),
        "ipc_per_sm": get("sm__inst_executed.avg.per_cycle_elapsed"),
        "warps_active_pct": get("sm__warps_active.avg.pct_of_peak_sustained_active"),
        "registers_per_thread": get("launch__registers_per_thread"),
    }
    # source page: instruction mix + stalls
    src = list(csv.reader(io.StringIO(ncu(["-i", a.report, "--page", "source", "--csv"]))))
    blocks, cur = [], None
    for r in src:
        if r and r[0] == "Kernel Name":
            cur = []
            blocks.append(cur)
            continue
        if cur is not None and r:
            cur.append(r)
    b = blocks[min(a.kernel_index, len(blocks) - 1)]
    h = b[0]
    data = [r for r in b[1:] if len(r) == len(h)]
    ix = {k: i for i, k in enumerate(h)}
    ops, smp, stalls = collections.Counter(), collections.Counter(), collections.Counter()
    for r in data:
        mm = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[ix["Source"]].strip())
        op = mm.group(2).split(".")[0] if mm else "?"
        ops[op] += int(r[ix["Instructions Executed"]])
        smp[op] += int(r[ix["# Samples"]])
        for k in ix:
            if k.startswith("stall_") and "Not Issued" not in k:
                stalls[k] += int(r[ix[k]])
    tot, ts = sum(ops.values()), max(sum(smp.values()), 1)
    with open(os.path.join(prof, a.tag + "_instmix.txt"), "w") as f:
        f.write(f"# executed SASS instruction mix + stall reasons (ncu source page), {kernel}, tree {sha}\n")
        f.write(f"total warp inst {tot}  thread inst per base {tot * 32 / a.bases:.2f}  samples {ts}\n")
        for op, c in ops.most_common(32):
            f.write(f"{op:14s} {c:12d} {100 * c / tot:5.1f}%  per base {c * 32 / a.bases:6.2f}  samples {100 * smp[op] / ts:5.1f}%\n")
        f.write("\n")
        for k, v in stalls.most_common(14):
            f.write(f"{k} {v} {100 * v / ts:.1f}%\n")
    alu_ops = sum(ops[o] for o in ("LOP3", "SHF", "IADD3", "ISETP", "SEL", "VIMNMX", "VIADDMNMX", "VIADD", "PRMT", "LEA", "PLOP3", "VIMNMX3"))
    intmix["alu_ops_per_base"] = alu_ops * 32 / a.bases
    intmix["imad_per_base"] = ops["IMAD"] * 32 / a.bases
    intmix["stall_no_inst_pct"] = 100 * stalls["stall_no_inst"] / ts
    if not a.no_json:
        with open(os.path.join(prof, "ncu_traffic.json"), "w") as f:
            json.dump({"kernel": kernel, "dram_bytes_per_launch": dram, "git_sha": sha, "report": os.path.basename(a.report),
                       "how": "dram__bytes_read.sum + dram__bytes_write.sum of one ncu --set full capture of the bench workload"}, f, indent=1)
        with open(os.path.join(prof, "ncu_intmix.json"), "w") as f:
            json.dump(intmix, f, indent=1)
    print(json.dumps(intmix, indent=1))
    print("dram bytes per launch", dram)


if __name__ == "__main__":
    sys.exit(main())
