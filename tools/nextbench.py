#!/usr/bin/env python
"""Rates of the SURVEY 8(f) "next" rows on one B200 (development / BASELINE.md figures; not the bench contract).

    python tools/nextbench.py [--n 100000000]

ingest (FASTA text -> bases + offsets, device in / device out; packed input end to end), payloads (records -> bucket-ordered
records + 2-bit stream, device out), sort / unique / Jaccard (two related inputs). Wall clock around the C-ABI call with a
device synchronise on both sides; the library's own kernel time beside it."""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=100_000_000)
    a = ap.parse_args()
    import torch
    import biolib_b200
    from biolib_b200 import synth, _capi
    ctx = biolib_b200.Context(0)
    n = a.n
    s = synth.generate(42, n)
    out = {}

    def timed(fn, reps=3):
        best = None
        for _ in range(reps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            r = fn()
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        return best, r

    # ---- ingest: FASTA with 80-column lines, reads of 10 kbp
    L, wrap = 10_000, 80
    nrec = n // L
    lines = s[: nrec * L].reshape(nrec, L // wrap, wrap)
    body = np.concatenate([lines, np.full((nrec, L // wrap, 1), 10, np.uint8)], axis=2).reshape(nrec, -1)
    hdr = np.frombuffer(b">read_0000000000\n", np.uint8)
    text = np.concatenate([np.tile(hdr, (nrec, 1)), body], axis=1).reshape(-1)
    d_text = torch.from_numpy(text).cuda()

    def parse():
        fx = ctx.parse_fastx_device(d_text.data_ptr(), text.size, 0)
        nb, ns = fx.n_bases, fx.n_seq
        ctx.free_fastx(fx)
        return nb, ns
    dt, (nb, ns) = timed(parse)
    assert nb == nrec * L and ns == nrec
    out["fastx_parse"] = {"text_bytes": int(text.size), "records": int(ns), "ms": dt * 1e3, "GB_per_s_text": text.size / dt / 1e9,
                          "kernel_launches": ctx.stats()["kernel_launches"], "note": "device text in, device bases + offsets out, incl. output allocation"}

    # ---- packed input end to end vs ASCII end to end (host buffers, compact records)
    codes, bs, bl = ctx.pack_bases(s)
    dt_a, ra = timed(lambda: ctx.super_kmers(s, 31, 19, 42, True, want_hash=False, pos32=True)["n"])
    st_a = ctx.stats()
    dt_p, rp = timed(lambda: ctx.super_kmers_packed(codes, n, bs, bl, 31, 19, 42, True, want_hash=False, pos32=True)["n"])
    st_p = ctx.stats()
    assert ra == rp
    out["packed_input_e2e"] = {"bases": n, "ascii_ms": dt_a * 1e3, "packed_ms": dt_p * 1e3, "ascii_h2d": st_a["bytes_h2d"], "packed_h2d": st_p["bytes_h2d"],
                               "d2h": st_p["bytes_d2h"], "note": "pageable numpy buffers, library-allocated pinned outputs, incl. the Python copy of the results"}

    # ---- the k-mer entry points end to end: pinned host bases in, library-owned pinned (value, position) arrays out
    hp = ctx.host_alloc(n)
    np.ctypeslib.as_array(C.cast(hp, C.POINTER(C.c_uint8)), shape=(n,))[:] = s
    pk = biolib_b200.context.Params(k=31, m=19, canonical=1, flags=0, seed=42)

    def kmer_call(which):
        o = _capi.KmerOut()
        if which == "kmers":
            st = ctx.L.blg_kmers(ctx.h, hp, n, None, 0, C.byref(pk), C.byref(o))
        elif which == "syncmers":
            st = ctx.L.blg_syncmers(ctx.h, hp, n, None, 0, C.byref(pk), 0, 12, C.byref(o))
        else:
            st = ctx.L.blg_hash_sampled_kmers(ctx.h, hp, n, None, 0, C.byref(pk), ctx.hash_sampler_threshold(0.1), C.byref(o))
        assert st == 0, st
        cnt = o.count
        ctx.L.blg_free_kmer_out(ctx.h, C.byref(o))
        return cnt
    pk_runs = biolib_b200.context.Params(k=31, m=19, canonical=1, flags=_capi.F_RUNS, seed=42)

    def kmer_runs_call():
        o = _capi.KmerOut()
        st = ctx.L.blg_kmers(ctx.h, hp, n, None, 0, C.byref(pk_runs), C.byref(o))
        assert st == 0, st
        cnt = o.count
        ctx.L.blg_free_kmer_out(ctx.h, C.byref(o))
        return cnt
    kmer_runs_call()
    dt, cnt = timed(kmer_runs_call)
    st = ctx.stats()
    out["kmers_runs_e2e_k31"] = {"bases": n, "records": int(cnt), "ms": dt * 1e3, "Gbases_per_s": n / dt / 1e9, "h2d": st["bytes_h2d"], "d2h": st["bytes_d2h"],
                                 "note": "BLG_F_RUNS: values + run list, no position array (8 B per k-mer over PCIe)"}
    for which in ("kmers", "syncmers", "hash_sampled"):
        kmer_call(which)
        dt, cnt = timed(lambda: kmer_call(which))
        st = ctx.stats()
        out[which + "_e2e_k31"] = {"bases": n, "records": int(cnt), "ms": dt * 1e3, "Gbases_per_s": n / dt / 1e9, "h2d": st["bytes_h2d"], "d2h": st["bytes_d2h"],
                                   "kernel_ms_all": st["kernel_ms"], "kernel_launches": st["kernel_launches"]}
    ctx.host_free(hp)

    # ---- payloads, device output
    p = biolib_b200.context.Params(k=31, m=19, canonical=1, flags=_capi.F_DEVICE_INPUT | _capi.F_DEVICE_OUTPUT, seed=42)
    d_s = torch.from_numpy(s).cuda()

    def payload():
        o = _capi.PayloadOut()
        st = ctx.L.blg_super_kmer_payloads(ctx.h, C.c_void_p(d_s.data_ptr()), n, None, 0, C.byref(p), 256, C.byref(o))
        assert st == 0
        r = (o.count, o.total_bases)
        ctx.L.blg_free_payload_out(ctx.h, C.byref(o))
        return r
    dt, (cnt, tb) = timed(payload)
    st = ctx.stats()
    algo = cnt * 18 + tb + cnt * 26 + tb / 4
    out["payloads_256_buckets"] = {"bases": n, "records": int(cnt), "payload_bases": int(tb), "ms_total_call": dt * 1e3, "kernel_ms_all": st["kernel_ms"],
                                   "algorithmic_bytes": algo, "note": "device in / out; the call = super-k-mer kernel + 10 payload launches + allocations; "
                                   "algorithmic bytes = 18 B/record in + 1 B/payload base read + 26 B/record + 0.25 B/payload base out"}

    # ---- sort / unique / Jaccard: two inputs of n/2 bases, 2 % substitutions
    h = n // 2
    a1 = s[:h]
    b1 = a1.copy()
    rng = np.random.default_rng(1)
    mut = rng.random(h) < 0.02
    b1[mut] = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=int(mut.sum()))
    d_a, d_b = torch.from_numpy(a1).cuda(), torch.from_numpy(b1).cuda()
    pj = biolib_b200.context.Params(k=31, m=0, canonical=1, flags=_capi.F_DEVICE_INPUT, seed=0)

    def jac():
        o = _capi.JaccardOut()
        st = ctx.L.blg_kmer_jaccard(ctx.h, C.c_void_p(d_a.data_ptr()), h, None, 0, C.c_void_p(d_b.data_ptr()), h, None, 0, C.byref(pj), C.byref(o))
        assert st == 0
        return o.intersection, o.union_size, o.size1, o.size2
    dt, r = timed(jac, reps=2)
    st = ctx.stats()
    out["jaccard_k31"] = {"bases_each": h, "result": list(map(int, r)), "ms_total_call": dt * 1e3, "kernel_ms_all": st["kernel_ms"],
                          "kmers_per_s": 2 * h / dt, "kernel_launches": st["kernel_launches"]}
    print(json.dumps(out, indent=1))
    ctx.close()


if __name__ == "__main__":
    main()
