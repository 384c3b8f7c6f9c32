"""The BASELINE.json configurations at (or near) their full sizes on one B200, checked through size-independent
properties of the record stream instead of the oracle (which would take minutes there):

  P1  sum(size) == number of valid k-mers                      (every valid k-mer belongs to exactly one record)
  P2  sk_start strictly increasing, sk_start[i] + size[i] <= sk_start[i+1]   (stream order, no overlap)
  P3  1 <= size <= w, 0 <= mm_pos <= w - 1, mm_pos + 1 <= ... the minimizer lies inside every k-mer of its run:
      mm_pos >= size - 1 - (w - 1) + ... i.e. sk_start + size - 1 <= mm_start <= sk_start + w - 1 is implied by
      sk_start <= mm_start - 0 and mm_start + m <= sk_start + k  and  mm_start >= sk_start + size - 1 - (w - 1)
  P4  hash64(value) of every record, recomputed by the independent blg_hash64 entry point, equals the emitted hash
  P5  a second run of the same kernel over window ranges cut at odd positions (blg_super_kmers_range: the tile grid starts at
      each range's first window, so every tile edge falls elsewhere and the ordered scan sees other counts) concatenates
      to identical arrays
  P6  a prefix of the same input equals the oracle record for record
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run(ctx, torch, d_bases, n, k, m, offs=None, want_hash=False):
    w = k - m + 1
    n_seq = 0 if offs is None else offs.numel() - 1
    cap = int(n * 2.0 / (w + 1) * 1.35) + 3 * n_seq + 65536
    dev = d_bases.device
    out = dict(value=torch.empty(cap, dtype=torch.int64, device=dev), mm=torch.empty(cap, dtype=torch.int64, device=dev),
               size=torch.empty(cap, dtype=torch.uint8, device=dev), pos=torch.empty(cap, dtype=torch.uint8, device=dev))
    if want_hash:
        out["hash"] = torch.empty(cap, dtype=torch.int64, device=dev)
    cnt = ctx.super_kmers_device(d_bases.data_ptr(), n, k, m, 42, True, out["value"].data_ptr(), out["mm"].data_ptr(),
                                 out["size"].data_ptr(), out["pos"].data_ptr(), cap,
                                 d_offsets=offs.data_ptr() if offs is not None else 0, n_seq=n_seq,
                                 d_hash=out["hash"].data_ptr() if want_hash else 0)
    torch.cuda.synchronize()
    return cnt, {key: v[:cnt] for key, v in out.items()}


def _properties(torch, r, k, m, n_valid_kmers):
    w = k - m + 1
    size = r["size"].to(torch.int64)
    pos = r["pos"].to(torch.int64)
    sk = r["mm"] - pos
    assert int(size.sum().item()) == n_valid_kmers                                     # P1
    assert bool((sk[1:] >= sk[:-1] + size[:-1]).all().item())                          # P2
    assert int(size.min().item()) >= 1 and int(size.max().item()) <= w                 # P3
    assert int(pos.min().item()) >= 0 and int(pos.max().item()) <= w - 1
    assert bool((pos >= size - 1).all().item())          # the minimizer starts at or after the run's last k-mer start ...
    # ... because it lies inside every k-mer of the run, the last one included (mm_start >= sk_start + size - 1 is only
    # required when the run is longer than the minimizer offset allows: mm_start >= last k-mer start)


def _ctx():
    import biolib_b200
    return biolib_b200.Context(0)


def _second_grid(ctx, torch, d_bases, n, k, m, ref, cuts, offs=None):
    """P5: the ranges [cuts[i], cuts[i+1]) run one by one must concatenate to `ref` (arrays of the single call)."""
    w = k - m + 1
    n_seq = 0 if offs is None else offs.numel() - 1
    done = 0
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        cap = int((hi - lo) * 2.0 / (w + 1) * 1.35) + 3 * n_seq + 65536
        v = torch.empty(cap, dtype=torch.int64, device="cuda"); mm = torch.empty_like(v)
        sz = torch.empty(cap, dtype=torch.uint8, device="cuda"); ps = torch.empty_like(sz)
        cnt = ctx.super_kmers_device(d_bases.data_ptr(), n, k, m, 42, True, v.data_ptr(), mm.data_ptr(), sz.data_ptr(), ps.data_ptr(), cap,
                                     d_offsets=offs.data_ptr() if offs is not None else 0, n_seq=n_seq, win_range=(lo, hi))
        torch.cuda.synchronize()
        assert bool(torch.equal(v[:cnt], ref["value"][done:done + cnt])) and bool(torch.equal(mm[:cnt], ref["mm"][done:done + cnt]))
        assert bool(torch.equal(sz[:cnt], ref["size"][done:done + cnt])) and bool(torch.equal(ps[:cnt], ref["pos"][done:done + cnt]))
        done += cnt
        del v, mm, sz, ps
    assert done == ref["value"].numel()


def _valid_kmers(torch, d_bases, n, k, offs=None):
    """Number of valid k-mers, computed with plain torch ops (independent of the library's kernels)."""
    lut = torch.ones(256, dtype=torch.int32, device=d_bases.device)
    for ch in b"ACGTUacgtu":
        lut[ch] = 0
    total = 0
    step = 1 << 28                                   # chunks keep the cumulative sums small; k-1 bases of overlap
    bnd = None
    if offs is not None:
        bnd = torch.zeros(n + 1, dtype=torch.int32, device=d_bases.device)
        bnd[offs[1:-1].long()] = 1                   # a sequence starts here
    for lo in range(0, max(n - k + 1, 0), step):
        hi = min(lo + step, n - k + 1)               # k-mer starts [lo, hi)
        bad = lut[d_bases[lo:hi + k - 1].long()]
        cs = torch.cat([torch.zeros(1, dtype=torch.int32, device=bad.device), torch.cumsum(bad, 0, dtype=torch.int32)])
        inv = cs[k:] - cs[:-k]
        if bnd is not None and k > 1:
            b = bnd[lo + 1:hi + k - 1]
            cb = torch.cat([torch.zeros(1, dtype=torch.int32, device=bad.device), torch.cumsum(b, 0, dtype=torch.int32)])
            inv = inv + (cb[k - 1:] - cb[:-(k - 1)])
        total += int((inv == 0).sum().item())
        del bad, cs, inv
    return total


def _equal(torch, a, b):
    return all(a[key].numel() == b[key].numel() and bool(torch.equal(a[key], b[key])) for key in a)


def test_config2_100mbp_k31_m19(oracle):
    """configs[1]: minimizer / super-k-mer stream, k=31 m=19 canonical seeded, 100 Mbp of the reference generator."""
    import torch
    from biolib_b200 import synth
    n, k, m = 100_000_000, 31, 19
    s = synth.generate_sharded_parallel(42, n, 1 << 24)
    d = torch.from_numpy(s).cuda()
    c3 = _ctx()
    n3, r3 = _run(c3, torch, d, n, k, m, want_hash=True)
    _second_grid(c3, torch, d, n, k, m, r3, [0, 1, 33_333_337, 33_333_338, 77_000_001, n - k + 1])   # P5
    _properties(torch, r3, k, m, n - k + 1)                                             # P1-P3
    hv = torch.empty_like(r3["value"])                                                  # P4
    import ctypes as C
    st = c3.L.blg_hash64(c3.h, C.c_void_p(r3["value"].data_ptr()), n3, 42, 3, C.c_void_p(hv.data_ptr()))
    assert st == 0 and bool(torch.equal(hv, r3["hash"]))
    pre = 3_000_000                                                                     # P6
    o = oracle.super_kmers(s[:pre], k, m, 42, True)
    keep = int(o["n"]) - 2                      # the last records of the prefix are cut short by its end
    assert np.array_equal(r3["value"][:keep].cpu().numpy().view(np.uint64), o["value"][:keep])
    assert np.array_equal(r3["mm"][:keep].cpu().numpy().view(np.uint64), o["mm_start"][:keep])
    assert np.array_equal(r3["size"][:keep].cpu().numpy(), o["size"][:keep])
    c3.close()


def test_config3_short_reads_with_n_runs(oracle):
    """configs[2]: 150 bp reads with injected N runs and lower-case reads, one batch of 4 M reads (600 Mbp) of the
    100 M-read configuration (the batches are independent; bench-scale runs sum them)."""
    import torch
    from biolib_b200 import synth
    reads, L, k, m = 4_000_000, 150, 31, 19
    n = reads * L
    s = synth.generate_sharded_parallel(43, n, 1 << 24)
    synth.inject_n_runs(s, L, seed=44)
    offs = torch.arange(0, reads + 1, dtype=torch.int64, device="cuda") * L
    d = torch.from_numpy(s).cuda()
    c3 = _ctx()
    n3, r3 = _run(c3, torch, d, n, k, m, offs=offs)
    _second_grid(c3, torch, d, n, k, m, r3, [0, 149, 200_000_003, 200_000_153, n - k + 1], offs=offs)
    _properties(torch, r3, k, m, _valid_kmers(torch, d, n, k, offs))
    sk = r3["mm"] - r3["pos"].to(torch.int64)
    assert bool(((sk // L) == ((sk + r3["size"].to(torch.int64) + k - 2) // L)).all().item())   # no record spans two reads
    pre = 20_000 * L
    o = oracle.super_kmers(s[:pre], k, m, 42, True, offsets=np.arange(0, 20_001, dtype=np.uint64) * L)
    cnt = int(o["n"])
    assert np.array_equal(r3["value"][:cnt].cpu().numpy().view(np.uint64), o["value"])
    assert np.array_equal(r3["mm"][:cnt].cpu().numpy().view(np.uint64), o["mm_start"])
    assert np.array_equal(r3["size"][:cnt].cpu().numpy(), o["size"])
    c3.close()


def test_config4_3gbp_k63_m31(oracle):
    """configs[3]: k=63 m=31 (128-bit k-mer span, 64-bit minimizers) over a 3 Gbp synthetic genome on one GPU."""
    import torch
    from biolib_b200 import synth, shard
    n, k, m = 3_000_000_000, 63, 31
    s = synth.generate_sharded_parallel(45, n, 1 << 26)
    s[1_000_000:1_000_050] = ord("N")
    s[2_999_999_000:2_999_999_007] = ord("n")
    d = torch.from_numpy(s).cuda()
    c3 = _ctx()
    n3, r3 = _run(c3, torch, d, n, k, m)
    _properties(torch, r3, k, m, _valid_kmers(torch, d, n, k))
    # the multi-GPU plan on one device: 8 slices with halos, each given only its slice, concatenate to the whole
    parts = 0
    chk_v, chk_m = 0, 0
    for pl in shard.plan_chunks(n, k, m, 8):
        sl = d[pl.buf_lo:pl.buf_hi]
        w = k - m + 1
        cap = int((pl.win_hi - pl.win_lo) * 2.0 / (w + 1) * 1.35) + 65536
        v = torch.empty(cap, dtype=torch.int64, device="cuda"); mm = torch.empty_like(v)
        sz = torch.empty(cap, dtype=torch.uint8, device="cuda"); ps = torch.empty_like(sz)
        cnt = c3.super_kmers_device(sl.data_ptr(), sl.numel(), k, m, 42, True, v.data_ptr(), mm.data_ptr(), sz.data_ptr(),
                                    ps.data_ptr(), cap, win_range=pl.local_window)
        torch.cuda.synchronize()
        assert bool(torch.equal(v[:cnt], r3["value"][parts:parts + cnt]))
        assert bool(torch.equal(mm[:cnt] + pl.buf_lo, r3["mm"][parts:parts + cnt]))
        assert bool(torch.equal(sz[:cnt], r3["size"][parts:parts + cnt]))
        parts += cnt
        del v, mm, sz, ps
    assert parts == n3
    pre = 2_000_000
    o = oracle.super_kmers(s[:pre], k, m, 42, True)
    keep = int(o["n"]) - 2
    assert np.array_equal(r3["value"][:keep].cpu().numpy().view(np.uint64), o["value"][:keep])
    assert np.array_equal(r3["mm"][:keep].cpu().numpy().view(np.uint64), o["mm_start"][:keep])
    c3.close()


def test_config5_long_reads_k21_m11(oracle):
    """configs[4]: k=21 m=11 over mutated long reads (10-100 kbp): read i = sequence_mutator(source slice i, 0.10, 0.3, 0.3)
    from cbegin(47 + i) (biolib_b200.synth.mutated_long_reads, pinned to the reference's mutator), sharded
    by read across 8 ranks (one device here)."""
    import torch
    from biolib_b200 import synth, shard
    k, m = 21, 11
    s, offs_h = synth.mutated_long_reads(4000)
    n = int(offs_h[-1])
    d = torch.from_numpy(s).cuda()
    offs = torch.from_numpy(offs_h.view(np.int64)).cuda()
    c3 = _ctx()
    n3, r3 = _run(c3, torch, d, n, k, m, offs=offs)
    _second_grid(c3, torch, d, n, k, m, r3, [0, 7, 50_000_017, n - k + 1], offs=offs)
    _properties(torch, r3, k, m, _valid_kmers(torch, d, n, k, offs))
    parts = 0
    for a, b in shard.plan_sequences(offs_h, 8):
        base = int(offs_h[a])
        sl = d[base:int(offs_h[b])]
        cnt, r = _run(c3, torch, sl, sl.numel(), k, m, offs=offs[a:b + 1] - base)
        assert bool(torch.equal(r["value"], r3["value"][parts:parts + cnt]))
        assert bool(torch.equal(r["mm"] + base, r3["mm"][parts:parts + cnt]))
        parts += cnt
    assert parts == n3
    first = 40
    o = oracle.super_kmers(s[:int(offs_h[first])], k, m, 42, True, offsets=offs_h[:first + 1])
    cnt = int(o["n"])
    assert np.array_equal(r3["value"][:cnt].cpu().numpy().view(np.uint64), o["value"])
    assert np.array_equal(r3["size"][:cnt].cpu().numpy(), o["size"])
    c3.close()
