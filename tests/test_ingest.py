"""Ingest: the 2-bit packed wire format and the FASTA / FASTQ record splitter.

CPU part: the host packer against a numpy restatement; the parser restatement (oracle) against the reference's own kseq.h
executed over the same text (oracle/_ref). GPU part: blg_super_kmers_packed == blg_super_kmers on the same bases;
blg_parse_fastx == the oracle (and == kseq when oracle/_ref travelled)."""
import numpy as np
import pytest


def rand_seq(rng, n, p_n=0.0):
    s = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=n)
    if p_n > 0:
        s = s.copy()
        s[rng.random(n) < p_n] = ord("N")
    return s


def make_fasta(rng, n_rec, wrap, crlf=False, last_newline=True, lens=None):
    nl = b"\r\n" if crlf else b"\n"
    out = []
    for i in range(n_rec):
        L = int(rng.integers(0, 700)) if lens is None else lens[i]
        s = rand_seq(rng, L, 0.01).tobytes()
        out.append(b">read%d some comment" % i + nl)
        if wrap:
            out += [s[j:j + wrap] + nl for j in range(0, L, wrap)]
        else:
            out.append(s + nl)
    t = b"".join(out)
    return t if last_newline else t.rstrip(b"\r\n")


def make_fastq(rng, n_rec, crlf=False):
    nl = b"\r\n" if crlf else b"\n"
    out = []
    for i in range(n_rec):
        L = int(rng.integers(1, 400))
        s = rand_seq(rng, L, 0.01).tobytes()
        q = rng.choice(np.frombuffer(b"!@>+IJ#5", np.uint8), size=L).tobytes()      # qualities that look like headers
        out += [b"@r%d/1" % i + nl, s + nl, b"+" + nl, q + nl]
    return b"".join(out)


def test_parser_restatement_vs_kseq(oracle, ref):
    rng = np.random.default_rng(1)
    texts = [make_fasta(rng, 40, 60), make_fasta(rng, 25, 0), make_fasta(rng, 30, 70, crlf=True), make_fasta(rng, 10, 61, last_newline=False),
             make_fasta(rng, 3, 80, lens=[0, 5000, 1]), make_fastq(rng, 50), make_fastq(rng, 20, crlf=True)]
    for t in texts:
        ob, oo = oracle.parse_fastx(t)
        rb, ro = ref.kseq_parse(t)
        assert np.array_equal(oo, ro) and np.array_equal(ob, rb)


def test_host_packer(oracle):
    import biolib_b200
    from biolib_b200 import _capi
    import ctypes as C
    L = _capi.lib()
    rng = np.random.default_rng(2)
    for n in (0, 1, 15, 16, 17, 1000, 40_003):
        s = rand_seq(rng, n, 0.02)
        if n > 100:
            s[10:40] = ord("N"); s[-3:] = ord("n"); s[50] = ord("u"); s[51] = ord("R")
        oc, os_, ol = oracle.pack_bases(s)
        codes = np.zeros((n + 15) // 16, np.uint32)
        bs, bl = np.zeros(n + 1, np.uint64), np.zeros(n + 1, np.uint32)
        nb = C.c_size_t()
        assert L.blg_pack_bases(s.ctypes.data, n, codes.ctypes.data, bs.ctypes.data, bl.ctypes.data, n + 1, C.byref(nb)) == 0
        assert np.array_equal(codes, oc) and np.array_equal(bs[:nb.value], os_) and np.array_equal(bl[:nb.value], ol)
        if nb.value > 1:
            assert L.blg_pack_bases(s.ctypes.data, n, codes.ctypes.data, bs.ctypes.data, bl.ctypes.data, 1, C.byref(nb)) == _capi.BLG_E_CAPACITY


@pytest.mark.gpu
@pytest.mark.parametrize("k,m", [(31, 19), (21, 11), (63, 31), (9, 5)])
def test_packed_input_equals_ascii(gpu, oracle, k, m):
    rng = np.random.default_rng(k)
    for n in (0, 5, 1000, 70_001):
        s = rand_seq(rng, n, 0.003)
        if n > 1000:
            s[300:420] = ord("N")
        codes, bs, bl = gpu.pack_bases(s)
        for canon in (False, True):
            a = gpu.super_kmers(s, k, m, 42, canon)
            b = gpu.super_kmers_packed(codes, n, bs, bl, k, m, 42, canon)
            assert a["n"] == b["n"]
            for key in ("value", "hash", "mm_start", "size", "mm_pos"):
                assert np.array_equal(a[key], b[key]), (key, k, m, n)
    o = oracle.super_kmers(s, k, m, 42, True)
    assert np.array_equal(b["value"], o["value"]) and np.array_equal(b["mm_start"], o["mm_start"])


@pytest.mark.gpu
def test_packed_input_pipelined_large(gpu):
    """>= 8 Mi windows: the chunked host path with staged (unpacked-on-device) input, both record layouts, quarter the H2D."""
    rng = np.random.default_rng(3)
    n = 20_000_000
    s = rand_seq(rng, n, 0.0001)
    offs = np.arange(0, n + 1, 20_000, dtype=np.uint64)
    codes, bs, bl = gpu.pack_bases(s)
    a = gpu.super_kmers(s, 31, 19, 42, True, want_hash=False, offsets=offs)
    for pos32 in (False, True):
        b = gpu.super_kmers_packed(codes, n, bs, bl, 31, 19, 42, True, want_hash=False, offsets=offs, pos32=pos32)
        st = gpu.stats()
        assert a["n"] == b["n"] > 2_000_000
        for key in ("value", "mm_start", "size", "mm_pos"):
            assert np.array_equal(a[key], b[key]), key
        assert st["bytes_h2d"] < n // 4 + bs.size * 12 + offs.size * 8 + 4096


@pytest.mark.gpu
def test_fastx_parser(gpu, oracle):
    import pyoracle
    ref = pyoracle.Ref() if pyoracle.Ref.available() else None
    rng = np.random.default_rng(4)
    texts = [make_fasta(rng, 400, 60), make_fasta(rng, 25, 0), make_fasta(rng, 300, 70, crlf=True), make_fasta(rng, 10, 61, last_newline=False),
             make_fasta(rng, 3, 80, lens=[0, 50_000, 1]), make_fasta(rng, 2, 0, lens=[30_000, 17]), make_fastq(rng, 500), make_fastq(rng, 200, crlf=True),
             b">only\nACGT", b"@r\nACGT\n+\nIIII\n"]
    for t in texts:
        gb, go = gpu.parse_fastx(t)
        ob, oo = oracle.parse_fastx(t)
        assert np.array_equal(go, oo) and np.array_equal(gb, ob)
        if ref is not None:
            rb, ro = ref.kseq_parse(t)
            assert np.array_equal(go, ro) and np.array_equal(gb, rb)
    with pytest.raises(Exception):
        gpu.parse_fastx(b"ACGT\n>late\nAC\n", 0)
    # parse -> super-k-mers entirely on the device
    import torch
    t = make_fasta(rng, 300, 80)
    d = torch.frombuffer(bytearray(t), dtype=torch.uint8).cuda()
    fx = gpu.parse_fastx_device(d.data_ptr(), len(t))
    try:
        ob, oo = oracle.parse_fastx(t)
        assert fx.n_bases == ob.size and fx.n_seq == len(oo) - 1
        cap = ob.size + 1024
        v = torch.empty(cap, dtype=torch.int64, device="cuda"); mm = torch.empty_like(v)
        sz = torch.empty(cap, dtype=torch.uint8, device="cuda"); ps = torch.empty_like(sz)
        cnt = gpu.super_kmers_device(fx.bases, fx.n_bases, 21, 11, 42, True, v.data_ptr(), mm.data_ptr(), sz.data_ptr(), ps.data_ptr(), cap,
                                     d_offsets=fx.seq_offsets, n_seq=fx.n_seq)
        o = oracle.super_kmers(ob, 21, 11, 42, True, offsets=oo)
        assert cnt == o["n"]
        assert np.array_equal(v[:cnt].cpu().numpy().view(np.uint64), o["value"]) and np.array_equal(mm[:cnt].cpu().numpy().view(np.uint64), o["mm_start"])
    finally:
        gpu.free_fastx(fx)


@pytest.mark.gpu
def test_packed_input_chunked_upload(gpu):
    """Packed input on the chunked host path is uploaded and unpacked piece by piece: break runs across the piece boundaries, the
    10-byte layout, caller arrays, and the fall-back to the plain path when the caller's arrays are too small."""
    import biolib_b200
    from biolib_b200 import _capi
    from biolib_b200.context import expand_runs
    rng = np.random.default_rng(33)
    n = 30_000_000
    s = rand_seq(rng, n, 0.00005)
    nwin = n - 30
    for cut in (nwin // 2 // 4, nwin // 2 // 4 + (nwin - nwin // 2 // 4), 12_000_000):      # around where the pieces meet
        c16 = cut // 16 * 16
        s[max(c16 - 3000, 0): min(c16 + 3000, n)] = ord("N")
        s[min(c16 + 5000, n - 1): min(c16 + 5001, n)] = ord("N")
    codes, bs, bl = gpu.pack_bases(s)
    a = gpu.super_kmers(s, 31, 19, 42, True, want_hash=False)
    cap = a["n"] + 100
    v, mm = np.zeros(cap, np.uint64), np.zeros(cap, np.uint64)
    sz, mp = np.zeros(cap, np.uint8), np.zeros(cap, np.uint8)
    rf, rs = np.zeros(1 << 16, np.uint64), np.zeros(1 << 16, np.uint64)
    cnt = gpu.super_kmers_packed_raw(codes.ctypes.data, n, bs.ctypes.data, bl.ctypes.data, bs.size, 31, 19, 42, True, v.ctypes.data, mm.ctypes.data,
                                     sz.ctypes.data, mp.ctypes.data, cap)
    st = gpu.stats()
    assert cnt == a["n"] and st["bytes_h2d"] < n // 4 + bs.size * 12 + 4096 and st["kernel_launches"] >= 4
    assert np.array_equal(v[:cnt], a["value"]) and np.array_equal(mm[:cnt], a["mm_start"]) and np.array_equal(sz[:cnt], a["size"])
    v[:] = 0
    cnt = gpu.super_kmers_packed_raw(codes.ctypes.data, n, bs.ctypes.data, bl.ctypes.data, bs.size, 31, 19, 42, True, v.ctypes.data, 0,
                                     sz.ctypes.data, mp.ctypes.data, cap, flags=_capi.F_RUNS, run_ptrs=(rf.ctypes.data, rs.ctypes.data), run_capacity=1 << 16)
    nr = gpu.last_n_runs
    assert cnt == a["n"] and 3 < nr < (1 << 16)
    assert np.array_equal(v[:cnt], a["value"]) and np.array_equal(expand_runs(rf[:nr], rs[:nr], sz[:cnt], mp[:cnt]), a["mm_start"])
    with pytest.raises(biolib_b200.BlgError) as e:
        gpu.super_kmers_packed_raw(codes.ctypes.data, n, bs.ctypes.data, bl.ctypes.data, bs.size, 31, 19, 42, True, v.ctypes.data, mm.ctypes.data,
                                   sz.ctypes.data, mp.ctypes.data, 1000)
    assert e.value.status == 3
    # a window range in the middle of the packed input
    lo, hi = 5_000_003, 29_000_000
    b = gpu.super_kmers(s, 31, 19, 42, True, want_hash=False, win_range=(lo, hi))
    cnt = gpu.super_kmers_packed_raw(codes.ctypes.data, n, bs.ctypes.data, bl.ctypes.data, bs.size, 31, 19, 42, True, v.ctypes.data, mm.ctypes.data,
                                     sz.ctypes.data, mp.ctypes.data, cap, win_range=(lo, hi))
    assert cnt == b["n"] and np.array_equal(v[:cnt], b["value"]) and np.array_equal(mm[:cnt], b["mm_start"])
