"""CUDA path vs the oracle, through the C ABI (bit-exact: everything here is integer work)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

WORKED = "ACGGTCATTGACNCGTAGGATCCAGTT"


def rand_seq(rng, n, p_n=0.0, lower=0.0):
    s = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=n)
    if p_n > 0:
        s = s.copy()
        s[rng.random(n) < p_n] = ord("N")
    if lower > 0:
        m = rng.random(n) < lower
        s = np.where(m, s | 0x20, s).astype(np.uint8)
    return s


def check_sk(gpu, oracle, seq, k, m, seed, canon, offsets=None):
    g = gpu.super_kmers(seq, k, m, seed, canon, offsets=offsets)
    o = oracle.super_kmers(seq, k, m, seed, canon, offsets=offsets)
    assert g["n"] == o["n"], (k, m, seed, canon, g["n"], o["n"])
    for key in ("value", "hash", "mm_start", "sk_start", "size"):
        assert np.array_equal(g[key], o[key]), (key, k, m, canon)
    assert np.array_equal(g["mm_pos"].astype(np.uint64), o["mm_start"] - o["sk_start"])


def test_worked_example(gpu, oracle):
    for canon in (False, True):
        check_sk(gpu, oracle, WORKED, 7, 4, 42, canon)
    g = gpu.super_kmers(WORKED, 7, 4, 42, False)
    assert g["value"].tolist() == [107, 173, 62, 178, 40, 212]
    assert g["mm_start"].tolist() == [1, 2, 6, 14, 16, 20]
    assert g["size"].tolist() == [2, 1, 3, 2, 2, 4]
    assert g["mm_pos"].tolist() == [1, 0, 3, 1, 1, 3]


@pytest.mark.parametrize("k,m", [(7, 4), (31, 19), (21, 11), (63, 31), (15, 10), (5, 5), (8, 4), (64, 32), (33, 1), (13, 12), (40, 9)])
@pytest.mark.parametrize("canon", [False, True])
def test_random_clean(gpu, oracle, k, m, canon):
    rng = np.random.default_rng(k * 100 + m)
    for n in (0, 1, k - 1, k, k + 1, 100, 5000, 70001):
        check_sk(gpu, oracle, rand_seq(rng, n), k, m, 42, canon)


def _sweep_pairs():
    """One (k, m) for every k - m in 0 .. 63 (the general kernel form splits k - m = q B + r with B = 16, 12, 8, 6 and
    picks 2- or 3-word lane streams from m), plus the corners of each block size and the largest m per stream width."""
    rng = np.random.default_rng(20240)
    out = []
    for d in range(0, 64):
        m = int(rng.integers(1, min(32, 64 - d) + 1))
        out.append((m + d, m))
    out += [(32 + 15, 32), (17 + 15, 17), (18 + 16, 18), (21 + 11, 21), (22 + 14, 22), (25 + 7, 25), (26 + 10, 26), (27 + 5, 27), (28 + 6, 28),
            (32 + 31, 32), (1 + 63, 1), (16 + 47, 16), (17 + 47, 17), (32 + 5, 32), (1 + 5, 1), (30 + 12, 30), (12 + 10, 17 - 5)]
    return sorted(set(out))


@pytest.mark.parametrize("k,m", _sweep_pairs())
def test_every_window_length(gpu, oracle, k, m):
    rng = np.random.default_rng(k * 131 + m)
    for canon in (False, True):
        check_sk(gpu, oracle, rand_seq(rng, 30011), k, m, 42, canon)
    s = rand_seq(rng, 40000, p_n=0.004, lower=0.05)
    check_sk(gpu, oracle, s, k, m, 7, True)
    lens = rng.integers(0, 3 * k, size=400)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    check_sk(gpu, oracle, rand_seq(rng, int(offsets[-1]), p_n=0.002), k, m, 42, True, offsets=offsets)
    check_sk(gpu, oracle, "ACGT" * 300 + "A" * 700 + "AT" * 300, k, m, 42, True)


@pytest.mark.parametrize("k,m", [(7, 4), (31, 19), (21, 11), (63, 31), (16, 8)])
def test_random_with_breaks(gpu, oracle, k, m):
    rng = np.random.default_rng(7 * k + m)
    for n, pn in ((300, 0.05), (20000, 0.01), (20000, 0.2), (150000, 0.002)):
        s = rand_seq(rng, n, p_n=pn, lower=0.1)
        for canon in (False, True):
            check_sk(gpu, oracle, s, k, m, 0x10000002A, canon)


def test_low_complexity_ties(gpu, oracle):
    # repeats give equal hashes inside one window: leftmost must win, and a repeated value at a new position cuts
    for s in ("A" * 500, "AT" * 400, "ACG" * 300, "A" * 100 + "N" + "A" * 100, "ACGT" * 50 + "TTTT" * 30):
        for k, m in ((7, 4), (31, 19), (12, 6), (9, 9)):
            for canon in (False, True):
                check_sk(gpu, oracle, s, k, m, 1, canon)


def test_all_byte_values(gpu, oracle):
    rng = np.random.default_rng(5)
    s = rng.integers(0, 256, size=40000, dtype=np.uint8)
    acgt = rand_seq(rng, 40000)
    s = np.where(rng.random(40000) < 0.97, acgt, s).astype(np.uint8)
    check_sk(gpu, oracle, s, 15, 7, 9, True)
    check_sk(gpu, oracle, s, 31, 19, 9, False)


@pytest.mark.parametrize("k,m", [(31, 19), (21, 11), (9, 5)])
def test_multi_sequence(gpu, oracle, k, m):
    rng = np.random.default_rng(11 + k)
    lens = rng.integers(0, 400, size=300)
    lens[::17] = 0
    lens[5] = k
    lens[6] = k - 1
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    s = rand_seq(rng, int(offsets[-1]), p_n=0.003)
    for canon in (False, True):
        check_sk(gpu, oracle, s, k, m, 42, canon, offsets=offsets)
    # 150 bp reads
    offsets = (np.arange(0, 2001) * 150).astype(np.uint64)
    s = rand_seq(rng, int(offsets[-1]), p_n=0.001)
    check_sk(gpu, oracle, s, k, m, 42, True, offsets=offsets)


def test_mutated_long_reads_config5_shape(gpu, oracle):
    """configs[4] in small: reads drawn by the reference generator, passed through the reference's sequence_mutator
    (restated in the oracle and pinned to the reference: substitutions, insertions, deletions; absent values dropped as
    tests/test_random_sequence_generation.cpp:44-48 does), k=21 m=11, one call over all reads."""
    rng = np.random.default_rng(55)
    reads = []
    for i in range(40):
        src = oracle.generate(1000 + i, int(rng.integers(2_000, 12_000)))
        out, _, _ = oracle.mutate(src, 0.1, 0.3, 0.2, 2000 + i)
        reads.append(np.asarray(out, dtype=np.uint8))
    offsets = np.concatenate([[0], np.cumsum([len(r) for r in reads])]).astype(np.uint64)
    s = np.concatenate(reads)
    assert len(s) > 100_000
    for canon in (False, True):
        check_sk(gpu, oracle, s, 21, 11, 42, canon, offsets=offsets)
    g = gpu.kmers(s, 21, True, offsets=offsets)
    o = oracle.kmers(s, 21, True, offsets=offsets)
    for key in ("lo", "pos"):
        assert np.array_equal(g[key], o[key])


def test_minimizers_entry_point(gpu, oracle):
    rng = np.random.default_rng(3)
    s = rand_seq(rng, 100000, p_n=0.001)
    g = gpu.minimizers(s, 31, 19, 42, True)
    o = oracle.super_kmers(s, 31, 19, 42, True)
    assert g["n"] == o["n"]
    assert np.array_equal(g["value"], o["value"]) and np.array_equal(g["mm_start"], o["mm_start"])
    assert np.array_equal(g["hash"], o["hash"])


def test_window_ranges_concatenate(gpu, oracle):
    rng = np.random.default_rng(4)
    s = rand_seq(rng, 200000, p_n=0.0005)
    full = gpu.super_kmers(s, 31, 19, 42, True)
    cuts = [0, 1, 777, 65536, 65537, 131072, 199000, 200000]
    parts = [gpu.super_kmers(s, 31, 19, 42, True, win_range=(a, b)) for a, b in zip(cuts[:-1], cuts[1:])]
    for key in ("value", "mm_start", "size", "mm_pos", "hash"):
        assert np.array_equal(np.concatenate([p[key] for p in parts]), full[key]), key


@pytest.mark.parametrize("k,m,world", [(31, 19, 2), (63, 31, 4), (21, 11, 8)])
def test_chunk_plan_slices_concatenate(gpu, oracle, k, m, world):
    """The multi-GPU chunk+halo plan (biolib_b200.shard) on one device: each 'rank' is given ONLY its slice."""
    from biolib_b200 import shard
    rng = np.random.default_rng(11)
    s = rand_seq(rng, 300_000, p_n=0.0004)
    o = oracle.super_kmers(s, k, m, 42, True)
    parts = []
    for pl in shard.plan_chunks(len(s), k, m, world):
        g = gpu.super_kmers(s[pl.buf_lo:pl.buf_hi], k, m, 42, True, win_range=pl.local_window)
        g["mm_start"] = g["mm_start"] + np.uint64(pl.buf_lo)
        g["sk_start"] = g["sk_start"] + np.uint64(pl.buf_lo)
        parts.append(g)
    assert sum(p["n"] for p in parts) == o["n"]
    for key in ("value", "hash", "mm_start", "sk_start", "size"):
        assert np.array_equal(np.concatenate([p[key] for p in parts]), o[key]), key


def test_sequence_plan_slices_concatenate(gpu, oracle):
    from biolib_b200 import shard
    rng = np.random.default_rng(12)
    lens = rng.integers(10, 600, 2000)
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    s = rand_seq(rng, int(offs[-1]), p_n=0.002, lower=0.01)
    o = oracle.super_kmers(s, 21, 11, 42, True, offsets=offs)
    parts = []
    for a, b in shard.plan_sequences(offs, 8):
        base = offs[a]
        g = gpu.super_kmers(s[int(base):int(offs[b])], 21, 11, 42, True, offsets=offs[a:b + 1] - base)
        g["mm_start"] = g["mm_start"] + base
        g["sk_start"] = g["sk_start"] + base
        parts.append(g)
    for key in ("value", "hash", "mm_start", "sk_start", "size"):
        assert np.array_equal(np.concatenate([p[key] for p in parts]), o[key]), key


@pytest.mark.parametrize("k,m,multi", [(31, 19, False), (21, 11, True)])
def test_pipelined_host_path_large_input(gpu, oracle, k, m, multi):
    """Inputs above 8 M k-mers take the chunked upload / kernel / download pipeline: same records as the oracle."""
    from biolib_b200 import synth
    n = 20_000_000
    s = synth.generate(77, n).copy()
    rng = np.random.default_rng(3)
    for p_ in rng.integers(0, n - 100, 200):
        s[p_:p_ + int(rng.integers(1, 40))] = ord("N")
    offs = None
    if multi:
        cuts = np.sort(rng.choice(np.arange(1, n), size=3000, replace=False))
        offs = np.concatenate([[0], cuts, [n]]).astype(np.uint64)
    g = gpu.super_kmers(s, k, m, 42, True, offsets=offs)
    assert gpu.stats()["kernel_launches"] >= 2          # it really was cut into chunks
    o = oracle.super_kmers(s, k, m, 42, True, offsets=offs)
    assert g["n"] == o["n"]
    for key in ("value", "hash", "mm_start", "sk_start", "size"):
        assert np.array_equal(g[key], o[key]), key
    # caller-provided arrays that are too small: the required count comes back with BLG_E_CAPACITY
    import biolib_b200
    buf = np.zeros(1000, np.uint64)
    buf8 = np.zeros(1000, np.uint8)
    with pytest.raises(biolib_b200.BlgError) as e:
        gpu.super_kmers_raw(s.ctypes.data, n, k, m, 42, True, buf.ctypes.data, buf.ctypes.data, buf8.ctypes.data, buf8.ctypes.data, 1000,
                            offsets_ptr=offs.ctypes.data if offs is not None else 0, n_seq=0 if offs is None else len(offs) - 1)
    assert e.value.status == 3


@pytest.mark.parametrize("k,m", [(31, 19), (21, 11), (9, 5), (32, 32), (32, 7), (12, 12), (3, 1), (32, 16)])
def test_minimizer_position_extractor(gpu, oracle, k, m):
    """blg_minimizer_positions vs hash::minimizer_position_extractor (reference include/kmer_view.hpp:266-283)."""
    rng = np.random.default_rng(80 + k)
    v = rng.integers(0, 2**63, size=100_000, dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, size=100_000, dtype=np.uint64)
    if k < 32:
        v &= np.uint64((1 << (2 * k)) - 1)
    v[:500] &= np.uint64(0xFF)
    got = gpu.minimizer_positions(v, k, m)
    assert np.array_equal(got.astype(np.uint64), oracle.minimizer_positions(v, k, m))
    # through the k-mers of a sequence: offset of the minimizer inside each canonical k-mer value
    s = rand_seq(rng, 5000)
    km = gpu.kmers(s, k, True)
    assert np.array_equal(gpu.minimizer_positions(km["lo"], k, m).astype(np.uint64), oracle.minimizer_positions(km["lo"], k, m))


def test_minimizer_position_extractor_errors(gpu):
    import biolib_b200
    for k, m in [(33, 5), (5, 6), (0, 0), (5, 0)]:
        with pytest.raises(biolib_b200.BlgError) as e:
            gpu.minimizer_positions(np.zeros(4, np.uint64), k, m)
        assert e.value.status == 1
    assert gpu.minimizer_positions(np.zeros(0, np.uint64), 31, 19).size == 0


def test_hash64(gpu, oracle):
    vals = np.array([0, 1, 0x1234567890ABCDEF, 0xFFFFFFFFFFFFFFFF], dtype=np.uint64)
    assert gpu.hash64(vals, 0).tolist() == [0x28DF63B7CC57C3CB, 0x004403B7FB05C44A, 0x495AE5FABBE9F68B, 0xA0E4B27A1ABAED73]
    assert gpu.hash64(vals, 42).tolist() == [0x434B9AFE2C3E9543, 0xD3FE46E112F04C44, 0x05E73E3913CB8728, 0xB6A86CA16A85A979]
    assert gpu.hash64(vals, 0x10000002A).tolist() == gpu.hash64(vals, 42).tolist()
    rng = np.random.default_rng(0)
    v = rng.integers(0, 2**63, size=10000, dtype=np.uint64)
    assert gpu.hash64(v, 7).tolist() == [oracle.hash64(int(x), 7) for x in v]
    # ragged sizes around the vector width and the unrolled stride of the kernel (two values per access, four accesses)
    big = rng.integers(0, 2**63, size=2_500_003, dtype=np.uint64)
    ref = gpu.hash64(big, 9)
    idx = rng.integers(0, big.size, 2000)
    assert [int(ref[i]) for i in idx] == [oracle.hash64(int(big[i]), 9) for i in idx]
    assert int(ref[-1]) == oracle.hash64(int(big[-1]), 9)
    for n in (0, 1, 2, 3, 7, 8, 9, 1001):
        assert gpu.hash64(big[:n], 9).tolist() == ref[:n].tolist()


@pytest.mark.parametrize("k", [1, 3, 15, 31, 32, 33, 47, 63, 64])
@pytest.mark.parametrize("canon", [False, True])
def test_kmers(gpu, oracle, k, canon):
    rng = np.random.default_rng(k)
    for n, pn in ((0, 0), (k - 1, 0), (k, 0), (k + 1, 0), (5000, 0.0), (70000, 0.01), (5000, 0.2)):
        s = rand_seq(rng, n, p_n=pn, lower=0.05)
        g = gpu.kmers(s, k, canon)
        o = oracle.kmers(s, k, canon)
        assert len(g["lo"]) == len(o["lo"]), (k, n)
        for key in ("lo", "hi", "pos"):
            assert np.array_equal(g[key], o[key]), (key, k, n, pn)


@pytest.mark.parametrize("k,m", [(31, 19), (21, 11), (32, 32), (32, 1), (9, 5), (3, 1), (15, 10), (32, 16)])
@pytest.mark.parametrize("canon", [False, True])
def test_syncmers(gpu, oracle, k, m, canon):
    """blg_syncmers vs sampler::syncmer_sampler over minimizer_position_extractor (reference include/syncmer_sampler.hpp:130-137,
    include/kmer_view.hpp:266-283; the restatement is pinned by tests/golden/syncmers.json)."""
    rng = np.random.default_rng(1000 * k + m)
    for n, pn in ((0, 0), (k - 1, 0), (k, 0), (k + 1, 0), (5000, 0.0), (70000, 0.01), (5000, 0.2), (300000, 0.0005)):
        s = rand_seq(rng, n, p_n=pn, lower=0.05)
        if n >= 5000:
            s[1000:1400] = ord("A")                     # equal m-mers: ties go to the leftmost
        for soff, eoff in ((0, k - m), (1 % (k - m + 1), 1 % (k - m + 1))):
            g = gpu.syncmers(s, k, m, soff, eoff, canon)
            o = oracle.syncmers(s, k, m, soff, eoff, canon)
            assert len(g["lo"]) == len(o["lo"]), (k, m, n, pn, soff)
            for key in ("lo", "pos"):
                assert np.array_equal(g[key], o[key]), (key, k, m, n, pn, soff)


def test_syncmers_multi_sequence_and_errors(gpu, oracle):
    rng = np.random.default_rng(88)
    lens = rng.integers(0, 300, size=400)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    s = rand_seq(rng, int(offsets[-1]), p_n=0.005)
    for k, m in ((31, 19), (21, 11)):
        g = gpu.syncmers(s, k, m, canonical=True, offsets=offsets)
        o = oracle.syncmers(s, k, m, 0, k - m, True, offsets=offsets)
        for key in ("lo", "pos"):
            assert np.array_equal(g[key], o[key]), (key, k)
    for k, m in ((33, 19), (31, 0), (19, 31)):
        with pytest.raises(Exception):
            gpu.syncmers("ACGT" * 100, k, m)
    # an offset no m-mer can have: nothing passes (as the reference's loop would run to the end)
    assert len(gpu.syncmers("ACGT" * 100, 11, 5, 9, 9)["lo"]) == 0


@pytest.mark.parametrize("k,canon", [(31, True), (21, False), (32, True), (5, False)])
def test_hash_sampled_kmers(gpu, oracle, k, canon):
    """blg_hash_sampled_kmers vs sampler::hash_sampler (reference include/hash_sampler.hpp:73-78,138-141; the restatement is
    pinned against the reference-executed sampler in tests/test_oracle_vs_ref.py and tests/golden/hash_sampler.json)."""
    rng = np.random.default_rng(k)
    for n, pn in ((0, 0), (k, 0), (5000, 0.0), (40000, 0.01)):
        s = rand_seq(rng, n, p_n=pn, lower=0.05)
        for seed, rate in ((42, 0.1), (0, 0.5), (2**32 + 42, 0.01), (7, 0.0), (7, 1.0)):
            thr = gpu.hash_sampler_threshold(rate)
            g = gpu.hash_sampled_kmers(s, k, seed, rate, canon)
            o = oracle.hash_sampled_kmers(s, k, seed, thr, canon)
            assert len(g["lo"]) == len(o["lo"]), (k, n, seed, rate)
            for key in ("lo", "pos"):
                assert np.array_equal(g[key], o[key]), (key, k, n, seed, rate)
    with pytest.raises(Exception):
        gpu.hash_sampled_kmers("ACGT" * 100, 33, 0, 0.5)
    with pytest.raises(ValueError):
        gpu.hash_sampled_kmers("ACGT" * 100, 31, 0, 1.5)


def test_hash_sampled_kmers_large_cross_check(gpu):
    """20 Mbp, two sequences: the fused filter equals blg_kmers filtered by blg_hash64."""
    rng = np.random.default_rng(98)
    s = rand_seq(rng, 20_000_000, p_n=0.0001)
    offsets = np.array([0, 7_000_001, 20_000_000], dtype=np.uint64)
    k, seed, rate = 31, 42, 0.125
    g = gpu.hash_sampled_kmers(s, k, seed, rate, True, offsets=offsets)
    km = gpu.kmers(s, k, True, offsets=offsets)
    keep = gpu.hash64(km["lo"], seed) < np.uint64(gpu.hash_sampler_threshold(rate))
    assert np.array_equal(g["lo"], km["lo"][keep]) and np.array_equal(g["pos"], km["pos"][keep])
    assert 0.12 < len(g["lo"]) / len(km["lo"]) < 0.13


def test_syncmers_large_cross_check(gpu):
    """20 Mbp: the fused filter equals blg_kmers filtered by the separate blg_minimizer_positions kernel."""
    rng = np.random.default_rng(99)
    s = rand_seq(rng, 20_000_000, p_n=0.0001)
    k, m = 31, 19
    g = gpu.syncmers(s, k, m, canonical=True)
    km = gpu.kmers(s, k, True)
    p = gpu.minimizer_positions(km["lo"], k, m)
    keep = (p == 0) | (p == k - m)
    assert np.array_equal(g["lo"], km["lo"][keep]) and np.array_equal(g["pos"], km["pos"][keep])
    assert 0.1 < len(g["lo"]) / len(km["lo"]) < 0.2         # closed syncmers: about 2 / (k - m + 1)


def test_kmers_multi_sequence(gpu, oracle):
    rng = np.random.default_rng(8)
    lens = rng.integers(0, 300, size=200)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    s = rand_seq(rng, int(offsets[-1]), p_n=0.005)
    for k in (31, 40):
        g = gpu.kmers(s, k, True, offsets=offsets)
        o = oracle.kmers(s, k, True, offsets=offsets)
        for key in ("lo", "hi", "pos"):
            assert np.array_equal(g[key], o[key]), (key, k)


@pytest.mark.parametrize("k,canon", [(31, True), (63, True), (17, False)])
def test_kmers_many_tiles(gpu, oracle, k, canon):
    """Hundreds of tiles of the k-mer kernel (clean tiles, tiles with N runs, read boundaries, a ragged end) against the
    oracle."""
    rng = np.random.default_rng(900 + k)
    n = 2_000_003
    s = rand_seq(rng, n)
    for p in rng.integers(0, n - 3000, 40):
        s[p:p + int(rng.integers(1, 2500))] = ord("N")
    lens = rng.integers(0, 40000, size=200)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    offsets = offsets[offsets < n]
    offsets = np.concatenate([offsets, [n]]).astype(np.uint64)
    for offs in (None, offsets):
        g = gpu.kmers(s, k, canon, offsets=offs)
        o = oracle.kmers(s, k, canon, offsets=offs)
        assert len(g["lo"]) == len(o["lo"]) > 1_000_000
        for key in ("lo", "hi", "pos"):
            assert np.array_equal(g[key], o[key]), (key, k)


def test_kmers_device_input_small_buffers(gpu, oracle):
    """device-resident input of 16 .. 64 KB and of exactly 2^k bytes (the next-tile prefetch of kmer2_kernel
    used to read past such buffers); k-mers and syncmers."""
    import torch
    rng = np.random.default_rng(4242)
    for n in (16_384, 20_000, 32_768, 40_961, 65_536):
        s = rand_seq(rng, n, p_n=0.001)
        d = torch.from_numpy(s).cuda()
        cap = n
        lo = torch.empty(cap, dtype=torch.int64, device="cuda"); pos = torch.empty_like(lo)
        for canon in (False, True):
            cnt = gpu.kmers_device(d.data_ptr(), n, 31, canon, lo.data_ptr(), 0, pos.data_ptr(), cap)
            o = oracle.kmers(s, 31, canon)
            assert cnt == len(o["lo"])
            assert np.array_equal(lo[:cnt].cpu().numpy().view(np.uint64), o["lo"]) and np.array_equal(pos[:cnt].cpu().numpy().view(np.uint64), o["pos"])
            cnt = gpu.kmers_device(d.data_ptr(), n, 31, canon, lo.data_ptr(), 0, pos.data_ptr(), cap, syncmer=(19, 0, 12))
            o = oracle.syncmers(s, 31, 19, 0, 12, canon)
            assert cnt == len(o["lo"])
            assert np.array_equal(lo[:cnt].cpu().numpy().view(np.uint64), o["lo"]) and np.array_equal(pos[:cnt].cpu().numpy().view(np.uint64), o["pos"])


def test_config1_kmer_view_1mbp(gpu, oracle):
    """BASELINE config 1: kmer_view k=31 canonical over the 1 Mbp sequence_generator(42) sequence."""
    s = oracle.generate(42, 1_000_000)
    g = gpu.kmers(s, 31, True)
    assert len(g["lo"]) == 1_000_000 - 31 + 1
    # the reference's iterator drops the last k-mer (K1): its 999 969 values have these checksums
    v = g["lo"][:-1]
    assert int(np.bitwise_xor.reduce(v)) == 4314334728908509064
    assert int(v.sum(dtype=np.uint64)) == 8341584218946238560
    assert int(v[0]) == 2246878431356894220 and int(v[-1]) == 1070654462582993322


def test_errors(gpu):
    import biolib_b200
    with pytest.raises(biolib_b200.BlgError) as e:
        gpu.super_kmers("ACGT", 3, 4, 0, False)
    assert e.value.status == biolib_b200._capi.BLG_E_INVALID
    with pytest.raises(biolib_b200.BlgError):
        gpu.super_kmers("ACGT", 65, 4, 0, False)
    with pytest.raises(biolib_b200.BlgError):
        gpu.super_kmers("ACGT", 40, 33, 0, False)
    with pytest.raises(biolib_b200.BlgError):
        gpu.kmers("ACGT", 0, False)
    with pytest.raises(biolib_b200.BlgError):
        gpu.super_kmers("ACGTACGT", 3, 2, 0, False, offsets=np.array([0, 9], dtype=np.uint64))


def oracle_ids(oracle, seq, m, seed, canon, mm_start, offsets=None):
    """D9 from the oracle's per-position validity: valid m-mer starts before mm_start inside its sequence."""
    valid = oracle.mmers(seq, m, seed, canon, offsets=offsets)["valid"].astype(np.int64)
    excl = np.concatenate([[0], np.cumsum(valid)])
    ids = excl[mm_start.astype(np.int64)]
    if offsets is not None:
        o = np.asarray(offsets, dtype=np.int64)
        q = np.searchsorted(o, mm_start.astype(np.int64), side="right") - 1
        ids = ids - excl[o[q]]
    return ids.astype(np.uint64)


@pytest.mark.parametrize("k,m", [(31, 19), (21, 11), (63, 31), (9, 5), (33, 1), (64, 32)])
def test_minimizer_ids(gpu, oracle, k, m):
    rng = np.random.default_rng(k + 3 * m)
    for n, pn in ((k, 0), (5000, 0.0), (70001, 0.01), (200000, 0.0005), (4000, 0.2)):
        s = rand_seq(rng, n, p_n=pn)
        for entry in (gpu.super_kmers, gpu.minimizers):
            g = entry(s, k, m, 42, True, want_id=True)
            assert np.array_equal(g["id"], oracle_ids(oracle, s, m, 42, True, g["mm_start"])), (k, m, n, pn)
    lens = rng.integers(0, 500, size=400)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    s = rand_seq(rng, int(offsets[-1]), p_n=0.003)
    g = gpu.super_kmers(s, k, m, 42, False, offsets=offsets, want_id=True)
    o = oracle.super_kmers(s, k, m, 42, False, offsets=offsets)
    assert np.array_equal(g["mm_start"], o["mm_start"])
    assert np.array_equal(g["id"], oracle_ids(oracle, s, m, 42, False, g["mm_start"], offsets))


CPP_DRIVER = r"""
#include "biolib_gpu.hpp"
namespace hash { class hash64 {}; }   // the reference's include/hash.hpp:45 defines it; the shim only needs the tag
#include <cstdio>
#include <fstream>
#include <iterator>
#include <string>
int main(int argc, char** argv) {
    std::ifstream f(argv[1], std::ios::binary);
    std::string s((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    const int k = std::atoi(argv[2]), m = std::atoi(argv[3]), canon = std::atoi(argv[4]);
    auto kv = biolib::gpu::kmer_view_from_cstr<uint64_t>(s.data(), s.size(), k, canon);
    for (auto it = kv.cbegin(); it != kv.cend(); ++it) {
        auto r = *it;
        std::printf("K %d %llu %zu %zu\n", r.value ? 1 : 0, (unsigned long long)(r.value ? *r.value : 0), r.position, r.id);
    }
    auto mv = biolib::gpu::minimizer_view_from_string<uint64_t, uint64_t, hash::hash64>(s, k, m, 42, canon);
    if (mv.get_k() != k || mv.get_m() != m) return 2;
    for (auto const& r : mv) std::printf("M %llu %zu %zu\n", (unsigned long long)r.value, r.position, r.id);
    biolib::gpu::super_kmer_view<uint64_t, uint64_t, hash::hash64> sk(s, k, m, canon, 42);
    for (auto it = sk.cbegin(); it != sk.cend(); ++it)
        std::printf("S %llu %zu %zu %u %u\n", (unsigned long long)(*it).minimizer.value, (*it).minimizer.position, (*it).minimizer.id,
                    (unsigned)(*it).mm_pos, (unsigned)(*it).size);
    {
        // streaming form: smallest records over PCIe, positions rebuilt by the forward iterator, id = the record's ordinal
        biolib::gpu::super_kmer_view<uint64_t, uint64_t, hash::hash64> st(s, k, m, canon, 42, false);
        for (auto it = st.cbegin(); it != st.cend(); ++it)
            std::printf("T %llu %zu %zu %u %u\n", (unsigned long long)(*it).minimizer.value, (*it).minimizer.position, (*it).minimizer.id,
                        (unsigned)(*it).mm_pos, (unsigned)(*it).size);
    }
    {
        biolib::gpu::minimizer_position_extractor ex(k, m);
        std::vector<uint64_t> vals;
        for (auto it = kv.cbegin(); it != kv.cend(); ++it) if ((*it).value) vals.push_back(*(*it).value);
        auto pos = ex(vals.data(), vals.size());
        for (std::size_t i = 0; i < vals.size(); ++i) std::printf("X %llu %u\n", (unsigned long long)vals[i], (unsigned)pos[i]);
        biolib::gpu::kmer_context_t<uint64_t> none{std::nullopt, 0, 0};
        if (ex(none) != (std::size_t)k + 1 || ex.get_k() != k || ex.get_m() != m) return 5;
        if (!vals.empty()) { biolib::gpu::kmer_context_t<uint64_t> one{vals[0], 0, 0}; if (ex(one) != pos[0]) return 6; }
    }
    if (k <= 32) {
        biolib::gpu::minimizer_position_extractor ex(k, m);
        biolib::gpu::syncmer_sampler<std::string::const_iterator> sy(s.cbegin(), s.cend(), ex, 0, k - m, canon);
        if (sy.get_offsets() != std::make_pair<uint16_t, uint16_t>(0, k - m)) return 7;
        for (auto const& r : sy) std::printf("Y %llu %zu\n", (unsigned long long)*r.value, r.position);
    }
    if (k <= 32) {
        biolib::gpu::hash_sampler<std::string::const_iterator, hash::hash64> hs(s.cbegin(), s.cend(), hash::hash64(), 42, 0.25, k, canon);
        if (hs.get_sampling_rate() != 0.25) return 8;
        for (auto const& r : hs) std::printf("H %llu %zu %zu\n", (unsigned long long)*r.value, r.position, r.id);
        try { biolib::gpu::hash_sampler<std::string::const_iterator, hash::hash64> bad(s.cbegin(), s.cend(), hash::hash64(), 42, 1.5, k); return 9; }
        catch (std::invalid_argument const&) {}
    }
    try { biolib::gpu::super_kmer_view<uint64_t, uint64_t, hash::hash64> bad(s, 3, 4); bad.cbegin(); return 4; }
    catch (std::invalid_argument const&) {}
    return 0;
}
"""


def test_cpp_shim_views(gpu, oracle, tmp_path):
    """The reference-facing C++ surface (include/biolib_gpu.hpp): record streams of the three views."""
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = tmp_path / "drv.cpp"
    src.write_text(CPP_DRIVER)
    exe = tmp_path / "drv"
    libdir = os.path.join(root, "biolib_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-I", os.path.join(root, "include"), str(src), "-o", str(exe),
                           "-L", libdir, "-lbiolib_gpu", f"-Wl,-rpath,{libdir}"])
    rng = np.random.default_rng(21)
    cases = [("ACGTNNNACGT", 3, 2), ("ACGTACGT", 3, 2), ("ACGTNA", 3, 3), ("ACGTNNNACG", 3, 1), ("AC", 3, 2), ("NNNN", 3, 2)]
    cases += [(rand_seq(rng, 3000, p_n=0.01, lower=0.05).tobytes().decode(), 31, 19),
              (rand_seq(rng, 2000, p_n=0.0).tobytes().decode(), 21, 11),
              (rand_seq(rng, 500, p_n=0.1).tobytes().decode(), 7, 4)]
    for seq, k, m in cases:
        for canon in (0, 1):
            f = tmp_path / "seq.txt"
            f.write_bytes(seq.encode())
            out = subprocess.check_output([str(exe), str(f), str(k), str(m), str(canon)]).decode().split("\n")
            K = [tuple(map(int, l.split()[1:])) for l in out if l.startswith("K ")]
            M = [tuple(map(int, l.split()[1:])) for l in out if l.startswith("M ")]
            S = [tuple(map(int, l.split()[1:])) for l in out if l.startswith("S ")]
            ks = oracle.kmer_stream(seq, k, canon)
            assert K == list(zip(ks["has"].tolist(), ks["lo"].tolist(), ks["pos"].tolist(), ks["id"].tolist())), (seq[:20], k, canon)
            o = oracle.super_kmers(seq, k, m, 42, canon)
            ids = oracle_ids(oracle, seq, m, 42, canon, o["mm_start"]).tolist() if o["n"] else []
            want_m = list(zip(o["value"].tolist(), (o["mm_start"] + 1).tolist(), ids))       # D7: position = start + 1
            assert M == want_m, (seq[:20], k, m, canon)
            assert S == [a + (int(p), int(z)) for a, p, z in zip(want_m, (o["mm_start"] - o["sk_start"]).tolist(), o["size"].tolist())]
            T = [tuple(map(int, l.split()[1:])) for l in out if l.startswith("T ")]
            assert T == [(v, p, i, mp, z) for i, ((v, p, _), mp, z) in enumerate(zip(want_m, (o["mm_start"] - o["sk_start"]).tolist(), o["size"].tolist()))], (seq[:20], k, m, canon)
            Y = [tuple(map(int, l.split()[1:])) for l in out if l.startswith("Y ")]
            sy = oracle.syncmer_stream(seq, k, m, 0, k - m, canon)
            assert Y == list(zip(sy["lo"].tolist(), sy["pos"].tolist())), (seq[:20], k, m, canon)
            H = [tuple(map(int, l.split()[1:])) for l in out if l.startswith("H ")]
            if k <= 32:
                thr = gpu.hash_sampler_threshold(0.25)
                valid = ks["has"] == 1                                              # kmer_view's stream already applies K1
                keep = np.array([oracle.hash64(int(v), 42) < thr for v in ks["lo"][valid]], dtype=bool)
                assert H == list(zip(ks["lo"][valid][keep].tolist(), ks["pos"][valid][keep].tolist(), range(int(keep.sum())))), (seq[:20], k, canon)
            X = [tuple(map(int, l.split()[1:])) for l in out if l.startswith("X ")]
            xv = np.array([v for v, _ in X], dtype=np.uint64)
            assert [p for _, p in X] == oracle.minimizer_positions(xv, k, m).tolist() if len(X) else True


@pytest.mark.parametrize("k,m", [(31, 19), (21, 11), (32, 20), (9, 5), (16, 8), (32, 1), (13, 12), (28, 12), (32, 17)])
def test_d4_window_argmin_vs_reference_extractor(gpu, oracle, k, m):
    """D4 of the fused kernels (sk4 / sk3 / sk2, whichever serves (k, m)) against hash::minimizer_position_extractor
    (reference include/kmer_view.hpp:266-283): the reference-compiled library when it travelled to this box
    (oracle/_ref), else the oracle restatement that tests/test_oracle_vs_ref.py pins to it."""
    import pyoracle
    from d4_util import check_d4
    ext = pyoracle.Ref() if pyoracle.Ref.available() else oracle
    rng = np.random.default_rng(k * 41 + m)
    low = np.concatenate([np.frombuffer(b"A" * 2000, np.uint8), np.frombuffer(b"AC" * 1500, np.uint8), rand_seq(rng, 3000),
                          np.frombuffer(b"ACGTTGCA" * 600, np.uint8), np.frombuffer(b"GATTACA" * 700, np.uint8)])
    for s in (rand_seq(rng, 60_000), rand_seq(rng, 30_000, 0.002, 0.05), low):
        rec = gpu.super_kmers(s, k, m, 0, False)
        km = gpu.kmers(s, k, False)
        check_d4(rec, km, np.asarray(ext.minimizer_positions(km["lo"], k, m)))


@pytest.mark.parametrize("k,m", [(31, 19), (21, 11), (63, 31), (40, 9), (8, 4)])
def test_compact_records_pos32(gpu, oracle, k, m):
    """BLG_F_POS32: 32-bit positions relative to the range start, every kernel (sk4 / sk3 / sk2), plain and pipelined host
    paths, whole buffer and window ranges; identical to the 64-bit layout."""
    rng = np.random.default_rng(k * 3 + m)
    s = rand_seq(rng, 300_000, 0.0005, 0.02)
    for canon in (False, True):
        a = gpu.super_kmers(s, k, m, 42, canon)
        b = gpu.super_kmers(s, k, m, 42, canon, pos32=True)
        for key in ("value", "hash", "mm_start", "sk_start", "size", "mm_pos"):
            assert np.array_equal(a[key], b[key]), (key, k, m)
    lo, hi = 100_003, 250_000
    a = gpu.super_kmers(s, k, m, 42, True, win_range=(lo, hi))
    b = gpu.super_kmers(s, k, m, 42, True, win_range=(lo, hi), pos32=True)
    assert a["n"] == b["n"] > 0 and np.array_equal(a["mm_start"], b["mm_start"]) and np.array_equal(a["size"], b["size"])
    o = oracle.super_kmers(s, k, m, 42, True)
    sel = (o["sk_start"] >= lo) & (o["sk_start"] < hi)
    assert np.array_equal(b["mm_start"], o["mm_start"][sel]) and np.array_equal(b["value"], o["value"][sel])


def test_compact_records_pos32_pipelined(gpu):
    """The chunked host path (>= 8 Mi windows) with the compact layout against the 64-bit layout."""
    rng = np.random.default_rng(77)
    s = rand_seq(rng, 20_000_000, 0.0001)
    a = gpu.super_kmers(s, 31, 19, 42, True, want_hash=False)
    b = gpu.super_kmers(s, 31, 19, 42, True, want_hash=False, pos32=True)
    assert a["n"] == b["n"] > 2_000_000
    for key in ("value", "mm_start", "size", "mm_pos"):
        assert np.array_equal(a[key], b[key]), key
    st = gpu.stats()
    assert st["bytes_d2h"] < a["n"] * 14 + 4096


@pytest.mark.parametrize("k,m", [(31, 19), (21, 11), (63, 31), (40, 9), (8, 4)])
def test_runs_layout(gpu, oracle, k, m):
    """BLG_F_RUNS: 10-byte records + the run list; positions rebuilt from the runs equal the 64-bit layout (plain host path)."""
    rng = np.random.default_rng(k * 5 + m)
    s = rand_seq(rng, 200_000, 0.0008, 0.02)
    lens = rng.integers(0, 3000, size=120)
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    offs = np.concatenate([offs[offs < s.size], [s.size]]).astype(np.uint64)
    for canon in (False, True):
        for o in (None, offs):
            a = gpu.super_kmers(s, k, m, 42, canon, offsets=o, want_hash=False)
            b = gpu.super_kmers(s, k, m, 42, canon, offsets=o, want_hash=False, runs=True)
            assert a["n"] == b["n"] and 0 < b["n_runs"] <= b["n"]
            for key in ("value", "mm_start", "sk_start", "size", "mm_pos"):
                assert np.array_equal(a[key], b[key]), (key, k, m)
    clean = rand_seq(rng, 50_000)
    assert gpu.super_kmers(clean, k, m, 42, True, want_hash=False, runs=True)["n_runs"] == 1
    assert gpu.super_kmers(np.zeros(0, np.uint8), k, m, 42, True, want_hash=False, runs=True)["n"] == 0


def test_runs_layout_pipelined(gpu):
    """The chunked host path with BLG_F_RUNS: 10 bytes per record over PCIe."""
    rng = np.random.default_rng(78)
    s = rand_seq(rng, 20_000_000, 0.00005)
    a = gpu.super_kmers(s, 31, 19, 42, True, want_hash=False)
    b = gpu.super_kmers(s, 31, 19, 42, True, want_hash=False, runs=True)
    st = gpu.stats()
    assert a["n"] == b["n"] > 2_000_000 and b["n_runs"] > 100
    for key in ("value", "mm_start", "size", "mm_pos"):
        assert np.array_equal(a[key], b[key]), key
    assert st["bytes_d2h"] < a["n"] * 10 + b["n_runs"] * 16 + 4096


@pytest.mark.parametrize("k,canon", [(31, True), (63, True), (5, False), (32, False)])
def test_kmers_runs_layout(gpu, oracle, k, canon):
    """BLG_F_RUNS for blg_kmers: values + the list of stretches of consecutive positions instead of a position per k-mer."""
    import biolib_b200
    rng = np.random.default_rng(k)
    s = rand_seq(rng, 120_000, 0.0006, 0.02)
    offs = np.concatenate([[0], np.cumsum(rng.integers(0, 2500, size=90))]).astype(np.uint64)
    offs = np.concatenate([offs[offs < s.size], [s.size]]).astype(np.uint64)
    for o in (None, offs):
        a = gpu.kmers(s, k, canon, offsets=o)
        b = gpu.kmers(s, k, canon, offsets=o, runs=True)
        st = gpu.stats()
        assert len(a["lo"]) == len(b["lo"]) > 0 and 0 < b["n_runs"] < len(b["lo"])
        for key in ("lo", "hi", "pos"):
            assert np.array_equal(a[key], b[key]), (key, k)
        assert st["bytes_d2h"] <= len(a["lo"]) * (16 if k > 32 else 8) + b["n_runs"] * 16 + 64
    ok = oracle.kmers(s, k, canon, offsets=offs)
    assert np.array_equal(b["lo"], ok["lo"]) and np.array_equal(b["pos"], ok["pos"])
    clean = rand_seq(rng, 40_000)
    assert gpu.kmers(clean, k, canon, runs=True)["n_runs"] == 1
    assert len(gpu.kmers(np.zeros(0, np.uint8), k, canon, runs=True)["lo"]) == 0
    with pytest.raises(biolib_b200.BlgError):
        gpu.kmers(clean, 31, canon, syncmer=(19, 0, 12), runs=True)


def test_kmers_runs_layout_pipelined(gpu):
    """The chunked host path of blg_kmers with BLG_F_RUNS: half the download; runs that cross a piece boundary are joined."""
    rng = np.random.default_rng(80)
    s = rand_seq(rng, 30_000_000, 0.00002)
    a = gpu.kmers(s, 31, True)
    b = gpu.kmers(s, 31, True, runs=True)
    st = gpu.stats()
    assert len(a["lo"]) == len(b["lo"]) > 25_000_000
    assert np.array_equal(a["lo"], b["lo"]) and np.array_equal(a["pos"], b["pos"])
    # maximal runs: no run starts where the previous one would have continued
    nb = int(np.count_nonzero(np.diff(a["pos"].astype(np.int64)) != 1)) + 1
    assert b["n_runs"] == nb
    assert st["bytes_d2h"] < len(a["lo"]) * 8 + b["n_runs"] * 16 + 4096


@pytest.mark.parametrize("k,m", [(31, 19), (21, 11), (31, 16), (35, 20), (8, 4), (16, 1)])
def test_tight_layout(gpu, oracle, k, m):
    """BLG_F_TIGHT: the BLG_F_RUNS records in 6 bytes (5 for m <= 16): value_lo32 / value_hi8 / (size - 1) | mm_pos << 4."""
    rng = np.random.default_rng(k * 7 + m)
    s = rand_seq(rng, 150_000, 0.0008, 0.02)
    offs = np.concatenate([[0], np.cumsum(rng.integers(0, 3000, size=90))]).astype(np.uint64)
    offs = np.concatenate([offs[offs < s.size], [s.size]]).astype(np.uint64)
    for canon in (False, True):
        for o in (None, offs):
            a = gpu.super_kmers(s, k, m, 42, canon, offsets=o, want_hash=False)
            b = gpu.super_kmers(s, k, m, 42, canon, offsets=o, want_hash=False, runs=True, tight=True)
            st = gpu.stats()
            assert a["n"] == b["n"] > 0
            for key in ("value", "mm_start", "sk_start", "size", "mm_pos"):
                assert np.array_equal(a[key], b[key]), (key, k, m)
            assert st["bytes_d2h"] <= a["n"] * (6 if m > 16 else 5) + b["n_runs"] * 16 + 64     # + the count words of the kernels that copy them
    o = oracle.super_kmers(s, k, m, 42, True)
    assert np.array_equal(b["value"], oracle.super_kmers(s, k, m, 42, True, offsets=offs)["value"]) and o["n"] > 0
    assert gpu.super_kmers(np.zeros(0, np.uint8), k, m, 42, True, want_hash=False, runs=True, tight=True)["n"] == 0


def test_tight_layout_rejects_what_does_not_fit(gpu):
    import biolib_b200
    s = rand_seq(np.random.default_rng(5), 5000)
    for k, m, kw in ((31, 21, dict(runs=True)), (40, 19, dict(runs=True)), (31, 19, dict(runs=False)), (31, 19, dict(runs=True, want_hash=True))):
        with pytest.raises(biolib_b200.BlgError):
            gpu.super_kmers(s, k, m, 42, True, tight=True, **{"want_hash": False, **kw})


def test_tight_layout_pipelined(gpu):
    """The chunked host path with BLG_F_TIGHT: 6 bytes per record over PCIe, packed piece by piece on the device."""
    rng = np.random.default_rng(79)
    s = rand_seq(rng, 20_000_000, 0.00005)
    a = gpu.super_kmers(s, 31, 19, 42, True, want_hash=False)
    b = gpu.super_kmers(s, 31, 19, 42, True, want_hash=False, runs=True, tight=True)
    st = gpu.stats()
    assert a["n"] == b["n"] > 2_000_000 and b["n_runs"] > 100
    for key in ("value", "mm_start", "size", "mm_pos"):
        assert np.array_equal(a[key], b[key]), key
    assert st["bytes_d2h"] < a["n"] * 6 + b["n_runs"] * 16 + 4096


@pytest.mark.parametrize("k,canon,multi", [(31, True, False), (63, True, True), (21, False, True)])
def test_kmers_pipelined_host_path(gpu, oracle, k, canon, multi):
    """blg_kmers / blg_syncmers / blg_hash_sampled_kmers above 8 Mi k-mer starts: upload, kernel and download overlap piece by
    piece through two alternating device buffers; the result equals the oracle's (and the small-input path's) records."""
    import biolib_b200
    from biolib_b200 import synth
    n = 30_000_000
    s = synth.generate(79 + k, n).copy()
    rng = np.random.default_rng(k)
    for p_ in rng.integers(0, n - 100, 300):
        s[p_:p_ + int(rng.integers(1, 40))] = ord("N")
    offs = None
    if multi:
        offs = np.concatenate([[0], np.sort(rng.choice(np.arange(1, n), size=2000, replace=False)), [n]]).astype(np.uint64)
    g = gpu.kmers(s, k, canon, offsets=offs)
    st = gpu.stats()
    assert st["kernel_launches"] >= 2 and st["bytes_h2d"] >= n
    o = oracle.kmers(s, k, canon, offsets=offs)
    assert len(g["lo"]) == len(o["lo"])
    for key in ("lo", "hi", "pos"):
        assert np.array_equal(g[key], o[key]), key
    if k <= 32:
        m = k - 12
        gs = gpu.syncmers(s, k, m, 0, k - m, canon, offsets=offs)
        assert gpu.stats()["kernel_launches"] >= 2
        os_ = oracle.syncmers(s, k, m, 0, k - m, canon, offsets=offs)
        assert np.array_equal(gs["lo"], os_["lo"]) and np.array_equal(gs["pos"], os_["pos"])
        thr = gpu.hash_sampler_threshold(0.05)
        gh = gpu.hash_sampled_kmers(s, k, 42, 0.05, canon, offsets=offs)
        oh = oracle.hash_sampled_kmers(s, k, 42, thr, canon, offsets=offs)
        assert np.array_equal(gh["lo"], oh["lo"]) and np.array_equal(gh["pos"], oh["pos"])


@pytest.mark.parametrize("k,m", [(31, 19), (21, 11), (32, 20), (32, 22), (11, 1), (13, 1), (15, 5), (16, 4), (17, 7), (28, 16), (29, 17), (27, 17)])
def test_syncmers_sk4_frame(gpu, oracle, k, m):
    """k - m = 12 or 10: the syncmer filter in the sk4 frame (sy4_kernel.cu), every word-width variant, both strands, arbitrary offset
    pairs, ties (low complexity: the exact redo of both rows), N runs, many sequences, k < 16 (full strand compare)."""
    rng = np.random.default_rng(7000 + 100 * k + m)
    w = k - m + 1
    s = rand_seq(rng, 400_000, 0.0005, 0.03)
    s[5000:9000] = ord("A")
    s[20000:24000] = np.frombuffer(b"AT" * 2000, np.uint8)            # reverse-complement palindromes: strands tie
    s[30000:33000] = np.frombuffer(b"ACG" * 1000, np.uint8)
    lens = rng.integers(0, 2500, size=300)
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    offs = np.concatenate([offs[offs < s.size], [s.size]]).astype(np.uint64)
    for canon in (True, False):
        for soff, eoff in ((0, w - 1), (3, 7), (w - 1, w - 1), (w, 0), (w + 5, w + 9)):
            for o in (None, offs):
                g = gpu.syncmers(s, k, m, soff, eoff, canon, offsets=o)
                r = oracle.syncmers(s, k, m, soff, eoff, canon, offsets=o)
                assert len(g["lo"]) == len(r["lo"]), (k, m, canon, soff, eoff, o is not None)
                assert np.array_equal(g["lo"], r["lo"]) and np.array_equal(g["pos"], r["pos"]), (k, m, canon, soff, eoff)
    for n in (0, k - 1, k, k + 1, 700, 733, 1500):
        t = rand_seq(rng, n)
        g = gpu.syncmers(t, k, m, 0, w - 1, True)
        r = oracle.syncmers(t, k, m, 0, w - 1, True)
        assert np.array_equal(g["lo"], r["lo"]) and np.array_equal(g["pos"], r["pos"]), (k, m, n)
