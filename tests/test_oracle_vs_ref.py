"""CPU: oracle.c (Tier B restatement) against the reference's own code executed live (Tier A, oracle/_ref).

oracle/_ref/libbiolib_ref.so is compiled from the sources where they lie under /root/reference (oracle/Makefile) and
travels to the GPU box as a built file; if it is absent these tests skip and tests/test_oracle_golden.py (committed
vectors from the same library) still pins the oracle.
"""
import numpy as np
import pytest


def rand_seq(rng, n, p_n=0.0, lower=0.0, junk=0.0):
    s = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=n).copy()
    if p_n:
        s[rng.random(n) < p_n] = ord("N")
    if lower:
        mk = rng.random(n) < lower
        s[mk] |= 0x20
    if junk:
        mk = rng.random(n) < junk
        s[mk] = rng.integers(0, 128, size=int(mk.sum()), dtype=np.uint8)
    return s


def test_nt4_and_hash(oracle, ref):
    assert np.array_equal(oracle.nt4_table(), ref.nt4_table())
    rng = np.random.default_rng(1)
    for v in rng.integers(0, 2**63, size=3000, dtype=np.uint64).tolist() + [0, 1, 2**64 - 1]:
        for seed in (0, 42, 2**32 + 42, 2**64 - 1):
            assert oracle.hash64(v, seed) == ref.hash64(v, seed)
    for v in rng.integers(0, 2**63, size=200, dtype=np.uint64).tolist():
        assert oracle.double_hash64(v, 7) == ref.double_hash64(v, 7)
        assert oracle.double_hash64_u128(v, v ^ 0x55, 7) == ref.double_hash64_u128(v, v ^ 0x55, 7)
        assert oracle.remix(v) == ref.remix(v)


@pytest.mark.parametrize("k", [1, 2, 3, 7, 15, 16, 21, 31, 32])
def test_kmer_view_u64(oracle, ref, k):
    rng = np.random.default_rng(100 + k)
    for n, pn in ((0, 0), (k - 1, 0), (k, 0), (k + 1, 0), (2 * k, 0), (400, 0.0), (400, 0.02), (2000, 0.005), (300, 0.2)):
        for trial in range(3):
            s = rand_seq(rng, max(n, 0), pn, 0.05, 0.002)
            for canon in (0, 1):
                if canon and k < 32:
                    # K5: with zero-initialised buffers `rev` carries no garbage, any k is fine
                    pass
                a, b = oracle.kmer_stream(s, k, canon), ref.kmer_stream(s, k, canon)
                for key in ("lo", "has", "pos", "id"):
                    assert np.array_equal(a[key], b[key]), (key, k, n, pn, canon)


@pytest.mark.parametrize("k", [33, 40, 63, 64])
def test_kmer_view_u128_non_canonical(oracle, ref, k):
    """K7: for k > 32 the reference is right only when not canonical; that is what is pinned."""
    rng = np.random.default_rng(k)
    for n, pn in ((k, 0), (k + 1, 0), (500, 0.0), (800, 0.01)):
        s = rand_seq(rng, n, pn, 0.05)
        a, b = oracle.kmer_stream(s, k, 0), ref.kmer_stream(s, k, 0, bits=128)
        for key in ("lo", "hi", "has", "pos", "id"):
            assert np.array_equal(a[key], b[key]), (key, k, n)


def test_kmer_view_u128_canonical_reference_is_wrong(oracle, ref):
    """Documents K7: `(3ULL ^ c) << shift` with shift >= 64 (include/kmer_view.hpp:195,223)."""
    rng = np.random.default_rng(9)
    s = rand_seq(rng, 300)
    a, b = oracle.kmer_stream(s, 41, 1), ref.kmer_stream(s, 41, 1, bits=128)
    assert len(a["lo"]) == len(b["lo"])
    assert not (np.array_equal(a["lo"], b["lo"]) and np.array_equal(a["hi"], b["hi"]))


@pytest.mark.parametrize("k,m", [(7, 4), (31, 19), (21, 11), (63, 31), (64, 32), (15, 10), (8, 4), (12, 6), (5, 5), (33, 1)])
def test_mmer_stream(oracle, ref, k, m):
    rng = np.random.default_rng(k * 64 + m)
    for n, pn in ((0, 0), (m - 1, 0), (m, 0), (300, 0), (1500, 0.01), (400, 0.1)):
        s = rand_seq(rng, max(n, 0), pn, 0.05, 0.003)
        for canon in (0, 1):
            r = ref.mmer_stream(s, k, m, 0x10000002A, canon)
            o = oracle.mmers(s, m, 42, canon)
            ends = np.nonzero(r["valid"])[0]
            starts = ends + 1 - m
            assert int(o["valid"].sum()) == len(ends)
            assert np.array_equal(r["idx"][ends], (starts + 1).astype(np.uint64))
            assert np.array_equal(o["value"][starts], r["value"][ends]) and np.array_equal(o["hash"][starts], r["hash"][ends])


def test_minimizer_view_literal_is_empty(ref):
    """F2, kept as a one-line regression fact: the reference's own minimizer_view yields no record."""
    rng = np.random.default_rng(0)
    s = rand_seq(rng, 2000)
    for k, m in ((7, 4), (31, 19), (21, 11), (15, 10)):
        assert ref.minimizer_view_literal_count(s, k, m, 42, 0) == 0
        assert ref.minimizer_view_literal_count(s, k, m, 42, 1) == 0


@pytest.mark.parametrize("k,m", [(31, 19), (21, 11), (63, 31), (9, 5), (16, 8), (64, 1), (64, 32), (40, 9), (33, 1)])
def test_super_kmers_linear_vs_brute_vs_reference_arithmetic(oracle, ref, k, m):
    rng = np.random.default_rng(k + m)
    for n, pn in ((3000, 0.0), (6000, 0.004), (2000, 0.05)):
        s = rand_seq(rng, n, pn, 0.05)
        for canon in (0, 1):
            a = oracle.super_kmers(s, k, m, 42, canon)
            b = oracle.super_kmers(s, k, m, 42, canon, brute=True)
            assert a["n"] == b["n"]
            for key in ("value", "hash", "mm_start", "sk_start", "size"):
                assert np.array_equal(a[key], b[key]), key
            # the bench's CPU arm: reference read_base()+hash64 verbatim + deque windowing
            cnt, sums = ref.super_kmers_sums(s, k, m, 42, canon)
            assert cnt == a["n"]
            assert int(sums[0]) == int(a["value"].sum(dtype=np.uint64)) and int(sums[1]) == int(a["mm_start"].sum(dtype=np.uint64))
            assert int(sums[2]) == int(a["size"].sum(dtype=np.uint64)) and int(sums[3]) == int(a["hash"].sum(dtype=np.uint64))


def test_super_kmers_multi_sequence_equals_per_sequence(oracle):
    rng = np.random.default_rng(77)
    lens = rng.integers(0, 300, size=120)
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    s = rand_seq(rng, int(offs[-1]), 0.004)
    for brute in (False, True):
        whole = oracle.super_kmers(s, 21, 11, 42, 1, offsets=offs, brute=brute)
        parts = {key: [] for key in ("value", "hash", "mm_start", "sk_start", "size")}
        for a, b in zip(offs[:-1], offs[1:]):
            a, b = int(a), int(b)
            r = oracle.super_kmers(s[a:b], 21, 11, 42, 1)
            for key in parts:
                parts[key].append(r[key] + (np.uint64(a) if key in ("mm_start", "sk_start") else 0))
        for key in parts:
            assert np.array_equal(np.concatenate(parts[key]).astype(whole[key].dtype), whole[key]), key


def test_generator_and_mutator(oracle, ref):
    for seed in (0, 42, 2**32 + 42):
        assert np.array_equal(oracle.generate(seed, 5000), ref.generate(seed, 5000))
    src = ref.generate(5, 3000)
    for seed, (mp, idf, ep) in enumerate(((0.1, 0.3, 0.3), (0.5, 0.5, 0.5), (1.0, 1.0, 0.9), (0.02, 0.0, 0.0), (0.0, 0.5, 0.5))):
        a, ca, ta = oracle.mutate(src, mp, idf, ep, seed)
        b, cb, tb = ref.mutate(src, mp, idf, ep, seed)
        assert np.array_equal(a, b) and np.array_equal(ca, cb) and ta == tb


def test_minimizer_position_extractor_live(oracle, ref):
    rng = np.random.default_rng(8)
    for k, m in [(31, 19), (21, 11), (9, 5), (32, 32), (32, 7), (12, 12), (3, 1)]:
        v = rng.integers(0, 2**63, size=5000, dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, size=5000, dtype=np.uint64)
        if k < 32:
            v &= np.uint64((1 << (2 * k)) - 1)
        v[:200] &= np.uint64(0xFF)                      # low-complexity values: hash ties between equal m-mers
        assert np.array_equal(oracle.minimizer_positions(v, k, m), ref.minimizer_positions(v, k, m)), (k, m)


def test_syncmer_sampler_live(oracle, ref):
    """The restated syncmer stream vs sampler::syncmer_sampler executed verbatim over the reference's kmer_view."""
    rng = np.random.default_rng(31)
    acgt = np.frombuffer(b"ACGT", np.uint8)
    for k, m in [(31, 19), (21, 11), (32, 32), (32, 1), (9, 5), (3, 1)]:
        for pn in (0.0, 0.01):
            s = rng.choice(acgt, size=20000).copy()
            s[rng.random(s.size) < pn] = ord("N")
            s[5000:5400] = ord("A")                     # low complexity: equal m-mers, ties resolved to the leftmost
            for canon in (0, 1):
                for soff, eoff in ((0, k - m), (2 % (k - m + 1), 2 % (k - m + 1))):
                    a = oracle.syncmer_stream(s, k, m, soff, eoff, canon)
                    b = ref.syncmer_stream(s, k, m, soff, eoff, canon)
                    for key in ("lo", "pos", "id"):
                        assert np.array_equal(a[key], b[key]), (key, k, m, pn, canon, soff)


def test_hash_sampler_live(oracle, ref):
    """hash64(value, seed) < threshold vs sampler::hash_sampler executed verbatim over a vector of values; the threshold the
    Python mirror derives from the rate equals the reference constructor's (include/hash_sampler.hpp:77)."""
    import biolib_b200.context as ctxmod
    rng = np.random.default_rng(77)
    v = rng.integers(0, 2**62, size=4000, dtype=np.uint64)
    for seed, rate in ((42, 0.1), (0, 0.5), (7, 0.01), (2**32 + 42, 0.25), (1, 0.0), (3, 0.999)):
        kept, thr = ref.hash_sampler(v, seed, rate)
        assert thr == ctxmod.Context.hash_sampler_threshold(rate)
        h = np.array([oracle.hash64(int(x), seed) for x in v], dtype=np.uint64)
        assert np.array_equal(kept, v[h < np.uint64(thr)]), (seed, rate)


@pytest.mark.parametrize("k,m", [(31, 19), (21, 11), (32, 20), (9, 5), (16, 8), (32, 1), (13, 12), (7, 7), (28, 12)])
def test_d4_window_argmin_pinned_to_reference_extractor(oracle, ref, k, m):
    """D4 (smallest hash, leftmost on ties) of the Tier B restatement against the reference's own
    hash::minimizer_position_extractor executed verbatim (include/kmer_view.hpp:266-283), incl. low-complexity input
    where ties decide."""
    from d4_util import check_d4
    rng = np.random.default_rng(k * 37 + m)
    low = np.concatenate([np.frombuffer(b"A" * 200, np.uint8), np.frombuffer(b"AC" * 150, np.uint8), rand_seq(rng, 300),
                          np.frombuffer(b"ACGTTGCA" * 60, np.uint8), np.frombuffer(b"GATTACA" * 70, np.uint8)])
    for s in (rand_seq(rng, 4000), rand_seq(rng, 3000, 0.004, 0.05), low):
        for brute in (False, True):
            rec = oracle.super_kmers(s, k, m, 0, 0, brute=brute)
            km = oracle.kmers(s, k, 0)
            check_d4(rec, km, ref.minimizer_positions(km["lo"], k, m))
