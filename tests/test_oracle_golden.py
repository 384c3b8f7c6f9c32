"""CPU: the plain-C oracle (oracle/oracle.c) against the committed golden vectors.

tests/golden/*.json were produced by tests/golden/make_golden.py from the reference's own code executed in the
build container (hash64 / kmer_view / read_base() / generator / mutator verbatim); `superkmers.json` is the
brute-force D1-D9 restatement over that reference-executed stream. These tests pin the oracle, they run without
a GPU and without /root/reference.
"""
import hashlib
import json
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    with open(os.path.join(GOLD, name)) as f:
        return json.load(f)


def test_nt4_table(oracle):
    g = load("hash_kat.json")
    assert oracle.nt4_table().tolist() == g["nt4_table"]
    t = g["nt4_table"]
    assert [t[ord(c)] for c in "ACGTUacgtuN\n>"] == [0, 1, 2, 3, 3, 0, 1, 2, 3, 3, 4, 4, 4] and t[0] == 4


def test_hash_known_answers(oracle):
    g = load("hash_kat.json")
    for k, s, h in g["hash64"]:
        assert oracle.hash64(k, s) == h, (k, s)
    for k, s, h in g["double_hash64_u64"]:
        assert list(oracle.double_hash64(k, s)) == h
    for lo, hi, s, h in g["double_hash64_u128"]:
        assert list(oracle.double_hash64_u128(lo, hi, s)) == h
    for z, r in g["remix"]:
        assert oracle.remix(z) == r
    # SURVEY 8(c) KATs, written out
    assert oracle.hash64(0, 0) == 0x28DF63B7CC57C3CB and oracle.hash64(0, 42) == 0x434B9AFE2C3E9543
    assert oracle.hash64(1, 0) == 0x004403B7FB05C44A and oracle.hash64(1, 42) == 0xD3FE46E112F04C44
    assert oracle.hash64(0x1234567890ABCDEF, 42) == 0x05E73E3913CB8728
    assert oracle.hash64(0xFFFFFFFFFFFFFFFF, 0x10000002A) == oracle.hash64(0xFFFFFFFFFFFFFFFF, 42) == 0xB6A86CA16A85A979
    assert oracle.remix(1) == 0x5692161D100B05E5


def test_kmer_view_stream(oracle):
    g = load("kmer_view.json")
    assert len(g["cases"]) >= 40
    for c in g["cases"]:
        r = oracle.kmer_stream(c["seq"], c["k"], c["canonical"])
        tag = (c["seq"][:20], c["k"], c["canonical"])
        assert r["lo"].tolist() == c["value"], tag
        assert r["has"].tolist() == c["has"] and r["pos"].tolist() == c["pos"] and r["id"].tolist() == c["id"], tag
        if c["bits"] == 128:
            assert r["hi"].tolist() == c["value_hi"], tag


def test_kmer_view_quirks_spelled_out(oracle):
    """K1-K4 of SURVEY 8(a) on the k=3 examples (values ACG=6 CGT=27 GTA=44 TAC=49)."""
    r = oracle.kmer_stream("ACGTACGT", 3, 0)
    assert r["lo"].tolist() == [6, 27, 44, 49, 6]                      # K1: the 6th k-mer is never yielded
    assert oracle.kmer_stream("ACGTACGT", 3, 1)["lo"].tolist() == [6, 6, 44, 44, 6]
    assert oracle.kmer_stream("ACGT", 3, 0)["lo"].tolist() == [6] and len(oracle.kmer_stream("ACG", 3, 0)["lo"]) == 0
    assert oracle.kmer_stream("ACGTN", 3, 0)["lo"].tolist() == [6, 27]  # K2
    r = oracle.kmer_stream("ACGTNNNACGT", 3, 0)                         # K3 + K4
    assert r["has"].tolist() == [1, 1, 0, 1] and r["pos"].tolist() == [0, 1, 2, 7] and r["id"].tolist() == [0, 1, 2, 3]


def test_config1_checksums(oracle):
    """BASELINE configs[0]: kmer_view k=31 canonical over sequence_generator(42), 1 Mbp."""
    g = load("kmer_view.json")["config1_1mbp_seed42_k31_canonical"]
    s = oracle.generate(42, 1_000_000)
    r = oracle.kmer_stream(s, 31, 1)
    assert len(r["lo"]) == g["n"] == 999_969
    assert int(np.bitwise_xor.reduce(r["lo"])) == g["xor"] == 4314334728908509064
    assert int(r["lo"].sum(dtype=np.uint64)) == g["sum"] == 8341584218946238560
    assert int(r["lo"][0]) == g["first"] and int(r["lo"][-1]) == g["last"]
    # the batched form keeps the k-mer the iterator drops
    b = oracle.kmers(s, 31, 1)
    assert len(b["lo"]) == 999_970 and np.array_equal(b["lo"][:-1], r["lo"])


def test_mmer_stream(oracle):
    g = load("mmer_stream.json")
    for c in g["cases"]:
        s, m = c["seq"], c["m"]
        o = oracle.mmers(s, m, c["seed"], c["canonical"])
        valid = np.array(c["valid"], dtype=bool)
        idx = np.array(c["idx"], dtype=np.uint64)
        ends = np.nonzero(valid)[0]
        starts = ends + 1 - m
        assert np.array_equal(idx[ends], (starts + 1).astype(np.uint64))          # D7: idx = 0-based start + 1
        assert int(o["valid"].sum()) == len(ends) and bool(o["valid"][starts].all())
        assert np.array_equal(o["value"][starts], np.array(c["value"], dtype=np.uint64)[ends]), (s[:20], m)
        assert np.array_equal(o["hash"][starts], np.array(c["hash"], dtype=np.uint64)[ends]), (s[:20], m)


@pytest.mark.parametrize("brute", [False, True])
def test_super_kmers(oracle, brute):
    g = load("superkmers.json")
    n_rec = 0
    for c in g["cases"]:
        o = oracle.super_kmers(c["seq"], c["k"], c["m"], c["seed"], c["canonical"], brute=brute)
        tag = (c["seq"][:20], c["k"], c["m"], c["canonical"])
        assert o["n"] == len(c["value"]), tag
        for key in ("value", "hash", "mm_start", "sk_start", "size"):
            assert o[key].tolist() == c[key], (key,) + tag
        n_rec += o["n"]
    assert n_rec > 1000


def test_worked_example(oracle):
    """SURVEY 8(a) worked golden, both strands' tables."""
    s = "ACGGTCATTGACNCGTAGGATCCAGTT"
    o = oracle.super_kmers(s, 7, 4, 42, 0)
    assert o["sk_start"].tolist() == [0, 2, 3, 13, 15, 17] and o["size"].tolist() == [2, 1, 3, 2, 2, 4]
    assert o["mm_start"].tolist() == [1, 2, 6, 14, 16, 20] and o["value"].tolist() == [107, 173, 62, 178, 40, 212]
    assert o["hash"].tolist() == [5635657261505292540, 11490186368348817871, 12361921081606499017, 1386823931139414920,
                                  4562025650448022362, 1228567025868838683]
    o = oracle.super_kmers(s, 7, 4, 42, 1)
    assert o["sk_start"].tolist() == [0, 2, 5, 13, 14, 17] and o["size"].tolist() == [2, 3, 1, 1, 3, 4]
    assert o["mm_start"].tolist() == [1, 4, 7, 13, 16, 20] and o["value"].tolist() == [22, 56, 208, 108, 40, 212]


def test_generator_and_mutator(oracle):
    from biolib_b200 import synth
    g = load("generator.json")
    for seed, n, s in g["generate"]:
        assert oracle.generate(seed, n).tobytes().decode() == s
        assert synth.generate(seed, n).tobytes().decode() == s
    assert g["generate"][2][2].startswith("CTTAGTGGACAAACTCGAGGAATGTTAAATAGCGGACACG")
    big = synth.generate(42, 1_000_000)
    assert hashlib.sha256(big.tobytes()).hexdigest() == g["generate_1m_seed42_sha"]
    assert np.array_equal(big, oracle.generate(42, 1_000_000))
    for c in g["mutate"]:
        out, counts, tot = oracle.mutate(c["src"], c["mutp"], c["indelf"], c["extp"], c["seed"])
        assert out.tobytes().decode() == c["out"], c["seed"]
        assert counts.tolist() == c["counts"] and tot == c["total_indel_len"]


def test_minimizer_position_extractor_golden(oracle):
    """hash::minimizer_position_extractor (reference include/kmer_view.hpp:266-283) against reference-executed vectors."""
    import json
    import os
    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "minpos.json")))
    assert g["nullopt_k7_m4"] == 8                      # k + 1 for a valueless record (kmer_view.hpp:270)
    for c in g["cases"]:
        got = oracle.minimizer_positions(np.array(c["values"], dtype=np.uint64), c["k"], c["m"])
        assert [int(x) for x in got] == c["positions"], (c["k"], c["m"])


def test_syncmer_sampler_golden(oracle):
    """sampler::syncmer_sampler over kmer_view (reference include/syncmer_sampler.hpp:130-137) against reference-executed
    record streams (values, positions, ids; K1-K4 included), and the batched form blg_syncmers returns against them."""
    import json
    import os
    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "syncmers.json")))
    assert len(g["cases"]) >= 20
    for c in g["cases"]:
        o = oracle.syncmer_stream(c["seq"], c["k"], c["m"], c["soff"], c["eoff"], c["canonical"])
        assert o["lo"].tolist() == c["value"] and o["pos"].tolist() == c["position"] and o["id"].tolist() == c["id"], (c["k"], c["m"])
        # the batched form holds the same records plus, at most, the final k-mer the reference's iteration drops (K1)
        b = oracle.syncmers(c["seq"], c["k"], c["m"], c["soff"], c["eoff"], c["canonical"])
        n = len(c["value"])
        assert b["lo"][:n].tolist() == c["value"] and b["pos"][:n].tolist() == c["position"]
        assert len(b["lo"]) - n in (0, 1)
        if len(b["lo"]) > n:
            assert int(b["pos"][-1]) == len(c["seq"]) - c["k"]


def test_hash_sampler_golden(oracle):
    """sampler::hash_sampler (reference include/hash_sampler.hpp:73-78,138-141) against reference-executed vectors."""
    import json
    import os
    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "hash_sampler.json")))
    v = np.array(g["values"], dtype=np.uint64)
    for c in g["cases"]:
        assert c["threshold"] == min(int(c["rate"] * float(2**64 - 1)), 2**64 - 1)
        h = np.array([oracle.hash64(int(x), c["seed"]) for x in v], dtype=np.uint64)
        assert v[h < np.uint64(c["threshold"])].tolist() == c["kept"], (c["seed"], c["rate"])
