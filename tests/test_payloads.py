"""Super-k-mer payloads grouped by minimizer bucket.

CPU: the oracle's packing against the reference's own bit::packed::vector<uint64_t>(2) (oracle/_ref), and a decode round
trip. GPU: blg_super_kmer_payloads against the oracle restatement, bit for bit."""
import numpy as np
import pytest


def rand_seq(rng, n, p_n=0.0):
    s = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=n)
    if p_n > 0:
        s = s.copy()
        s[rng.random(n) < p_n] = ord("N")
    return s


def decode(words, start, length):
    e = np.arange(start, start + length, dtype=np.int64)
    return ((words[e // 32] >> (62 - 2 * (e % 32)).astype(np.uint64)) & np.uint64(3)).astype(np.uint8)


def test_oracle_packing_is_the_references_packed_vector(oracle, ref):
    rng = np.random.default_rng(1)
    s = rand_seq(rng, 6000, 0.003)
    for k, m, nb in ((31, 19, 7), (21, 11, 64), (9, 5, 1)):
        o = oracle.payloads(s, k, m, 42, True, nb)
        assert np.array_equal(o["words"], ref.packed_vector_words(o["codes"], 2))
        # every payload decodes to the bases of its super-k-mer
        code_of = oracle.nt4_table()
        for i in rng.integers(0, o["n"], 50):
            L = int(o["size"][i]) + k - 1
            a = int(o["sk_start"][i])
            assert np.array_equal(decode(o["words"], int(o["payload_start"][i]), L), code_of[s[a:a + L]])
        assert o["bucket_offsets"][-1] == o["n"] and np.all(np.diff(o["bucket_offsets"].astype(np.int64)) >= 0)


@pytest.mark.gpu
@pytest.mark.parametrize("k,m,nb", [(31, 19, 256), (21, 11, 1), (63, 31, 4096), (9, 5, 13), (31, 19, 2)])
def test_payloads_vs_oracle(gpu, oracle, k, m, nb):
    rng = np.random.default_rng(k + nb)
    for n in (0, k, 3000, 150_003):
        s = rand_seq(rng, n, 0.002)
        for canon in (False, True):
            g = gpu.super_kmer_payloads(s, k, m, 42, canon, nb)
            o = oracle.payloads(s, k, m, 42, canon, nb)
            assert g["n"] == o["n"] and g["total_bases"] == o["total_bases"]
            for key in ("bucket_offsets", "value", "mm_start", "size", "mm_pos", "payload_start", "words"):
                assert np.array_equal(g[key], o[key]), (key, k, m, nb, n)


@pytest.mark.gpu
def test_payloads_multi_sequence_and_errors(gpu, oracle):
    rng = np.random.default_rng(9)
    lens = rng.integers(0, 500, size=300)
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    s = rand_seq(rng, int(offs[-1]), 0.002)
    g = gpu.super_kmer_payloads(s, 21, 11, 7, True, 32, offsets=offs)
    o = oracle.payloads(s, 21, 11, 7, True, 32, offsets=offs)
    for key in ("bucket_offsets", "value", "mm_start", "size", "payload_start", "words"):
        assert np.array_equal(g[key], o[key]), key
    with pytest.raises(Exception):
        gpu.super_kmer_payloads(s, 21, 11, 7, True, 0)
    with pytest.raises(Exception):
        gpu.super_kmer_payloads(s, 21, 11, 7, True, 5000)
