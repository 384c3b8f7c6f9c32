"""Sort / unique / Jaccard of k-mers: the pipeline of the reference's tests/test_jaccard.cpp.

CPU: the oracle restatement against the reference pieces executed verbatim (kmer_view values incl. its dropped last k-mer,
sampler::ordered_unique_sampler, algorithm::jaccard; numpy's sort stands for the external-memory vector's sorted order).
GPU: blg_sorted_unique_kmers / blg_kmer_jaccard against the oracle."""
import numpy as np
import pytest


def rand_seq(rng, n, p_n=0.0):
    s = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=n)
    if p_n > 0:
        s = s.copy()
        s[rng.random(n) < p_n] = ord("N")
    return s


def related_pair(rng, n, p_mut=0.05):
    a = rand_seq(rng, n, 0.001)
    b = a.copy()
    mut = rng.random(n) < p_mut
    b[mut] = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=int(mut.sum()))
    return a, b


def test_oracle_vs_reference_pipeline(oracle, ref):
    rng = np.random.default_rng(1)
    for k, canon in ((15, True), (31, True), (9, False), (32, False), (4, True)):
        a, b = related_pair(rng, 4000)
        offs = np.array([0, 700, 700, 2500, 4000], np.uint64)
        for o in (None, offs):
            va, vb = np.sort(ref.kmer_values(a, k, canon, o)), np.sort(ref.kmer_values(b, k, canon, o))
            assert np.array_equal(ref.ordered_unique(va), oracle.sorted_unique_kmers(a, k, canon, o, drop_last=True))
            assert ref.jaccard_sorted(va, vb) == oracle.jaccard(a, b, k, canon, o, o, drop_last=True)


@pytest.mark.gpu
@pytest.mark.parametrize("k,canon", [(15, True), (31, True), (9, False), (32, False), (4, True), (1, False), (21, True)])
def test_sorted_unique_and_jaccard(gpu, oracle, k, canon):
    rng = np.random.default_rng(k)
    for n in (0, k, 5000, 300_007):
        a, b = related_pair(rng, n)
        for drop in (False, True):
            assert np.array_equal(gpu.sorted_unique_kmers(a, k, canon, drop_last=drop), oracle.sorted_unique_kmers(a, k, canon, drop_last=drop))
            assert gpu.kmer_jaccard(a, b, k, canon, drop_last=drop) == oracle.jaccard(a, b, k, canon, drop_last=drop)
    lens = rng.integers(0, 900, size=200)
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    a, b = related_pair(rng, int(offs[-1]))
    for drop in (False, True):
        assert np.array_equal(gpu.sorted_unique_kmers(a, k, canon, offsets=offs, drop_last=drop), oracle.sorted_unique_kmers(a, k, canon, offs, drop_last=drop))
        assert gpu.kmer_jaccard(a, b, k, canon, offs, offs, drop_last=drop) == oracle.jaccard(a, b, k, canon, offs, offs, drop_last=drop)


@pytest.mark.gpu
def test_jaccard_reference_pipeline_on_gpu_box(gpu, oracle):
    """With oracle/_ref on the box: the GPU numbers against the reference's own sampler + jaccard executed there."""
    import pyoracle
    if not pyoracle.Ref.available():
        pytest.skip("oracle/_ref did not travel")
    ref = pyoracle.Ref()
    rng = np.random.default_rng(7)
    a, b = related_pair(rng, 20_000, 0.02)
    for k, canon in ((15, True), (31, False)):
        va, vb = np.sort(ref.kmer_values(a, k, canon)), np.sort(ref.kmer_values(b, k, canon))
        assert gpu.kmer_jaccard(a, b, k, canon, drop_last=True) == ref.jaccard_sorted(va, vb)
        assert np.array_equal(gpu.sorted_unique_kmers(a, k, canon, drop_last=True), ref.ordered_unique(va))


@pytest.mark.gpu
def test_sort_many_tiles_low_entropy(gpu, oracle):
    """Many radix tiles, heavy duplication (every digit group large and skewed)."""
    rng = np.random.default_rng(11)
    motif = rand_seq(rng, 3000)
    s = np.concatenate([motif] * 400 + [rand_seq(rng, 200_000)])
    assert np.array_equal(gpu.sorted_unique_kmers(s, 17, True), oracle.sorted_unique_kmers(s, 17, True))
