"""D4 pinned to reference-executed code.

For seed 0 and non-canonical m-mers the reference's own hash::minimizer_position_extractor
(/root/reference/include/kmer_view.hpp:266-283: hash64 seed 0 of every m-mer bit field of ONE k-mer value, scanned
right to left with `>=`, i.e. smallest hash, leftmost on ties) must agree with the window argmin of the super-k-mer
records: for every k-mer start s inside a record, extractor(forward k-mer at s) == mm_start - s.
"""
import numpy as np


def argmin_offset_per_kmer(records: dict) -> tuple:
    """Expand records to one (k-mer start, mm_start - start) pair per k-mer, in position order."""
    size = records["size"].astype(np.int64)
    sk = records["sk_start"].astype(np.int64)
    mm = records["mm_start"].astype(np.int64)
    total = int(size.sum())
    first = np.repeat(sk, size)
    within = np.arange(total, dtype=np.int64) - np.repeat(np.cumsum(size) - size, size)
    starts = first + within
    return starts, np.repeat(mm, size) - starts


def check_d4(records: dict, kmers: dict, extractor_offsets: np.ndarray):
    """records: super-k-mers (seed 0, non-canonical); kmers: every valid forward k-mer as (lo, pos); extractor_offsets:
    minimizer_position_extractor of kmers['lo']."""
    starts, off = argmin_offset_per_kmer(records)
    assert np.array_equal(starts, kmers["pos"].astype(np.int64)), "records must cover exactly the valid k-mers, in order"
    assert np.array_equal(off, extractor_offsets.astype(np.int64)), "window argmin differs from the reference extractor"
