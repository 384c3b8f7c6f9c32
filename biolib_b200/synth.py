"""Synthetic inputs for the benchmark configurations (host side, numpy; not part of the hot path).

``generate`` reproduces sequence::generator::nucleic<packed::uniform>::get_sequence (reference
include/sequence_generator.hpp:14-23,34-39): std::mt19937 seeded with the 32-bit-narrowed seed, one draw per
base through std::uniform_int_distribution<uint8_t>(0,3), mapped through constants::bases = "ACGT"
(include/constants.hpp:23). With libstdc++ and a 32-bit engine that distribution is the top two bits of each
draw (Lemire multiply-shift with range 4, never rejecting), so the sequence is "ACGT"[draw >> 30];
tests/test_synth.py pins this against the reference executed here (seed 42 -> CTTAGTGGACAAACTCGAGG...).
numpy's legacy MT19937 seeding is init_genrand(seed), i.e. exactly std::mt19937(seed).

The reference's sequence_mutator never produces N (include/sequence_mutator.hpp:133-168), so the read
configurations use ``inject_n_runs``, a builder-owned seeded N injector.
"""
from __future__ import annotations

import numpy as np

_BASES = np.frombuffer(b"ACGT", dtype=np.uint8)


def _raw_stream(seed: int):
    bg = np.random.MT19937()
    bg._legacy_seeding(int(seed) & 0xFFFFFFFF)
    return bg


def generate(seed: int, length: int, out: np.ndarray | None = None, chunk: int = 1 << 24) -> np.ndarray:
    """``length`` bases of the reference generator's stream for ``seed`` (uint8 ASCII)."""
    if out is None:
        out = np.empty(length, dtype=np.uint8)
    bg = _raw_stream(seed)
    for lo in range(0, length, chunk):
        n = min(chunk, length - lo)
        raw = bg.random_raw(n)  # the engine's 32-bit outputs, widened to uint64
        np.take(_BASES, (raw >> np.uint64(30)).astype(np.intp), out=out[lo:lo + n])
    return out


def generate_sharded(seed0: int, length: int, shard: int, first_shard: int = 0, out: np.ndarray | None = None) -> np.ndarray:
    """A long genome as consecutive shards of ``shard`` bases, shard i drawn from uniform(seed0 + i).

    Deterministic given (seed0, shard); lets ranks build just their slice (plus halo) of one global sequence.
    ``length`` bases are produced starting at shard ``first_shard``.
    """
    if out is None:
        out = np.empty(length, dtype=np.uint8)
    done, i = 0, first_shard
    while done < length:
        n = min(shard, length - done)
        generate(seed0 + i, n, out=out[done:done + n])
        done += n
        i += 1
    return out


def inject_n_runs(reads: np.ndarray, read_len: int, seed: int = 44, p_read: float = 0.05, p_geom: float = 0.5,
                  max_run: int = 20, lower_frac: float = 0.01) -> np.ndarray:
    """In place: per read Bernoulli(p_read) one N run, start ~U[0, read_len), length 1 + Geom(p_geom) capped at
    max_run (clipped to the read); ``lower_frac`` of the reads are lower-cased (exercises the nt4 table)."""
    n_reads = reads.size // read_len
    rng = np.random.Generator(np.random.MT19937(seed))
    has = np.nonzero(rng.random(n_reads) < p_read)[0]
    start = rng.integers(0, read_len, size=has.size)
    run = np.minimum(rng.geometric(p_geom, size=has.size), max_run)
    view = reads[: n_reads * read_len].reshape(n_reads, read_len)
    for r, s, l in zip(has, start, run):
        view[r, s:min(s + l, read_len)] = ord("N")
    low = np.nonzero(rng.random(n_reads) < lower_frac)[0]
    view[low] |= 0x20
    return reads


def generate_sharded_parallel(seed0: int, length: int, shard: int = 1 << 26, threads: int | None = None) -> np.ndarray:
    """generate_sharded with the shards drawn concurrently (numpy's generator releases the GIL): a 3 Gbp genome in
    seconds instead of a minute. Same bytes as generate_sharded(seed0, length, shard)."""
    import os
    from concurrent.futures import ThreadPoolExecutor
    out = np.empty(length, dtype=np.uint8)
    jobs = [(i, lo, min(shard, length - lo)) for i, lo in enumerate(range(0, length, shard))]
    with ThreadPoolExecutor(threads or min(32, os.cpu_count() or 1)) as ex:
        list(ex.map(lambda j: generate(seed0 + j[0], j[2], out=out[j[1]:j[1] + j[2]]), jobs))
    return out
