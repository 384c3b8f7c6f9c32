"""Synthetic inputs for the benchmark configurations (host side, numpy; not part of the hot path).

``generate`` reproduces sequence::generator::nucleic<packed::uniform>::get_sequence (reference
include/sequence_generator.hpp:14-23,34-39): std::mt19937 seeded with the 32-bit-narrowed seed, one draw per
base through std::uniform_int_distribution<uint8_t>(0,3), mapped through constants::bases = "ACGT"
(include/constants.hpp:23). With libstdc++ and a 32-bit engine that distribution is the top two bits of each
draw (Lemire multiply-shift with range 4, never rejecting), so the sequence is "ACGT"[draw >> 30];
tests/test_host_side.py pins this against the reference executed here (seed 42 -> CTTAGTGGACAAACTCGAGG...).
numpy's legacy MT19937 seeding is init_genrand(seed), i.e. exactly std::mt19937(seed).

``mutate`` is wrapper::sequence_mutator (reference include/sequence_mutator.hpp:109-169) through the library's host
entry point blg_synth_mutate (csrc/synth_host.cpp: libstdc++'s std::mt19937 + std::uniform_real_distribution, the
reference's own draws), pinned against the reference executed here by tests/test_host_side.py.

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


def mutate(seq, mutp: float, indelf: float, extp: float, seed: int, with_report: bool = False):
    """The reference's ``sequence_mutator(seq, mutp, indelf, extp)`` iterated from ``cbegin(seed)`` to ``cend()`` with the
    absent values (deleted bases) dropped: uint8 ASCII out. ``seq`` must hold AaCcGgTtUu only."""
    import ctypes as C
    from . import _capi
    L = _capi.lib()
    src = np.ascontiguousarray(np.frombuffer(seq.encode() if isinstance(seq, str) else seq, dtype=np.uint8)
                               if not isinstance(seq, np.ndarray) else seq, dtype=np.uint8)
    cap = int(src.size * (1.0 + 2.0 * mutp) + 1024)
    counts = (C.c_uint64 * 3)()
    tot = C.c_int64()
    n_out = C.c_size_t()
    for _ in range(2):
        out = np.empty(cap, dtype=np.uint8)
        rc = L.blg_synth_mutate(src.ctypes.data, src.size, mutp, indelf, extp, int(seed) & (2**64 - 1), out.ctypes.data, cap,
                                C.byref(n_out), C.cast(counts, C.c_void_p), C.cast(C.pointer(tot), C.c_void_p))
        if rc != _capi.BLG_E_CAPACITY:
            break
        cap = n_out.value
    if rc != _capi.BLG_OK:
        raise ValueError(f"blg_synth_mutate: {L.blg_strerror(rc).decode()} (source must be ACGT/U only)")
    res = out[: n_out.value].copy()
    if with_report:
        return res, dict(substitutions=counts[0], insertions=counts[1], deletions=counts[2], total_indel_len=tot.value)
    return res


def mutated_long_reads(n_reads: int, seed_len: int = 46, seed_src: int = 46, seed_mut0: int = 47, lo: int = 10_000, hi: int = 100_000,
                       mutp: float = 0.10, indelf: float = 0.3, extp: float = 0.3, threads: int | None = None):
    """BASELINE config 5 as SURVEY 8(d) states it: read lengths ~U[lo, hi] (mt19937_64(seed_len)), read i =
    sequence_mutator(source slice i, 0.10, 0.3, 0.3) from cbegin(seed_mut0 + i), the source slices being consecutive pieces of
    the uniform(seed_src) stream. Returns (bases uint8, offsets uint64[n_reads + 1])."""
    import os
    from concurrent.futures import ThreadPoolExecutor
    rng = np.random.Generator(np.random.MT19937(seed_len))
    lens = rng.integers(lo, hi + 1, size=n_reads)
    src = generate(seed_src, int(lens.sum()))
    starts = np.concatenate([[0], np.cumsum(lens)])
    with ThreadPoolExecutor(threads or min(32, os.cpu_count() or 1)) as ex:
        reads = list(ex.map(lambda i: mutate(src[starts[i]:starts[i + 1]], mutp, indelf, extp, seed_mut0 + i), range(n_reads)))
    offs = np.concatenate([[0], np.cumsum([r.size for r in reads])]).astype(np.uint64)
    return (np.concatenate(reads) if reads else np.zeros(0, np.uint8)), offs


def related_genomes_slice(lo: int, hi: int, seed_base: int = 71, sub_rate: float = 0.02, shard_len: int = 1 << 24):
    """Bases [lo, hi) of two related synthetic genomes: the first is the generator's stream in shards uniform(seed_base + i)
    of shard_len bases, the second is the first with `sub_rate` random substitutions. Both the shards and the substitutions
    are seeded per shard, so any slice of them is the same bytes however the genome is cut over ranks (the Jaccard
    benchmark's input)."""
    first = lo // shard_len
    s1 = generate_sharded_parallel(seed_base + first, hi - first * shard_len, shard_len)[lo - first * shard_len:]
    s2 = s1.copy()
    for sh in range(first, (hi + shard_len - 1) // shard_len):
        g = np.random.default_rng(5000 + seed_base + sh)
        mut = np.flatnonzero(g.random(shard_len) < sub_rate) + sh * shard_len
        sub = g.choice(np.frombuffer(b"ACGT", np.uint8), size=mut.size)
        keep = (mut >= lo) & (mut < hi)
        s2[mut[keep] - lo] = sub[keep]
    return s1, s2
