"""Sharding plans for the multi-GPU driver (host logic only; no compute here).

The path shards into independent units: every record depends on at most 2k-m+1 consecutive
bases of one sequence, so ranks never exchange bases. Two plans:

* ``plan_chunks``     one long sequence cut into ranges of k-mer starts, each rank reading its range plus halos
                      (configs 2 and 4). A record belongs to the rank that owns its FIRST k-mer
                      (``blg_super_kmers_range``), so the per-rank outputs are disjoint and concatenate in rank
                      order into exactly the single-GPU output.
* ``plan_sequences``  a batch of reads cut at read boundaries into contiguous ranges balanced by bases
                      (configs 3 and 5); no halo.

The only collective of the path is the final sum of per-rank record / k-mer counts (16 bytes per rank).
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

LEFT_HALO = 16    # bases before the first owned k-mer start: 1 is enough to know whether the previous k-mer's
                  # minimizer differs; 16 keeps the slice 16-byte aligned for the kernel's vector loads


@dataclass(frozen=True)
class ChunkPlan:
    rank: int
    buf_lo: int      # first base of the slice this rank needs (global coordinate)
    buf_hi: int      # one past the last base of the slice
    win_lo: int      # owned k-mer starts [win_lo, win_hi), global coordinates
    win_hi: int

    @property
    def local_window(self):
        """The window range in slice coordinates, as blg_super_kmers_range wants it."""
        return self.win_lo - self.buf_lo, self.win_hi - self.buf_lo


def right_halo(k: int, m: int) -> int:
    """Bases after the last owned k-mer start that its record can reach: the run has at most w k-mers, the last of
    them ends k-1 bases after its start => (w - 1) + (k - 1) = k + w - 2 (the north star's halo figure)."""
    return k + (k - m + 1) - 2


def plan_chunks(n_bases: int, k: int, m: int, world: int, align: int = 16) -> list[ChunkPlan]:
    """Cut the k-mer starts [0, n_bases-k+1) of one sequence into ``world`` contiguous ranges of equal size
    (boundaries rounded down to ``align``), with the halos each rank must be given."""
    nwin = max(n_bases - k + 1, 0)
    plans = []
    for r in range(world):
        lo = (nwin * r // world) // align * align if r > 0 else 0
        hi = (nwin * (r + 1) // world) // align * align if r < world - 1 else nwin
        buf_lo = max(lo - LEFT_HALO, 0)
        buf_hi = min(hi + right_halo(k, m), n_bases) if hi > lo else buf_lo
        plans.append(ChunkPlan(r, buf_lo, buf_hi, lo, hi))
    return plans


def plan_sequences(offsets: np.ndarray, world: int) -> list[tuple[int, int]]:
    """Contiguous ranges [first_seq, last_seq) per rank, balanced by total bases (offsets has n_seq+1 entries)."""
    offsets = np.asarray(offsets, dtype=np.uint64)
    n_seq = len(offsets) - 1
    total = int(offsets[-1])
    cuts = [0]
    for r in range(1, world):
        target = total * r // world
        cuts.append(max(int(np.searchsorted(offsets, np.uint64(target), side="left")), cuts[-1]))
    cuts.append(n_seq)
    cuts = [min(c, n_seq) for c in cuts]
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


def global_record_offsets(counts) -> list[int]:
    """Exclusive scan of the gathered per-rank record counts: where each rank's records start in the
    concatenated result."""
    out, acc = [], 0
    for c in counts:
        out.append(acc)
        acc += int(c)
    return out
