"""Fourth-generation fused kernel (biolib_b200/csrc/sk4_device.cuh): 32-bit window keys + exact redo of tiles with ties.

Bit-exact against the oracle through the C ABI, on the inputs that stress what is new in this generation: tile edges
(a tile is (32 NB - 3) B windows), every m the lane-local streams admit, inputs where EVERY tile has ties in the hash
part of the keys (low-complexity sequence: the exact path must run and must agree), N runs, many short sequences.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from test_gpu_parity import check_sk, rand_seq  # noqa: E402

TILE12 = (32 * 2 - 3) * 12       # k - m = 12 (two blocks of w - 1 starts per lane)
TILE10 = (32 * 2 - 3) * 10       # k - m = 10
TILE32 = (32 * 2 - 5) * 16       # k - m = 32: a window spans three blocks of 16 (w = 33), two more margin blocks


@pytest.mark.parametrize("canon", [False, True])
@pytest.mark.parametrize("k,m,tile", [(31, 19, TILE12), (21, 11, TILE10), (63, 31, TILE32), (40, 8, TILE32)])
def test_tile_edges(gpu, oracle, k, m, tile, canon):
    rng = np.random.default_rng(k + m + canon)
    for nwin in (1, 2, tile - 1, tile, tile + 1, 2 * tile - 1, 2 * tile, 2 * tile + 1, 33 * tile, 33 * tile + 7, 65 * tile - 3):
        check_sk(gpu, oracle, rand_seq(rng, nwin + k - 1), k, m, 42, canon)


@pytest.mark.parametrize("d", [12, 10, 32])
@pytest.mark.parametrize("m", list(range(1, 33)))
def test_every_supported_m(gpu, oracle, m, d):
    """k - m = 12, 10 and 32 for every m: one-word m-mers (m <= 16), two-word with 2- and 3-word block streams."""
    k = m + d
    rng = np.random.default_rng(m * 2 + d)
    for canon in (False, True):
        check_sk(gpu, oracle, rand_seq(rng, 20011), k, m, 7, canon)
        check_sk(gpu, oracle, rand_seq(rng, 9000, p_n=0.002, lower=0.1), k, m, 0, canon)


@pytest.mark.parametrize("k,m", [(31, 19), (21, 11), (63, 31), (37, 5)])
def test_ties_take_the_exact_path(gpu, oracle, k, m):
    """Low-complexity sequence: equal m-mers inside one window => equal keys => exact redo. Results must not change and
    the library must report that the exact path ran."""
    rng = np.random.default_rng(5)
    parts = [np.frombuffer(b"A" * 3000, np.uint8), np.frombuffer(b"AC" * 2500, np.uint8), np.frombuffer(b"ACG" * 1500, np.uint8),
             rand_seq(rng, 4000), np.frombuffer(b"ACGTTGCA" * 700, np.uint8), np.frombuffer(b"T" * 100, np.uint8),
             rand_seq(rng, 3000), np.frombuffer(b"GATTACA" * 1000, np.uint8)]
    s = np.concatenate(parts)
    for canon in (False, True):
        check_sk(gpu, oracle, s, k, m, 42, canon)
        assert gpu.stats()["exact_tiles"] > 0
    # random sequence: at m = 19 practically no tile needs it; at m = 11 reverse-complement neighbours (a 12-base
    # reverse palindrome makes two adjacent canonical 11-mers equal: about 5 per 10^4 positions) flag one tile in eight
    r = rand_seq(rng, 400_000)
    check_sk(gpu, oracle, r, k, m, 42, True)
    assert gpu.stats()["exact_tiles"] <= (2 if m >= 19 else 430)


@pytest.mark.parametrize("k,m", [(31, 19), (21, 11), (63, 31)])
def test_breaks_and_long_n_runs(gpu, oracle, k, m):
    rng = np.random.default_rng(11)
    s = rand_seq(rng, 120_000, p_n=0.001, lower=0.05)
    s[10_000:10_900] = ord("N")          # longer than a tile's span of one lane
    s[50_000:50_001] = ord("N")
    s[77_000:77_000 + k] = ord("N")
    s[-3:] = ord("N")
    for canon in (False, True):
        check_sk(gpu, oracle, s, k, m, 42, canon)
    allbad = np.full(5000, ord("N"), np.uint8)
    check_sk(gpu, oracle, allbad, k, m, 42, True)


@pytest.mark.parametrize("k,m", [(31, 19), (21, 11), (63, 31)])
def test_many_short_sequences(gpu, oracle, k, m):
    rng = np.random.default_rng(13)
    lens = rng.integers(0, 400, size=900)
    lens[::50] = 150
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    s = rand_seq(rng, int(offs[-1]), p_n=0.001)
    for canon in (False, True):
        check_sk(gpu, oracle, s, k, m, 42, canon, offsets=offs)


def test_window_ranges(gpu, oracle):
    """blg_super_kmers_range over cuts that are not tile multiples concatenates to the single call."""
    k, m = 31, 19
    rng = np.random.default_rng(17)
    s = rand_seq(rng, 300_000, p_n=0.0005)
    whole = gpu.super_kmers(s, k, m, 42, True)
    cuts = [0, 1, 733, 100_003, 100_004, 250_000, len(s)]
    parts = [gpu.super_kmers(s, k, m, 42, True, win_range=(cuts[i], cuts[i + 1])) for i in range(len(cuts) - 1)]
    for key in ("value", "mm_start", "size", "mm_pos", "hash"):
        assert np.array_equal(np.concatenate([p[key] for p in parts]), whole[key]), key
    o = oracle.super_kmers(s, k, m, 42, True)
    assert np.array_equal(whole["mm_start"], o["mm_start"]) and np.array_equal(whole["size"], o["size"])
