"""The N > 1 host logic on CPU: two gloo ranks shard one input with biolib_b200.shard, each extracts the records it
owns from ITS slice only, the counts go through the path's single collective (a sum) and the rank-ordered
concatenation must equal the unsharded result. The per-rank engine here is the oracle (tests may use it as the
checker's stand-in; the GPU tests run the same plans through blg_super_kmers_range).
"""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

K, M, SEED = 31, 19, 42


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _make_input(n):
    from biolib_b200 import synth
    s = synth.generate(123, n).copy()
    rng = np.random.default_rng(5)
    for p in rng.integers(0, n - 40, 25):
        s[p:p + int(rng.integers(1, 30))] = ord("N")
    return s


def _records(orc, seq, k, m, offsets=None):
    r = orc.super_kmers(seq, k, m, SEED, True, offsets=offsets)
    return {key: np.asarray(r[key]) for key in ("value", "hash", "mm_start", "sk_start", "size")}


def _worker(rank, world, port, mode, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import torch
    import torch.distributed as dist
    import pyoracle
    from biolib_b200 import shard

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    orc = pyoracle.Oracle()
    try:
        if mode == "chunks":
            n = 60_000
            seq = _make_input(n)
            plan = shard.plan_chunks(n, K, M, world)[rank]
            sl = seq[plan.buf_lo:plan.buf_hi]                       # all this rank is ever given
            rec = _records(orc, sl, K, M)
            lo, hi = plan.local_window
            keep = (rec["sk_start"] >= lo) & (rec["sk_start"] < hi)  # first k-mer owned by this rank
            rec = {k_: v[keep] for k_, v in rec.items()}
            rec["mm_start"] = rec["mm_start"] + np.uint64(plan.buf_lo)
            rec["sk_start"] = rec["sk_start"] + np.uint64(plan.buf_lo)
            owned_kmers = plan.win_hi - plan.win_lo
        else:
            rng = np.random.default_rng(9)
            lens = rng.integers(20, 400, 300)
            offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
            seq = _make_input(int(offs[-1]))
            a, b = shard.plan_sequences(offs, world)[rank]
            base = int(offs[a])
            sl = seq[base:int(offs[b])]
            rec = _records(orc, sl, K, M, offsets=offs[a:b + 1] - np.uint64(base))
            rec["mm_start"] = rec["mm_start"] + np.uint64(base)
            rec["sk_start"] = rec["sk_start"] + np.uint64(base)
            owned_kmers = int(np.maximum(lens[a:b].astype(np.int64) - K + 1, 0).sum())
        # the path's only collective: sum of the per-rank counts
        t = torch.tensor([len(rec["value"]), owned_kmers], dtype=torch.int64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        counts = [None] * world
        dist.all_gather_object(counts, len(rec["value"]))
        parts = [None] * world
        dist.all_gather_object(parts, rec)
        if rank == 0:
            whole = _records(orc, seq, K, M, offsets=None if mode == "chunks" else offs)
            cat = {k_: np.concatenate([p[k_] for p in parts]) for k_ in whole}
            ok = all(np.array_equal(cat[k_], whole[k_]) for k_ in whole)
            offsets_ok = shard.global_record_offsets(counts) == [0] + list(np.cumsum(counts)[:-1])
            kmers_expected = (len(seq) - K + 1) if mode == "chunks" else int(np.maximum(lens.astype(np.int64) - K + 1, 0).sum())
            q.put((ok, offsets_ok, int(t[0]) == len(whole["value"]), int(t[1]) == kmers_expected, len(whole["value"])))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["chunks", "sequences"])
def test_two_ranks_concatenate_to_the_unsharded_result(mode):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, mode, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    ok, offsets_ok, count_ok, kmers_ok, n = q.get(timeout=5)
    assert n > 100
    assert ok and offsets_ok and count_ok and kmers_ok


def test_chunk_plan_properties():
    from biolib_b200 import shard
    for n, world in [(1000, 1), (1000, 3), (100_000_000, 8), (50, 4), (10, 2)]:
        plans = shard.plan_chunks(n, K, M, world)
        nwin = max(n - K + 1, 0)
        assert plans[0].win_lo == 0 and plans[-1].win_hi == nwin
        for a, b in zip(plans, plans[1:]):
            assert a.win_hi == b.win_lo                      # disjoint, gap-free
        for p in plans:
            assert p.buf_lo <= p.win_lo and p.buf_hi <= n
            if p.win_hi > p.win_lo:
                assert p.buf_hi >= min(p.win_hi + shard.right_halo(K, M), n)
                assert p.buf_lo % 16 == 0
    assert shard.right_halo(31, 19) == 31 + 13 - 2


def test_sequence_plan_balances_by_bases():
    from biolib_b200 import shard
    lens = np.array([150] * 1000 + [100_000] + [150] * 1000)
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    plan = shard.plan_sequences(offs, 4)
    assert plan[0][0] == 0 and plan[-1][1] == len(lens)
    for (a, b), (c, d) in zip(plan, plan[1:]):
        assert b == c
    loads = [int(offs[b] - offs[a]) for a, b in plan]
    assert max(loads) <= 100_000 + 150 * 1000
