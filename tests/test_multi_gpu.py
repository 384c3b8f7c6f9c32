"""Sort / unique / Jaccard across ranks (biolib_b200/multi_gpu.py; SURVEY 8(e) x 8(f) rank 3).

CPU: the host logic -- splitters, cut sizes, the all-to-all by key range, the merge and the final sum -- over gloo with 2 and 3
ranks; the per-rank work is done by a numpy stand-in over the oracle's k-mers (test infrastructure, defined here), the result
must equal the oracle's single-process pipeline. GPU: the product ops (CudaKmerOps: C ABI on the device) through the same
functions in a one-rank NCCL group (the driver's GPU box has one GPU; tools/multi_gpu_jaccard.py is the N-GPU run).
"""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
K = 21


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class NumpyKmerOps:
    """Stand-in for CudaKmerOps on CPU tensors (tests only): same interface, oracle k-mers + numpy set operations."""

    def __init__(self, oracle, torch):
        self.oracle, self.torch = oracle, torch

    def _t(self, a):
        return self.torch.from_numpy(np.ascontiguousarray(a, dtype=np.uint64).view(np.int64).copy())

    @staticmethod
    def _u(t):
        return t.numpy().view(np.uint64)

    def empty(self, n):
        return self.torch.empty(int(n), dtype=self.torch.int64)

    def local_sorted_unique(self, bases, k, canonical, offsets=None):
        return self._t(self.oracle.sorted_unique_kmers(bases, k, canonical, offsets))

    def local_kmers(self, bases, k, canonical, offsets=None):
        return self._t(self.oracle.kmers(bases, k, canonical, offsets=offsets)["lo"])

    def partition(self, keys, bits):
        u = self._u(keys)
        top = ((u >> np.uint64(max(bits - 8, 0))) & np.uint64(255)).astype(np.int64)
        return self._t(u[np.argsort(top, kind="stable")]), np.bincount(top, minlength=256).astype(np.uint64)

    def sample(self, keys, idx):
        return self._u(keys)[np.asarray(idx, np.int64)].copy()

    def lower_bound(self, keys, queries):
        return np.searchsorted(self._u(keys), np.asarray(queries, np.uint64), side="left").astype(np.int64)

    def sort_unique(self, keys, bits):
        return self._t(np.unique(self._u(keys)))

    def set_sizes(self, a, b, bits):
        ua, ub = self._u(a), self._u(b)
        inter = int(np.intersect1d(ua, ub, assume_unique=True).size)
        return inter, int(ua.size + ub.size - inter), int(ua.size), int(ub.size)

    def counts_tensor(self, values):
        return self.torch.tensor([int(v) for v in values], dtype=self.torch.int64)


def _inputs(n, seed):
    from biolib_b200 import synth
    a = synth.generate(seed, n).copy()
    b = a.copy()
    rng = np.random.default_rng(seed)
    mut = rng.random(n) < 0.03
    b[mut] = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=int(mut.sum()))
    for p in rng.integers(0, n - 40, 20):
        a[p:p + int(rng.integers(1, 30))] = ord("N")
    a[n // 3: n // 3 + 2000] = ord("A")            # many equal k-mers: the sorted lists are not uniform
    return a, b


def _worker(rank, world, port, backend, n, strategy, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import torch
    import torch.distributed as dist
    import pyoracle
    from biolib_b200 import multi_gpu as mg

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    orc = pyoracle.Oracle()
    ctx = None
    if backend == "nccl":
        import biolib_b200
        torch.cuda.set_device(0)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", 0))
        ctx = biolib_b200.Context(0)
        ops = mg.CudaKmerOps(ctx, torch, torch.device("cuda", 0))
    else:
        dist.init_process_group("gloo", rank=rank, world_size=world)
        ops = NumpyKmerOps(orc, torch)
    try:
        a, b = _inputs(n, 31)
        la, ha = mg.kmer_chunk(n, K, world, rank)
        ja = mg.distributed_jaccard(ops, dist, a[la:ha], b[la:ha], K, True, strategy=strategy)
        su = mg.distributed_sorted_unique(ops, dist, a[la:ha], K, True, strategy=strategy)
        part = su["part"].cpu().numpy().view(np.uint64)
        parts = [None] * world
        dist.all_gather_object(parts, part)
        if rank == 0:
            want = orc.jaccard(a, b, K, True)
            whole = orc.sorted_unique_kmers(a, K, True)
            cat = np.concatenate(parts)
            sizes = [len(p) for p in parts]
            q.put({"jaccard_ok": (ja["intersection"], ja["union"], ja["size1"], ja["size2"]) == tuple(want),
                   "unique_ok": bool(np.array_equal(cat, whole)), "total_ok": su["total"] == len(whole),
                   "sizes": sizes, "exchanged": ja["keys_exchanged"], "want": want})
    finally:
        dist.destroy_process_group()
        if ctx is not None:
            ctx.close()


def _run(world, backend, n, strategy="partition"):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, backend, n, strategy, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    return q.get(timeout=5)


@pytest.mark.parametrize("world,strategy", [(2, "partition"), (3, "partition"), (2, "sorted"), (3, "sorted")])
def test_key_range_exchange_over_gloo(world, strategy):
    r = _run(world, "gloo", 60_000, strategy)
    assert r["jaccard_ok"] and r["unique_ok"] and r["total_ok"], r
    assert r["want"][0] > 1000 and r["exchanged"] > 0
    # the splitters balance the ranges: no rank holds more than twice its share
    assert max(r["sizes"]) < 2 * sum(r["sizes"]) / world, r["sizes"]


def test_splitters_and_chunks():
    from biolib_b200 import multi_gpu as mg
    for n, k, world in [(1000, 21, 1), (1000, 21, 3), (10, 21, 2), (100_000_000, 31, 8)]:
        spans = [mg.kmer_chunk(n, k, world, r) for r in range(world)]
        starts = [(a, max(b - k + 1, a)) for a, b in spans]
        assert starts[0][0] == 0 and starts[-1][1] == max(n - k + 1, 0)
        for (a, b), (c, d) in zip(starts, starts[1:]):
            assert b == c
    s = mg.choose_splitters(np.array([5, 1, 9, 3, 7, 2, 8, 4], np.uint64), 4)
    assert list(s) == [3, 5, 8] and len(mg.choose_splitters(np.zeros(0, np.uint64), 3)) == 2
    assert len(mg.choose_splitters(np.arange(4, dtype=np.uint64), 1)) == 0
    assert list(mg.sample_positions(10, 4)) == [0, 2, 5, 7] and len(mg.sample_positions(3)) == 3 and len(mg.sample_positions(0)) == 0
    c = np.zeros(256, np.uint64)
    c[10:20] = 100
    b = mg.deal_groups(c, 4)
    assert b[0] == 0 and b[-1] == 256 and all(x <= y for x, y in zip(b, b[1:]))
    assert [int(c[b[r]:b[r + 1]].sum()) for r in range(4)] in ([300, 200, 300, 200], [200, 300, 200, 300], [300, 200, 200, 300], [300, 300, 200, 200])
    assert list(mg.deal_groups(np.zeros(256, np.uint64), 3)) == [0, 256, 256, 256] and list(mg.deal_groups(c, 1)) == [0, 256]


@pytest.mark.gpu
@pytest.mark.parametrize("strategy", ["partition", "sorted"])
def test_product_ops_one_rank_nccl(strategy):
    """CudaKmerOps through the same driver functions in a one-rank NCCL group against the oracle."""
    r = _run(1, "nccl", 3_000_000, strategy)
    assert r["jaccard_ok"] and r["unique_ok"] and r["total_ok"], r


@pytest.mark.gpu
def test_partition_by_top_byte(gpu):
    """blg_partition_by_top_byte: a stable grouping by (key >> (bits - 8)) & 255 and the group sizes."""
    import ctypes as C
    import torch
    rng = np.random.default_rng(11)
    for n, bits in ((0, 62), (1, 62), (5000, 62), (1_000_003, 62), (300_000, 20), (70_000, 6), (4096 * 3, 64)):
        keys = rng.integers(0, 2**bits if bits < 64 else 2**64, size=n, dtype=np.uint64)
        t = torch.from_numpy(keys.view(np.int64).copy()).cuda()
        out = torch.empty_like(t)
        counts = np.zeros(256, np.uint64)
        gpu._check(gpu.L.blg_partition_by_top_byte(gpu.h, C.c_void_p(t.data_ptr()), n, bits, C.c_void_p(out.data_ptr()), counts.ctypes.data))
        top = ((keys >> np.uint64(max(bits - 8, 0))) & np.uint64(255)).astype(np.int64)
        assert np.array_equal(counts, np.bincount(top, minlength=256).astype(np.uint64)), (n, bits)
        assert np.array_equal(out.cpu().numpy().view(np.uint64), keys[np.argsort(top, kind="stable")]), (n, bits)


@pytest.mark.gpu
def test_device_building_blocks(gpu, oracle):
    """blg_sort_unique_u64 / blg_sorted_sets_jaccard / blg_lower_bound_u64 on arbitrary 64-bit keys."""
    import ctypes as C
    import torch
    from biolib_b200 import _capi
    rng = np.random.default_rng(4)
    for n, bits in ((0, 64), (1, 64), (1000, 10), (300_000, 64), (2_000_000, 42)):
        keys = rng.integers(0, 2**bits if bits < 64 else 2**64, size=n, dtype=np.uint64)
        t = torch.from_numpy(keys.view(np.int64).copy()).cuda()
        cnt = C.c_uint64(0)
        gpu._check(gpu.L.blg_sort_unique_u64(gpu.h, C.c_void_p(t.data_ptr()), n, bits, C.byref(cnt)))
        want = np.unique(keys)
        assert cnt.value == want.size and np.array_equal(t[: cnt.value].cpu().numpy().view(np.uint64), want)
        if n:
            q = np.concatenate([want[:: max(want.size // 7, 1)], np.array([0, 2**64 - 1], np.uint64), want[-1:] + np.uint64(1 if want[-1] < 2**64 - 1 else 0)])
            out = np.zeros(q.size, np.uint64)
            gpu._check(gpu.L.blg_lower_bound_u64(gpu.h, C.c_void_p(t.data_ptr()), int(cnt.value), q.ctypes.data, q.size, out.ctypes.data))
            assert np.array_equal(out, np.searchsorted(want, q, side="left").astype(np.uint64))
        other = np.unique(np.concatenate([want[::3], rng.integers(0, 2**bits if bits < 64 else 2**64, size=n // 2, dtype=np.uint64)]))
        tb = torch.from_numpy(other.view(np.int64).copy()).cuda()
        jo = _capi.JaccardOut()
        gpu._check(gpu.L.blg_sorted_sets_jaccard(gpu.h, C.c_void_p(t.data_ptr()), int(cnt.value), C.c_void_p(tb.data_ptr()), other.size, bits, C.byref(jo)))
        inter = np.intersect1d(want, other, assume_unique=True).size
        assert (jo.intersection, jo.union_size, jo.size1, jo.size2) == (inter, want.size + other.size - inter, want.size, other.size)
