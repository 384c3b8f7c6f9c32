"""Sort / unique / Jaccard of k-mers across GPUs: SURVEY 8(f) rank 3 at N > 1, the one step of the path where ranks exchange data.

The reference's pipeline (tests/test_jaccard.cpp:55-130) is kmer_view -> sorted external-memory vector ->
ordered_unique_sampler -> algorithm::jaccard. On one GPU that is blg_sorted_unique_kmers / blg_kmer_jaccard. On N GPUs
(one process per GPU, torch.distributed for the plumbing) equal k-mers found by different ranks have to meet, so the key
space is cut into N ranges and every key travels to the rank that owns its range: ONE all-to-all of 8-byte keys over NVLink
per set (NCCL ``all_to_all_single``). Rank r ends up with the distinct k-mers of range r of the WHOLE input, ranges ascending
with the rank: the concatenation in rank order is the single-GPU result. Two ways to get there:

  strategy "partition" (default)  exchange first, sort once. Every rank extracts its k-mers (blg_kmers), groups them by their
      top byte in one stable radix pass (blg_partition_by_top_byte), the per-byte counts of all ranks are summed (all-reduce of
      256 counters) and the 256 groups are dealt to the ranks as contiguous runs of about equal size; a rank's share is then a
      contiguous piece of every sender's grouped array. After the all-to-all each rank sorts and deduplicates what it received
      (blg_sort_unique_u64): the sort work per rank is 1/N of the whole, which is what makes the step scale.
  strategy "sorted"  sort first, exchange the distinct keys (a sample sort): every rank sorts and deduplicates locally, the ranks
      agree on N - 1 splitters from a regular sample of everybody's sorted list, every rank cuts its list at the splitters
      (blg_lower_bound_u64), and after the all-to-all each rank merges what it received with another sort. Every key is sorted
      twice, but only distinct keys travel: the choice for inputs that repeat themselves many times over (deep read sets).

Jaccard: both sets use the same ranges, each rank intersects its two ranges (blg_sorted_sets_jaccard) and the counters are
summed (one all-reduce).

The host logic (ranges, cut sizes, the exchange) is independent of where the keys live: `ops` does the per-rank work.
`CudaKmerOps` is the product implementation (C ABI on the rank's GPU, torch tensors as device memory); the CPU tests drive
the same functions over gloo with a numpy stand-in that lives in tests/. There is no CPU implementation in the product.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

SAMPLES_PER_RANK = 64


def kmer_chunk(n_bases: int, k: int, world: int, rank: int) -> tuple[int, int]:
    """The bases [lo, hi) rank `rank` needs so that the ranks' k-mer starts partition [0, n_bases - k + 1): its starts plus
    the k - 1 bases its last k-mer reaches into."""
    nwin = max(n_bases - k + 1, 0)
    a, b = nwin * rank // world, nwin * (rank + 1) // world
    return (a, b + k - 1) if b > a else (a, a)


def choose_splitters(samples: np.ndarray, world: int) -> np.ndarray:
    """world - 1 keys cutting the sorted sample into equal parts: rank r owns the keys v with splitters[r-1] <= v < splitters[r].
    Every rank computes this from the same gathered sample, so all agree. An empty sample gives empty ranges to ranks > 0."""
    s = np.sort(np.asarray(samples, dtype=np.uint64))
    if world <= 1:
        return np.zeros(0, np.uint64)
    if s.size == 0:
        return np.full(world - 1, np.iinfo(np.uint64).max, np.uint64)
    return s[[min(s.size * r // world, s.size - 1) for r in range(1, world)]]


def sample_positions(n: int, count: int = SAMPLES_PER_RANK) -> np.ndarray:
    """`count` regularly spaced indices into a list of n entries (fewer when n is small)."""
    if n <= 0:
        return np.zeros(0, np.int64)
    c = min(count, n)
    return (np.arange(c, dtype=np.int64) * n) // c


class CudaKmerOps:
    """Per-rank work on the rank's GPU through the C ABI. Keys are uint64 bit patterns held in torch int64 tensors."""

    def __init__(self, ctx, torch, device):
        self.ctx, self.torch, self.device = ctx, torch, device

    def empty(self, n: int):
        return self.torch.empty(int(n), dtype=self.torch.int64, device=self.device)

    def local_sorted_unique(self, bases, k: int, canonical: bool, offsets=None):
        """Distinct k-mer values of this rank's bases, sorted, in a device tensor (local_kmers + blg_sort_unique_u64 in place)."""
        return self.sort_unique(self.local_kmers(bases, k, canonical, offsets), 2 * k)

    def local_kmers(self, bases, k: int, canonical: bool, offsets=None):
        """The k-mer values of this rank's bases in input order, in a device tensor: blg_kmers into torch-owned memory (values
        only, no positions). `bases`: host uint8 array or a device uint8 tensor."""
        if k > 32:
            raise ValueError("sorted / unique / Jaccard: k <= 32 (64-bit k-mer values)")
        torch = self.torch
        d = bases if hasattr(bases, "data_ptr") else torch.from_numpy(np.ascontiguousarray(bases, dtype=np.uint8)).to(self.device)
        n = int(d.numel())
        d_offs = None if offsets is None else torch.from_numpy(np.ascontiguousarray(offsets, dtype=np.uint64).view(np.int64)).to(self.device)
        keys = self.empty(max(n - k + 1, 0))
        torch.cuda.synchronize(self.device)
        cnt = self.ctx.kmers_device(d.data_ptr(), n, k, canonical, keys.data_ptr(), 0, 0, max(int(keys.numel()), 1) if n >= k else 1,
                                    d_offsets=0 if d_offs is None else d_offs.data_ptr(), n_seq=0 if d_offs is None else int(d_offs.numel()) - 1) \
            if n >= k else 0
        return keys[:cnt]

    def partition(self, keys, bits: int):
        """(keys grouped stably by their top byte, keys per group as a host array of 256): blg_partition_by_top_byte."""
        out = self.empty(int(keys.numel()))
        counts = np.zeros(256, np.uint64)
        self.torch.cuda.synchronize(self.device)
        self.ctx._check(self.ctx.L.blg_partition_by_top_byte(self.ctx.h, C.c_void_p(keys.data_ptr()), int(keys.numel()), int(bits),
                                                             C.c_void_p(out.data_ptr()), counts.ctypes.data))
        return out, counts

    def sample(self, keys, idx: np.ndarray) -> np.ndarray:
        if len(idx) == 0:
            return np.zeros(0, np.uint64)
        sel = keys[self.torch.from_numpy(np.asarray(idx, np.int64)).to(self.device)]
        return sel.cpu().numpy().view(np.uint64)

    def lower_bound(self, keys, queries: np.ndarray) -> np.ndarray:
        q = np.ascontiguousarray(queries, dtype=np.uint64)
        out = np.zeros(q.size, np.uint64)
        if q.size:
            self.ctx._check(self.ctx.L.blg_lower_bound_u64(self.ctx.h, C.c_void_p(keys.data_ptr()), int(keys.numel()), q.ctypes.data, q.size,
                                                           out.ctypes.data))
        return out.astype(np.int64)

    def sort_unique(self, keys, bits: int):
        cnt = C.c_uint64(0)
        self.torch.cuda.synchronize(self.device)             # the exchange ran on torch's stream, the sort runs on the context's
        self.ctx._check(self.ctx.L.blg_sort_unique_u64(self.ctx.h, C.c_void_p(keys.data_ptr()), int(keys.numel()), int(bits), C.byref(cnt)))
        return keys[: cnt.value]

    def set_sizes(self, a, b, bits: int):
        from . import _capi
        out = _capi.JaccardOut()
        self.ctx._check(self.ctx.L.blg_sorted_sets_jaccard(self.ctx.h, C.c_void_p(a.data_ptr()), int(a.numel()), C.c_void_p(b.data_ptr()),
                                                           int(b.numel()), int(bits), C.byref(out)))
        return out.intersection, out.union_size, out.size1, out.size2

    def counts_tensor(self, values):
        return self.torch.tensor(list(values), dtype=self.torch.int64, device=self.device)


def _world(dist, group=None) -> int:
    """dist = None (or an uninitialised torch.distributed) means a single process: every step degenerates to the local one."""
    return dist.get_world_size(group) if dist is not None and dist.is_initialized() else 1


def agree_on_splitters(ops, dist, sorted_lists, group=None) -> np.ndarray:
    """Step 2: a regular sample of every rank's sorted lists -> the same world - 1 splitters on every rank."""
    world = _world(dist, group)
    if world == 1:
        return np.zeros(0, np.uint64)
    mine = np.concatenate([ops.sample(t, sample_positions(len(t))) for t in sorted_lists]) if sorted_lists else np.zeros(0, np.uint64)
    gathered = [None] * world
    dist.all_gather_object(gathered, mine, group=group)
    return choose_splitters(np.concatenate(gathered), world)


def exchange_by_key_range(ops, dist, local_sorted, splitters: np.ndarray, bits: int, group=None):
    """Steps 3 and 4 for one set: cut at the splitters, all-to-all, merge. Returns (this rank's key range of the global distinct
    set as a sorted tensor, keys sent, keys received)."""
    world = _world(dist, group)
    n = len(local_sorted)
    if world == 1:
        return local_sorted, 0, 0                      # one rank owns every key: its sorted distinct list is the result
    rank = dist.get_rank(group)
    cuts = np.concatenate([[0], ops.lower_bound(local_sorted, splitters), [n]]).astype(np.int64)
    send = np.diff(cuts)
    assert send.min() >= 0 and int(send.sum()) == n
    send_t = ops.counts_tensor(send)
    recv_t = ops.counts_tensor([0] * world)
    dist.all_to_all_single(recv_t, send_t, group=group)                         # who sends me how much
    recv = [int(x) for x in recv_t.tolist()]
    inbox = ops.empty(sum(recv))
    dist.all_to_all_single(inbox, local_sorted, recv, [int(x) for x in send], group=group)   # the keys themselves
    # keys that crossed a link: everything but the piece a rank keeps for itself
    return ops.sort_unique(inbox, bits), n - int(send[rank]), sum(recv) - recv[rank]


def deal_groups(global_counts: np.ndarray, world: int) -> np.ndarray:
    """The 256 top-byte groups dealt to the ranks as contiguous runs of about equal size: bounds[r] .. bounds[r+1] are rank r's
    groups (bounds has world + 1 entries, bounds[0] = 0, bounds[world] = 256). Computed from the SUMMED counts, so all ranks agree."""
    c = np.asarray(global_counts, dtype=np.uint64).astype(np.float64)
    cum = np.cumsum(c)
    total = cum[-1] if cum.size else 0.0
    bounds = [0]
    for r in range(1, world):
        b = int(np.searchsorted(cum, total * r / world, side="left")) + 1 if total > 0 else 256
        bounds.append(min(max(b, bounds[-1]), 256))
    bounds.append(256)
    return np.asarray(bounds, dtype=np.int64)


def exchange_groups(ops, dist, grouped, counts: np.ndarray, bounds: np.ndarray, bits: int, group=None):
    """strategy "partition", one set: the grouped keys of this rank go to the owners of their groups; what arrives is sorted and
    deduplicated. Returns (this rank's range of the global distinct set, keys sent to other ranks, keys received from them)."""
    world = _world(dist, group)
    rank = dist.get_rank(group)
    send = [int(np.asarray(counts[bounds[r]:bounds[r + 1]], dtype=np.uint64).sum()) for r in range(world)]
    assert sum(send) == len(grouped)
    recv_t = ops.counts_tensor([0] * world)
    dist.all_to_all_single(recv_t, ops.counts_tensor(send), group=group)
    recv = [int(x) for x in recv_t.tolist()]
    inbox = ops.empty(sum(recv))
    dist.all_to_all_single(inbox, grouped, recv, send, group=group)
    return ops.sort_unique(inbox, bits), sum(send) - send[rank], sum(recv) - recv[rank]


def _partitioned_sets(ops, dist, key_lists, bits: int, group=None):
    """strategy "partition" for one or two sets that share their ranges: [(range of set i, sent, received)]."""
    if _world(dist, group) == 1:
        return [(ops.sort_unique(k, bits), 0, 0) for k in key_lists]
    parts = [ops.partition(k, bits) for k in key_lists]
    tot = ops.counts_tensor(np.sum([c for _, c in parts], axis=0, dtype=np.uint64).astype(np.int64))
    dist.all_reduce(tot, op=dist.ReduceOp.SUM, group=group)
    bounds = deal_groups(np.asarray(tot.tolist(), dtype=np.uint64), _world(dist, group))
    return [exchange_groups(ops, dist, g, c, bounds, bits, group) for g, c in parts]


def distributed_sorted_unique(ops, dist, bases, k: int, canonical: bool, offsets=None, group=None, strategy: str = "partition"):
    """The distinct k-mers of the union of all ranks' inputs, partitioned by key range over the ranks. Returns
    dict(part=this rank's sorted range, total=number of distinct k-mers overall, ...)."""
    bits = 2 * k
    if strategy == "partition":
        (part, sent, received), = _partitioned_sets(ops, dist, [ops.local_kmers(bases, k, canonical, offsets)], bits, group)
    elif strategy == "sorted":
        local = ops.local_sorted_unique(bases, k, canonical, offsets)
        spl = agree_on_splitters(ops, dist, [local], group)
        part, sent, received = exchange_by_key_range(ops, dist, local, spl, bits, group)
    else:
        raise ValueError("strategy: 'partition' or 'sorted'")
    tot = ops.counts_tensor([len(part)])
    if _world(dist, group) > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.SUM, group=group)
    return {"part": part, "total": int(tot[0]), "sent": sent, "received": received}


def distributed_jaccard(ops, dist, bases1, bases2, k: int, canonical: bool, offsets1=None, offsets2=None, group=None,
                        strategy: str = "partition") -> dict:
    """algorithm::jaccard of the distinct k-mer sets of two inputs that are both spread over the ranks."""
    bits = 2 * k
    if strategy == "partition":
        (pa, sa, ra), (pb, sb, rb) = _partitioned_sets(ops, dist, [ops.local_kmers(bases1, k, canonical, offsets1),
                                                                   ops.local_kmers(bases2, k, canonical, offsets2)], bits, group)
    elif strategy == "sorted":
        a = ops.local_sorted_unique(bases1, k, canonical, offsets1)
        b = ops.local_sorted_unique(bases2, k, canonical, offsets2)
        spl = agree_on_splitters(ops, dist, [a, b], group)
        pa, sa, ra = exchange_by_key_range(ops, dist, a, spl, bits, group)
        pb, sb, rb = exchange_by_key_range(ops, dist, b, spl, bits, group)
    else:
        raise ValueError("strategy: 'partition' or 'sorted'")
    inter, union, n1, n2 = ops.set_sizes(pa, pb, bits)
    t = ops.counts_tensor([inter, union, n1, n2, sa + sb, ra + rb])
    if _world(dist, group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    inter, union, n1, n2, sent, received = (int(x) for x in t.tolist())
    return {"intersection": inter, "union": union, "size1": n1, "size2": n2, "jaccard": inter / union if union else 0.0,
            "keys_exchanged": sent, "keys_received": received, "bytes_exchanged": 8 * sent, "strategy": strategy}
