"""Sort / unique / Jaccard of k-mers across GPUs: SURVEY 8(f) rank 3 at N > 1, the one step of the path where ranks exchange data.

The reference's pipeline (tests/test_jaccard.cpp:55-130) is kmer_view -> sorted external-memory vector ->
ordered_unique_sampler -> algorithm::jaccard. On one GPU that is blg_sorted_unique_kmers / blg_kmer_jaccard. On N GPUs
(one process per GPU, torch.distributed for the plumbing):

  1. every rank extracts, sorts and deduplicates the k-mers of ITS share of the input (blg_kmers into device memory, then
     blg_sort_unique_u64 in place);
  2. the ranks agree on N - 1 key splitters from a regular sample of everybody's sorted list (sample sort; 64 samples per
     rank and list, gathered as host objects);
  3. every rank cuts its sorted list at the splitters (blg_lower_bound_u64) and the pieces travel to the rank that owns the
     key range: ONE all-to-all of 8-byte keys over NVLink per set (NCCL ``all_to_all_single`` with the cut sizes);
  4. every rank merges what it received (blg_sort_unique_u64): rank r now holds the distinct k-mers of key range r of the
     WHOLE input, ranges ascending with the rank, so the concatenation in rank order is the single-GPU result;
  5. Jaccard: both sets use the same splitters, each rank intersects its two ranges (blg_sorted_sets_jaccard) and the
     four counters are summed (one all-reduce of 32 bytes).

The host logic (splitters, cut sizes, the exchange) is independent of where the keys live: `ops` does the per-rank work.
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
        """Distinct k-mer values of this rank's bases, sorted, in a device tensor: blg_kmers into torch-owned memory (values
        only, no positions), then blg_sort_unique_u64 in place. `bases`: host uint8 array or a device uint8 tensor."""
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
        return self.sort_unique(keys[:cnt], 2 * k)

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


def distributed_sorted_unique(ops, dist, bases, k: int, canonical: bool, offsets=None, group=None):
    """The distinct k-mers of the union of all ranks' inputs, partitioned by key range over the ranks. Returns
    dict(part=this rank's sorted range, total=number of distinct k-mers overall, splitters=...)."""
    bits = 2 * k
    local = ops.local_sorted_unique(bases, k, canonical, offsets)
    spl = agree_on_splitters(ops, dist, [local], group)
    part, sent, received = exchange_by_key_range(ops, dist, local, spl, bits, group)
    tot = ops.counts_tensor([len(part)])
    if _world(dist, group) > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.SUM, group=group)
    return {"part": part, "total": int(tot[0]), "splitters": spl, "sent": sent, "received": received, "local_distinct": len(local)}


def distributed_jaccard(ops, dist, bases1, bases2, k: int, canonical: bool, offsets1=None, offsets2=None, group=None) -> dict:
    """algorithm::jaccard of the distinct k-mer sets of two inputs that are both spread over the ranks."""
    bits = 2 * k
    a = ops.local_sorted_unique(bases1, k, canonical, offsets1)
    b = ops.local_sorted_unique(bases2, k, canonical, offsets2)
    spl = agree_on_splitters(ops, dist, [a, b], group)
    pa, sa, ra = exchange_by_key_range(ops, dist, a, spl, bits, group)
    pb, sb, rb = exchange_by_key_range(ops, dist, b, spl, bits, group)
    inter, union, n1, n2 = ops.set_sizes(pa, pb, bits)
    t = ops.counts_tensor([inter, union, n1, n2, sa + sb, ra + rb])
    if _world(dist, group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    inter, union, n1, n2, sent, received = (int(x) for x in t.tolist())
    return {"intersection": inter, "union": union, "size1": n1, "size2": n2, "jaccard": inter / union if union else 0.0,
            "keys_exchanged": sent, "keys_received": received, "bytes_exchanged": 8 * sent}
