"""Host-side convenience over the C ABI: numpy in, numpy out (tests, bench, smoke).

The reference-facing host layer is the C++ shim include/biolib_gpu.hpp (same constructor surface as
wrapper::kmer_view / minimizer_view / super_kmer_view); this module is the Python harness equivalent and
calls exactly the same entry points.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _capi
from ._capi import F_DEVICE_INPUT, F_DEVICE_OUTPUT, F_EMIT_HASH, KmerOut, Params, SkOut, Stats


class BlgError(RuntimeError):
    def __init__(self, status: int, msg: str):
        super().__init__(f"biolib_gpu status {status}: {msg}")
        self.status = status


def _as_bytes_array(seq) -> np.ndarray:
    if isinstance(seq, np.ndarray):
        return np.ascontiguousarray(seq.view(np.uint8) if seq.dtype != np.uint8 else seq)
    if isinstance(seq, str):
        seq = seq.encode("latin-1")
    return np.frombuffer(bytes(seq), dtype=np.uint8)


class Context:
    """One GPU, one stream (blg_ctx). Not thread-safe."""

    def __init__(self, device: int = 0):
        self.L = _capi.lib()
        h = C.c_void_p()
        st = self.L.blg_create(device, C.byref(h))
        if st != _capi.BLG_OK:
            raise BlgError(st, self.L.blg_strerror(st).decode())
        self.h = h
        self.device = device

    def close(self):
        if getattr(self, "h", None):
            self.L.blg_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, st: int):
        if st != _capi.BLG_OK:
            raise BlgError(st, f"{self.L.blg_strerror(st).decode()}: {self.L.blg_last_error(self.h).decode()}")

    def set_stream(self, cuda_stream: int | None):
        self._check(self.L.blg_set_stream(self.h, C.c_void_p(cuda_stream or 0)))

    def stats(self) -> dict:
        s = Stats()
        self._check(self.L.blg_last_stats(self.h, C.byref(s)))
        return {f: getattr(s, f) for f, _ in Stats._fields_}

    def device_info(self):
        n = C.c_int()
        buf = C.create_string_buffer(128)
        self._check(self.L.blg_device_info(self.h, C.byref(n), buf, 128))
        return n.value, buf.value.decode()

    # ------------------------------------------------------------------ host in, host out
    def kmers(self, seq, k: int, canonical: bool = False, offsets=None, syncmer=None, hash_sample=None, runs=False) -> dict:
        """Valid k-mers in position order; syncmer=(m, start_offset, end_offset) keeps only those that
        sampler::syncmer_sampler over hash::minimizer_position_extractor(k, m) would (blg_syncmers).
        runs: BLG_F_RUNS (blg_kmers only): no position array comes back, it is rebuilt here from the run list."""
        b = _as_bytes_array(seq)
        o = None if offsets is None else np.ascontiguousarray(offsets, dtype=np.uint64)
        p = Params(k=k, m=0 if syncmer is None else syncmer[0], canonical=int(canonical), flags=_capi.F_RUNS if runs else 0,
                   seed=0 if hash_sample is None else hash_sample[0] & (2**64 - 1))
        out = KmerOut()
        args = (self.h, b.ctypes.data, b.size, o.ctypes.data if o is not None else None, 0 if o is None else len(o) - 1, C.byref(p))
        if hash_sample is not None:
            self._check(self.L.blg_hash_sampled_kmers(*args, int(hash_sample[1]), C.byref(out)))
        elif syncmer is None:
            self._check(self.L.blg_kmers(*args, C.byref(out)))
        else:
            self._check(self.L.blg_syncmers(*args, int(syncmer[1]), int(syncmer[2]), C.byref(out)))
        n = out.count
        try:
            if runs:
                rf, rs = _copy(out.run_first_record, out.n_runs, np.uint64), _copy(out.run_start, out.n_runs, np.uint64)
                pos = expand_kmer_runs(rf, rs, n)
            else:
                pos = _copy(out.position, n, np.uint64)
            res = dict(lo=_copy(out.value_lo, n, np.uint64), pos=pos,
                       hi=_copy(out.value_hi, n, np.uint64) if out.value_hi else np.zeros(n, np.uint64))
            if runs:
                res["n_runs"] = int(out.n_runs)
        finally:
            self.L.blg_free_kmer_out(self.h, C.byref(out))
        return res

    def super_kmers(self, seq, k: int, m: int, seed: int = 0, canonical: bool = False, offsets=None, want_hash=True,
                    win_range=None, minimizers_only=False, want_id=False, pos32=False, runs=False, tight=False) -> dict:
        """pos32: ask for the compact record layout (BLG_F_POS32: 32-bit positions relative to the range start); the
        returned mm_start / sk_start are widened back to absolute uint64 so that both layouts compare equal.
        tight (with runs): BLG_F_TIGHT, 6-byte records; unpacked here into the same value / size / mm_pos arrays."""
        b = _as_bytes_array(seq)
        o = None if offsets is None else np.ascontiguousarray(offsets, dtype=np.uint64)
        p = Params(k=k, m=m, canonical=int(canonical), flags=(F_EMIT_HASH if want_hash else 0) | (_capi.F_EMIT_ID if want_id else 0) | (_capi.F_POS32 if pos32 else 0) |
                   (_capi.F_RUNS if runs else 0) | (_capi.F_TIGHT if tight else 0), seed=seed & (2**64 - 1))
        out = SkOut()
        args = (self.h, b.ctypes.data, b.size, o.ctypes.data if o is not None else None, 0 if o is None else len(o) - 1)
        if win_range is not None:
            st = self.L.blg_super_kmers_range(*args, int(win_range[0]), int(win_range[1]), C.byref(p), C.byref(out))
        elif minimizers_only:
            st = self.L.blg_minimizers(*args, C.byref(p), C.byref(out))
        else:
            st = self.L.blg_super_kmers(*args, C.byref(p), C.byref(out))
        self._check(st)
        n = out.count
        try:
            if tight:
                # BLG_F_TIGHT: value = lo32 | hi8 << 32, size_pos = (size - 1) | mm_pos << 4
                value, sz, mp = unpack_tight(_copy(out.value_lo32, n, np.uint32), _copy(out.value_hi8, n, np.uint8) if out.value_hi8 else None,
                                             _copy(out.size_pos, n, np.uint8))
            if runs:
                # BLG_F_RUNS: no positions came back; rebuild them from the run list (what a streaming consumer does on the fly)
                rf, rs = _copy(out.run_first_record, out.n_runs, np.uint64), _copy(out.run_start, out.n_runs, np.uint64)
                if not tight:
                    sz, mp = _copy(out.size, n, np.uint8), _copy(out.mm_pos, n, np.uint8)
                mm = expand_runs(rf, rs, sz, mp)
            elif pos32:
                base = 0 if win_range is None else min(int(win_range[0]), max(b.size - k + 1, 0))
                mm = _copy(out.mm_start, n, np.uint32).astype(np.uint64) + np.uint64(base)
            else:
                mm = _copy(out.mm_start, n, np.uint64)
            res = dict(n=n, value=value if tight else _copy(out.value, n, np.uint64), mm_start=mm)
            if runs:
                res["n_runs"] = int(out.n_runs)
            if not minimizers_only:
                res["size"] = sz if tight else _copy(out.size, n, np.uint8)
                res["mm_pos"] = mp if tight else _copy(out.mm_pos, n, np.uint8)
                res["sk_start"] = res["mm_start"] - res["mm_pos"].astype(np.uint64)
            if want_hash:
                res["hash"] = _copy(out.hash, n, np.uint64)
            if want_id:
                res["id"] = _copy(out.id, n, np.uint64)
        finally:
            self.L.blg_free_sk_out(self.h, C.byref(out))
        return res

    # ------------------------------------------------------------------ sort / unique / Jaccard
    def sorted_unique_kmers(self, seq, k: int, canonical: bool = False, offsets=None, drop_last: bool = False) -> np.ndarray:
        """blg_sorted_unique_kmers: the distinct k-mer values in increasing order (external_memory_vector + ordered_unique_sampler)."""
        b = _as_bytes_array(seq)
        o = None if offsets is None else np.ascontiguousarray(offsets, dtype=np.uint64)
        p = Params(k=k, m=0, canonical=int(canonical), flags=_capi.F_REF_DROP_LAST if drop_last else 0, seed=0)
        out = KmerOut()
        self._check(self.L.blg_sorted_unique_kmers(self.h, b.ctypes.data, b.size, o.ctypes.data if o is not None else None,
                                                   0 if o is None else len(o) - 1, C.byref(p), C.byref(out)))
        try:
            return _copy(out.value_lo, out.count, np.uint64)
        finally:
            self.L.blg_free_kmer_out(self.h, C.byref(out))

    def kmer_jaccard(self, seq1, seq2, k: int, canonical: bool = False, offsets1=None, offsets2=None, drop_last: bool = False):
        """blg_kmer_jaccard: (intersection, union, size1, size2) of the two distinct k-mer sets (algorithm::jaccard)."""
        b1, b2 = _as_bytes_array(seq1), _as_bytes_array(seq2)
        o1 = None if offsets1 is None else np.ascontiguousarray(offsets1, dtype=np.uint64)
        o2 = None if offsets2 is None else np.ascontiguousarray(offsets2, dtype=np.uint64)
        p = Params(k=k, m=0, canonical=int(canonical), flags=_capi.F_REF_DROP_LAST if drop_last else 0, seed=0)
        out = _capi.JaccardOut()
        self._check(self.L.blg_kmer_jaccard(self.h, b1.ctypes.data, b1.size, o1.ctypes.data if o1 is not None else None, 0 if o1 is None else len(o1) - 1,
                                            b2.ctypes.data, b2.size, o2.ctypes.data if o2 is not None else None, 0 if o2 is None else len(o2) - 1,
                                            C.byref(p), C.byref(out)))
        return out.intersection, out.union_size, out.size1, out.size2

    # ------------------------------------------------------------------ payloads
    def super_kmer_payloads(self, seq, k: int, m: int, seed: int, canonical: bool, n_buckets: int, offsets=None) -> dict:
        """blg_super_kmer_payloads: records grouped stably by hash % n_buckets + their bases as one 2-bit packed stream
        (bit::packed::vector<uint64_t>(2) layout)."""
        b = _as_bytes_array(seq)
        o = None if offsets is None else np.ascontiguousarray(offsets, dtype=np.uint64)
        p = Params(k=k, m=m, canonical=int(canonical), flags=0, seed=seed & (2**64 - 1))
        out = _capi.PayloadOut()
        self._check(self.L.blg_super_kmer_payloads(self.h, b.ctypes.data, b.size, o.ctypes.data if o is not None else None,
                                                   0 if o is None else len(o) - 1, C.byref(p), n_buckets, C.byref(out)))
        n = out.count
        try:
            return dict(n=n, total_bases=out.total_bases, bucket_offsets=_copy(out.bucket_offsets, n_buckets + 1, np.uint64),
                        value=_copy(out.value, n, np.uint64), mm_start=_copy(out.mm_start, n, np.uint64), size=_copy(out.size, n, np.uint8),
                        mm_pos=_copy(out.mm_pos, n, np.uint8), payload_start=_copy(out.payload_start, n, np.uint64),
                        words=_copy(out.words, out.n_words, np.uint64))
        finally:
            self.L.blg_free_payload_out(self.h, C.byref(out))

    # ------------------------------------------------------------------ ingest
    def pack_bases(self, seq):
        """Host-side packer blg_pack_bases: (codes uint32[ceil(n/16)], break_start uint64[], break_len uint32[])."""
        b = _as_bytes_array(seq)
        codes = np.empty((b.size + 15) // 16, np.uint32)
        cap = 1024
        while True:
            bs, bl = np.empty(cap, np.uint64), np.empty(cap, np.uint32)
            nb = C.c_size_t()
            st = self.L.blg_pack_bases(b.ctypes.data, b.size, codes.ctypes.data, bs.ctypes.data, bl.ctypes.data, cap, C.byref(nb))
            if st == _capi.BLG_E_CAPACITY:
                cap = nb.value
                continue
            self._check(st)
            return codes, bs[:nb.value].copy(), bl[:nb.value].copy()

    def super_kmers_packed(self, codes, n_bases, break_start, break_len, k, m, seed=0, canonical=False, offsets=None, want_hash=True,
                           win_range=None, pos32=False) -> dict:
        """blg_super_kmers_packed: 2-bit packed input (0.25 B per base over PCIe) + break intervals."""
        codes = np.ascontiguousarray(codes, np.uint32)
        bs, bl = np.ascontiguousarray(break_start, np.uint64), np.ascontiguousarray(break_len, np.uint32)
        o = None if offsets is None else np.ascontiguousarray(offsets, dtype=np.uint64)
        seq = _capi.PackedSeq(codes=codes.ctypes.data, n_bases=n_bases, break_start=bs.ctypes.data if bs.size else None,
                              break_len=bl.ctypes.data if bl.size else None, n_breaks=bs.size)
        p = Params(k=k, m=m, canonical=int(canonical), flags=(F_EMIT_HASH if want_hash else 0) | (_capi.F_POS32 if pos32 else 0), seed=seed & (2**64 - 1))
        out = SkOut()
        lo, hi = (0, 2**64 - 1) if win_range is None else win_range
        self._check(self.L.blg_super_kmers_packed(self.h, C.byref(seq), o.ctypes.data if o is not None else None, 0 if o is None else len(o) - 1,
                                                  int(lo), int(hi), C.byref(p), C.byref(out)))
        n = out.count
        try:
            if pos32:
                base = 0 if win_range is None else min(int(win_range[0]), max(n_bases - k + 1, 0))
                mm = _copy(out.mm_start, n, np.uint32).astype(np.uint64) + np.uint64(base)
            else:
                mm = _copy(out.mm_start, n, np.uint64)
            res = dict(n=n, value=_copy(out.value, n, np.uint64), mm_start=mm, size=_copy(out.size, n, np.uint8), mm_pos=_copy(out.mm_pos, n, np.uint8))
            res["sk_start"] = res["mm_start"] - res["mm_pos"].astype(np.uint64)
            if want_hash:
                res["hash"] = _copy(out.hash, n, np.uint64)
        finally:
            self.L.blg_free_sk_out(self.h, C.byref(out))
        return res

    def super_kmers_packed_raw(self, codes_ptr: int, n_bases: int, break_start_ptr: int, break_len_ptr: int, n_breaks: int, k: int, m: int,
                               seed: int, canonical: bool, value_ptr: int, mm_start_ptr: int, size_ptr: int, mm_pos_ptr: int, capacity: int,
                               flags: int = 0, win_range=None, run_ptrs=None, run_capacity: int = 0, tight_ptrs=None) -> int:
        """blg_super_kmers_packed with caller-owned (pinned) host arrays; with F_RUNS the number of runs is left in self.last_n_runs.
        With F_TIGHT pass tight_ptrs=(value_lo32_ptr, value_hi8_ptr, size_pos_ptr) (value / size / mm_pos pointers are ignored)."""
        seq = _capi.PackedSeq(codes=codes_ptr, n_bases=n_bases, break_start=break_start_ptr or None, break_len=break_len_ptr or None,
                              n_breaks=n_breaks)
        p = Params(k=k, m=m, canonical=int(canonical), flags=flags, seed=seed & (2**64 - 1))
        out = SkOut(count=0, capacity=capacity, value=value_ptr, mm_start=mm_start_ptr or None, size=size_ptr or None,
                    mm_pos=mm_pos_ptr or None, hash=None, id=None, owner=None, n_runs=0, run_capacity=run_capacity,
                    run_first_record=run_ptrs[0] if run_ptrs else None, run_start=run_ptrs[1] if run_ptrs else None,
                    value_lo32=tight_ptrs[0] if tight_ptrs else None, value_hi8=(tight_ptrs[1] or None) if tight_ptrs else None,
                    size_pos=tight_ptrs[2] if tight_ptrs else None)
        lo, hi = (0, 2**64 - 1) if win_range is None else win_range
        self._check(self.L.blg_super_kmers_packed(self.h, C.byref(seq), None, 0, int(lo), int(hi), C.byref(p), C.byref(out)))
        self.last_n_runs = int(out.n_runs)
        return out.count

    def parse_fastx(self, text, fmt: int = -1):
        """blg_parse_fastx (device record splitter): FASTA / FASTQ text -> (bases uint8, seq_offsets uint64[n_seq + 1])."""
        t = _as_bytes_array(text)
        out = _capi.FastxOut()
        self._check(self.L.blg_parse_fastx(self.h, t.ctypes.data, t.size, fmt, 0, C.byref(out)))
        try:
            return _copy(out.bases, out.n_bases, np.uint8), _copy(out.seq_offsets, out.n_seq + 1, np.uint64)
        finally:
            self.L.blg_free_fastx_out(self.h, C.byref(out))

    def parse_fastx_device(self, text_ptr: int, n_text: int, fmt: int = -1, device_input: bool = True):
        """Device-resident form: returns a FastxOut whose bases / seq_offsets are device pointers (free with free_fastx)."""
        out = _capi.FastxOut()
        flags = (F_DEVICE_INPUT if device_input else 0) | F_DEVICE_OUTPUT
        self._check(self.L.blg_parse_fastx(self.h, C.c_void_p(text_ptr), n_text, fmt, flags, C.byref(out)))
        return out

    def free_fastx(self, out):
        self.L.blg_free_fastx_out(self.h, C.byref(out))

    def minimizers(self, seq, k, m, seed=0, canonical=False, offsets=None, want_hash=True, want_id=False) -> dict:
        return self.super_kmers(seq, k, m, seed, canonical, offsets, want_hash, minimizers_only=True, want_id=want_id)

    def hash64(self, values, seed: int) -> np.ndarray:
        v = np.ascontiguousarray(values, dtype=np.uint64)
        out = np.empty_like(v)
        self._check(self.L.blg_hash64(self.h, v.ctypes.data, v.size, seed & (2**64 - 1), 0, out.ctypes.data))
        return out

    def syncmers(self, seq, k: int, m: int, start_offset: int = 0, end_offset=None, canonical: bool = False, offsets=None) -> dict:
        """sampler::syncmer_sampler (reference include/syncmer_sampler.hpp:130-137); closed syncmers by default."""
        return self.kmers(seq, k, canonical, offsets, syncmer=(m, start_offset, k - m if end_offset is None else end_offset))

    @staticmethod
    def hash_sampler_threshold(sampling_rate: float) -> int:
        """hash_sampler's constructor (reference include/hash_sampler.hpp:76-77): same check, same double product. At rate 1.0
        the reference's cast of 2^64 to uint64_t is out of range (undefined behaviour; 0 with g++ -O2 here, i.e. nothing kept):
        this helper clamps to 2^64 - 1 instead -- a deliberate divergence."""
        if sampling_rate > 1 or sampling_rate < 0:
            raise ValueError("[hash_sampler] Invalid sampling rate")
        return min(int(sampling_rate * float(2**64 - 1)), 2**64 - 1)

    def hash_sampled_kmers(self, seq, k: int, seed: int, sampling_rate: float, canonical: bool = False, offsets=None) -> dict:
        """sampler::hash_sampler over the k-mer values (reference include/hash_sampler.hpp:138-141)."""
        return self.kmers(seq, k, canonical, offsets, hash_sample=(seed, self.hash_sampler_threshold(sampling_rate)))

    def minimizer_positions(self, kmer_values, k: int, m: int) -> np.ndarray:
        """hash::minimizer_position_extractor for an array of k-mer values (k <= 32)."""
        v = np.ascontiguousarray(kmer_values, dtype=np.uint64)
        out = np.empty(v.size, dtype=np.uint8)
        self._check(self.L.blg_minimizer_positions(self.h, v.ctypes.data, v.size, k, m, 0, out.ctypes.data))
        return out

    # ------------------------------------------------------------------ device resident (raw pointers)
    def super_kmers_device(self, d_bases: int, n_bases: int, k: int, m: int, seed: int, canonical: bool, d_value: int,
                           d_mm_start: int, d_size: int, d_mm_pos: int, capacity: int, d_offsets: int = 0, n_seq: int = 0,
                           win_range=None, d_hash: int = 0, async_call: bool = False) -> int:
        """Everything already in HBM: enqueue on the context stream, return the record count (0 with async_call: the kernel is
        only enqueued, collect the count with async_count())."""
        flags = F_DEVICE_INPUT | F_DEVICE_OUTPUT | (F_EMIT_HASH if d_hash else 0) | (_capi.F_ASYNC if async_call else 0)
        if not d_size:
            flags |= _capi.F_NO_SIZE
        p = Params(k=k, m=m, canonical=int(canonical), flags=flags, seed=seed & (2**64 - 1))
        out = SkOut(count=0, capacity=capacity, value=d_value, mm_start=d_mm_start, size=d_size or None,
                    mm_pos=d_mm_pos or None, hash=d_hash or None, id=None, owner=None)
        lo, hi = (0, 2**64 - 1) if win_range is None else win_range
        st = self.L.blg_super_kmers_range(self.h, C.c_void_p(d_bases), n_bases, C.c_void_p(d_offsets or 0), n_seq,
                                          int(lo), int(hi), C.byref(p), C.byref(out))
        self._check(st)
        return out.count

    def async_count(self) -> int:
        """blg_async_count: wait for the context's stream, records of the last async call."""
        n = C.c_uint64()
        self._check(self.L.blg_async_count(self.h, C.byref(n)))
        return n.value

    def super_kmers_raw(self, bases_ptr: int, n_bases: int, k: int, m: int, seed: int, canonical: bool, value_ptr: int,
                        mm_start_ptr: int, size_ptr: int, mm_pos_ptr: int, capacity: int, flags: int = 0,
                        offsets_ptr: int = 0, n_seq: int = 0, win_range=None, run_ptrs=None, run_capacity: int = 0, tight_ptrs=None) -> int:
        """Raw-pointer call with caller-owned arrays (host pointers unless the DEVICE flags are set). With F_RUNS pass
        run_ptrs=(first_record_ptr, start_ptr) and run_capacity; the number of runs is left in self.last_n_runs. With F_TIGHT pass
        tight_ptrs=(value_lo32_ptr, value_hi8_ptr, size_pos_ptr) (value / size / mm_pos pointers are ignored)."""
        p = Params(k=k, m=m, canonical=int(canonical), flags=flags, seed=seed & (2**64 - 1))
        out = SkOut(count=0, capacity=capacity, value=value_ptr, mm_start=mm_start_ptr or None, size=size_ptr or None,
                    mm_pos=mm_pos_ptr or None, hash=None, id=None, owner=None, n_runs=0, run_capacity=run_capacity,
                    run_first_record=run_ptrs[0] if run_ptrs else None, run_start=run_ptrs[1] if run_ptrs else None,
                    value_lo32=tight_ptrs[0] if tight_ptrs else None, value_hi8=(tight_ptrs[1] or None) if tight_ptrs else None,
                    size_pos=tight_ptrs[2] if tight_ptrs else None)
        lo, hi = (0, 2**64 - 1) if win_range is None else win_range
        self._check(self.L.blg_super_kmers_range(self.h, C.c_void_p(bases_ptr), n_bases, C.c_void_p(offsets_ptr or 0), n_seq,
                                                 int(lo), int(hi), C.byref(p), C.byref(out)))
        self.last_n_runs = int(out.n_runs)
        return out.count

    def host_alloc(self, nbytes: int) -> int:
        p = C.c_void_p()
        self._check(self.L.blg_host_alloc(self.h, nbytes, C.byref(p)))
        return p.value

    def host_free(self, ptr: int):
        self._check(self.L.blg_host_free(self.h, C.c_void_p(ptr)))

    def kmers_device(self, d_bases: int, n_bases: int, k: int, canonical: bool, d_lo: int, d_hi: int, d_pos: int,
                     capacity: int, d_offsets: int = 0, n_seq: int = 0, syncmer=None, hash_sample=None) -> int:
        p = Params(k=k, m=0 if syncmer is None else syncmer[0], canonical=int(canonical), flags=F_DEVICE_INPUT | F_DEVICE_OUTPUT,
                   seed=0 if hash_sample is None else hash_sample[0] & (2**64 - 1))
        out = KmerOut(count=0, capacity=capacity, value_lo=d_lo, value_hi=d_hi or None, position=d_pos or None, owner=None)
        args = (self.h, C.c_void_p(d_bases), n_bases, C.c_void_p(d_offsets or 0), n_seq, C.byref(p))
        if hash_sample is not None:
            self._check(self.L.blg_hash_sampled_kmers(*args, int(hash_sample[1]), C.byref(out)))
        elif syncmer is None:
            self._check(self.L.blg_kmers(*args, C.byref(out)))
        else:
            self._check(self.L.blg_syncmers(*args, int(syncmer[1]), int(syncmer[2]), C.byref(out)))
        return out.count


def expand_kmer_runs(run_first_record, run_start, n) -> np.ndarray:
    """position of every k-mer from the BLG_F_RUNS outputs of blg_kmers: consecutive positions inside a run."""
    if n == 0:
        return np.zeros(0, np.uint64)
    first = run_first_record.astype(np.int64)
    run_of = np.searchsorted(first, np.arange(n), side="right") - 1
    return (run_start.astype(np.int64)[run_of] + (np.arange(n) - first[run_of])).astype(np.uint64)


def unpack_tight(value_lo32, value_hi8, size_pos):
    """BLG_F_TIGHT records -> (value uint64, size uint8, mm_pos uint8)."""
    value = value_lo32.astype(np.uint64)
    if value_hi8 is not None:
        value |= value_hi8.astype(np.uint64) << np.uint64(32)
    return value, (size_pos & np.uint8(15)) + np.uint8(1), size_pos >> np.uint8(4)


def expand_runs(run_first_record, run_start, size, mm_pos) -> np.ndarray:
    """mm_start of every record from the BLG_F_RUNS outputs: inside a run every record starts where the previous one ended."""
    n = len(size)
    if n == 0:
        return np.zeros(0, np.uint64)
    sz = size.astype(np.int64)
    before = np.cumsum(sz) - sz                                   # bases (k-mer starts) consumed before record r, over all records
    run_of = np.searchsorted(run_first_record.astype(np.int64), np.arange(n), side="right") - 1
    first = run_first_record.astype(np.int64)[run_of]
    sk = run_start.astype(np.int64)[run_of] + before - before[first]
    return (sk + mm_pos.astype(np.int64)).astype(np.uint64)


def _copy(ptr, n, dtype) -> np.ndarray:
    if not n or not ptr:
        return np.zeros(0, dtype)
    nbytes = int(n) * np.dtype(dtype).itemsize
    buf = (C.c_uint8 * nbytes).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype, count=int(n)).copy()
