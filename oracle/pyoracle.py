"""TEST INFRASTRUCTURE ONLY: ctypes access to the two oracle libraries.

* ``Oracle``  -> oracle/liboracle.so       (plain-C restatement, oracle.c; travels to the GPU box)
* ``Ref``     -> oracle/_ref/libbiolib_ref.so (the reference's own code, ref_tier_a.cpp; built only
                 where /root/reference exists, but the built .so travels)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / ``--impl reference`` legs may
import this module. The product package (biolib_b200) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libbiolib_ref.so")

_u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")


def build(force: bool = False) -> None:
    """Compile oracle.c (always possible) and the Tier A library (only if the reference is here)."""
    if force or not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(os.path.join(HERE, "oracle.c")):
        subprocess.check_call(["make", "-C", HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    if os.path.isdir("/root/reference/include"):
        src = os.path.join(HERE, "ref_tier_a.cpp")
        if force or not os.path.exists(REF_SO) or os.path.getmtime(REF_SO) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", HERE, "_ref/libbiolib_ref.so"], stdout=subprocess.DEVNULL)


def _buf(seq) -> bytes:
    if isinstance(seq, str):
        return seq.encode("latin-1")
    if isinstance(seq, np.ndarray):
        return seq.tobytes()
    return bytes(seq)


def _offs(offsets):
    if offsets is None:
        return None, 0
    o = np.ascontiguousarray(offsets, dtype=np.uint64)
    return o, len(o) - 1


class Oracle:
    """oracle.c"""

    def __init__(self):
        build()
        L = self.lib = C.CDLL(ORACLE_SO)
        L.orc_nt4.restype = C.c_uint8
        L.orc_nt4.argtypes = [C.c_uint8]
        L.orc_hash64_u64.restype = C.c_uint64
        L.orc_hash64_u64.argtypes = [C.c_uint64, C.c_uint64]
        L.orc_remix.restype = C.c_uint64
        L.orc_remix.argtypes = [C.c_uint64]
        L.orc_double_hash64_u64.argtypes = [C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64)]
        L.orc_double_hash64_u128.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64)]
        L.orc_kmer_stream.restype = C.c_size_t
        L.orc_kmer_stream.argtypes = [C.c_char_p, C.c_size_t, C.c_uint, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
        L.orc_kmers.restype = C.c_size_t
        L.orc_kmers.argtypes = [C.c_char_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_uint, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
        L.orc_mmers.argtypes = [C.c_char_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_uint, C.c_uint64, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        for f in (L.orc_super_kmers, L.orc_super_kmers_brute):
            f.restype = C.c_size_t
            f.argtypes = [C.c_char_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_uint, C.c_uint, C.c_uint64, C.c_int,
                          C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
        L.orc_generate.argtypes = [C.c_uint64, C.c_size_t, C.c_void_p]
        L.orc_mutate.restype = C.c_size_t
        L.orc_mutate.argtypes = [C.c_char_p, C.c_size_t, C.c_double, C.c_double, C.c_double, C.c_uint64, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]

    def nt4_table(self):
        return np.array([self.lib.orc_nt4(i) for i in range(256)], dtype=np.uint8)

    def hash64(self, v, seed):
        return int(self.lib.orc_hash64_u64(v & (2**64 - 1), seed & (2**64 - 1)))

    def double_hash64(self, v, seed):
        out = (C.c_uint64 * 2)()
        self.lib.orc_double_hash64_u64(v, seed, out)
        return int(out[0]), int(out[1])

    def double_hash64_u128(self, lo, hi, seed):
        out = (C.c_uint64 * 2)()
        self.lib.orc_double_hash64_u128(lo, hi, seed, out)
        return int(out[0]), int(out[1])

    def remix(self, z):
        return int(self.lib.orc_remix(z))

    def minimizer_positions(self, values, k, m):
        v = np.ascontiguousarray(values, dtype=np.uint64)
        out = np.empty(v.size, dtype=np.uint64)
        self.lib.orc_minimizer_positions.argtypes = [C.c_void_p, C.c_size_t, C.c_uint, C.c_uint, C.c_void_p]
        self.lib.orc_minimizer_positions.restype = None
        self.lib.orc_minimizer_positions(v.ctypes.data, v.size, k, m, out.ctypes.data)
        return out

    def kmer_stream(self, seq, k, canonical):
        b = _buf(seq)
        cap = len(b) + 1
        lo = np.zeros(cap, np.uint64); hi = np.zeros(cap, np.uint64); has = np.zeros(cap, np.uint8)
        pos = np.zeros(cap, np.uint64); idv = np.zeros(cap, np.uint64)
        n = self.lib.orc_kmer_stream(b, len(b), k, int(canonical), lo.ctypes.data, hi.ctypes.data, has.ctypes.data, pos.ctypes.data, idv.ctypes.data, cap)
        return dict(lo=lo[:n], hi=hi[:n], has=has[:n], pos=pos[:n], id=idv[:n])

    def kmers(self, seq, k, canonical, offsets=None):
        b = _buf(seq)
        o, ns = _offs(offsets)
        cap = len(b) + 1
        lo = np.zeros(cap, np.uint64); hi = np.zeros(cap, np.uint64); pos = np.zeros(cap, np.uint64)
        n = self.lib.orc_kmers(b, len(b), o.ctypes.data if o is not None else None, ns, k, int(canonical), lo.ctypes.data, hi.ctypes.data, pos.ctypes.data, cap)
        return dict(lo=lo[:n], hi=hi[:n], pos=pos[:n])

    def syncmers(self, seq, k, m, soff, eoff, canonical, offsets=None):
        """sampler::syncmer_sampler::const_iterator::find_first_syncmer (reference include/syncmer_sampler.hpp:130-137)
        in the batched form blg_syncmers returns: the valid k-mers whose minimizer_position_extractor(k, m) offset
        (include/kmer_view.hpp:266-283) is soff or eoff, in position order."""
        km = self.kmers(seq, k, canonical, offsets)
        p = self.minimizer_positions(km["lo"], k, m)
        keep = (p == soff) | (p == eoff)
        return {key: v[keep] for key, v in km.items()}

    def hash_sampled_kmers(self, seq, k, seed, threshold, canonical, offsets=None):
        """sampler::hash_sampler::const_iterator::find_first_kmer (reference include/hash_sampler.hpp:138-141) over the valid
        k-mer values: kept iff hash64(value, seed) < threshold (threshold as the constructor computes it, :77)."""
        km = self.kmers(seq, k, canonical, offsets)
        h = np.array([self.hash64(int(x), seed) for x in km["lo"]], dtype=np.uint64) if len(km["lo"]) else np.zeros(0, np.uint64)
        keep = h < np.uint64(threshold)
        return {key: v[keep] for key, v in km.items()}

    def syncmer_stream(self, seq, k, m, soff, eoff, canonical):
        """The reference's record stream for one sequence: syncmer_sampler over kmer_view (K1-K4 apply: the final k-mer
        of a sequence ending in a valid base never appears; nullopt records have offset k + 1 and never pass)."""
        ks = self.kmer_stream(seq, k, canonical)
        p = self.minimizer_positions(ks["lo"], k, m)
        keep = (ks["has"] == 1) & ((p == soff) | (p == eoff))
        return dict(lo=ks["lo"][keep], pos=ks["pos"][keep], id=ks["id"][keep])

    def mmers(self, seq, m, seed, canonical, offsets=None):
        b = _buf(seq)
        o, ns = _offs(offsets)
        n = len(b)
        val = np.zeros(n, np.uint64); h = np.zeros(n, np.uint64); ok = np.zeros(n, np.uint8)
        self.lib.orc_mmers(b, n, o.ctypes.data if o is not None else None, ns, m, seed, int(canonical), val.ctypes.data, h.ctypes.data, ok.ctypes.data)
        return dict(value=val, hash=h, valid=ok)

    def super_kmers(self, seq, k, m, seed, canonical, offsets=None, brute=False, want=("value", "hash", "mm_start", "sk_start", "size"), cap=None):
        b = _buf(seq)
        o, ns = _offs(offsets)
        n = len(b)
        if cap is None:
            cap = max(n - k + 1, 0) + 1
        arr = dict(value=np.zeros(cap if "value" in want else 0, np.uint64), hash=np.zeros(cap if "hash" in want else 0, np.uint64),
                   mm_start=np.zeros(cap if "mm_start" in want else 0, np.uint64), sk_start=np.zeros(cap if "sk_start" in want else 0, np.uint64),
                   size=np.zeros(cap if "size" in want else 0, np.uint8))
        f = self.lib.orc_super_kmers_brute if brute else self.lib.orc_super_kmers
        ptr = lambda k_: arr[k_].ctypes.data if k_ in want else None
        cnt = f(b, n, o.ctypes.data if o is not None else None, ns, k, m, seed, int(canonical),
                ptr("value"), ptr("hash"), ptr("mm_start"), ptr("sk_start"), ptr("size"), cap if want else 0)
        out = {k_: v[:min(cnt, cap)] for k_, v in arr.items() if k_ in want}
        out["n"] = cnt
        return out

    def generate(self, seed, length):
        out = np.empty(length, np.uint8)
        self.lib.orc_generate(seed, length, out.ctypes.data)
        return out

    def parse_fastx(self, text, fastq=None):
        """Restatement of what kseq_read() yields for well-formed input (reference tests/kseq.h:185-240): FASTA = a line whose
        first byte is '>' starts a record, every other line is sequence; FASTQ = four lines per record, the second is the
        sequence. '\\n' / '\\r' dropped. Returns (bases uint8, offsets uint64[n_seq + 1]). Pure Python: small inputs only."""
        t = bytes(_buf(text))
        if fastq is None:
            fastq = t[:1] == b"@"
        lines = t.split(b"\n")
        if lines and lines[-1] == b"":
            lines.pop()
        seqs, cur = [], None
        if fastq:
            for i, ln in enumerate(lines):
                if i % 4 == 0:
                    seqs.append(b"")
                elif i % 4 == 1:
                    seqs[-1] = ln.replace(b"\r", b"")
        else:
            for ln in lines:
                if ln[:1] == b">":
                    seqs.append(b"")
                elif seqs:
                    seqs[-1] += ln.replace(b"\r", b"")
        offs = np.concatenate([[0], np.cumsum([len(x) for x in seqs])]).astype(np.uint64)
        return np.frombuffer(b"".join(seqs), np.uint8).copy(), offs

    def sorted_unique_kmers(self, seq, k, canonical, offsets=None, drop_last=False):
        """Restatement of test_jaccard's vector -> sort -> ordered_unique_sampler: distinct k-mer values in increasing order.
        drop_last: leave out the last k-mer of every sequence that ends in a valid base (kmer_view's K1)."""
        r = self.kmers(seq, k, canonical, offsets=offsets)
        lo, pos = r["lo"], r["pos"].astype(np.int64)
        if drop_last:
            n = len(_buf(seq))
            ends = np.array([n], np.int64) if offsets is None else np.asarray(offsets, np.int64)[1:]
            lo = lo[~np.isin(pos + k, ends)]
        return np.unique(lo)

    def jaccard(self, seq1, seq2, k, canonical, offsets1=None, offsets2=None, drop_last=False):
        a = self.sorted_unique_kmers(seq1, k, canonical, offsets1, drop_last)
        b = self.sorted_unique_kmers(seq2, k, canonical, offsets2, drop_last)
        inter = int(np.intersect1d(a, b, assume_unique=True).size)
        return inter, int(a.size + b.size - inter), int(a.size), int(b.size)

    def payloads(self, seq, k, m, seed, canonical, n_buckets, offsets=None):
        """Restatement of blg_super_kmer_payloads over this oracle's records: stable grouping by hash % n_buckets, every
        record's bases [sk_start, sk_start + size + k - 1) as 2-bit codes appended to one stream packed like
        bit::packed::vector<uint64_t>(2) (reference include/packed_vector.hpp:134-160,311-318)."""
        r = self.super_kmers(seq, k, m, seed, canonical, offsets=offsets)
        b = np.frombuffer(bytes(_buf(seq)), np.uint8)
        bucket = (r["hash"] % np.uint64(n_buckets)).astype(np.int64)
        order = np.argsort(bucket, kind="stable")
        res = {key: r[key][order] for key in ("value", "mm_start", "sk_start", "size")}
        res["mm_pos"] = (res["mm_start"] - res["sk_start"]).astype(np.uint8)
        res["bucket_offsets"] = np.concatenate([[0], np.cumsum(np.bincount(bucket, minlength=n_buckets))]).astype(np.uint64)
        lens = res["size"].astype(np.int64) + (k - 1)
        res["payload_start"] = (np.cumsum(lens) - lens).astype(np.uint64)
        total = int(lens.sum())
        idx = np.repeat(res["sk_start"].astype(np.int64), lens) + (np.arange(total) - np.repeat(np.cumsum(lens) - lens, lens))
        codes = self.nt4_table()[b[idx]].astype(np.uint64)
        res["codes"] = codes.astype(np.uint8)
        words = np.zeros(total * 2 // 64 + 1, np.uint64)
        e = np.arange(total, dtype=np.int64)
        np.bitwise_or.at(words, e // 32, codes << (62 - 2 * (e % 32)).astype(np.uint64))
        res["words"], res["total_bases"], res["n"] = words, total, len(order)
        return res

    def pack_bases(self, seq):
        """2-bit wire format of blg_super_kmers_packed: (codes uint32, break_start uint64, break_len uint32)."""
        b = np.frombuffer(bytes(_buf(seq)), np.uint8)
        code = self.nt4_table()[b].astype(np.uint32)
        bad = code > 3
        c2 = np.where(bad, 0, code).astype(np.uint32)
        pad = (-len(c2)) % 16
        c2 = np.concatenate([c2, np.zeros(pad, np.uint32)]).reshape(-1, 16)
        codes = (c2 << (2 * np.arange(16, dtype=np.uint32))).sum(axis=1).astype(np.uint32)
        d = np.diff(np.concatenate([[0], bad.astype(np.int8), [0]]))
        st, en = np.nonzero(d == 1)[0], np.nonzero(d == -1)[0]
        return codes, st.astype(np.uint64), (en - st).astype(np.uint32)

    def mutate(self, seq, mutp, indelf, extp, seed):
        b = _buf(seq)
        cap = len(b) * 4 + 64
        out = np.empty(cap, np.uint8)
        counts = np.zeros(3, np.uint64); tot = C.c_int64(0)
        n = self.lib.orc_mutate(b, len(b), mutp, indelf, extp, seed, out.ctypes.data, cap, counts.ctypes.data, C.byref(tot))
        if n > cap:  # long insertions: the call reports the size it needs
            cap = n
            out = np.empty(cap, np.uint8)
            n = self.lib.orc_mutate(b, len(b), mutp, indelf, extp, seed, out.ctypes.data, cap, counts.ctypes.data, C.byref(tot))
        return out[:n].copy(), counts, tot.value


class Ref:
    """ref_tier_a.cpp: the reference executed verbatim. ``Ref.available()`` is False on a box where
    neither /root/reference nor a prebuilt oracle/_ref exists."""

    @staticmethod
    def available() -> bool:
        try:
            build()
        except Exception:
            pass
        return os.path.exists(REF_SO)

    def __init__(self):
        build()
        L = self.lib = C.CDLL(REF_SO)
        L.ref_hash64_u64.restype = C.c_uint64
        L.ref_hash64_u64.argtypes = [C.c_uint64, C.c_uint64]
        L.ref_remix.restype = C.c_uint64
        L.ref_remix.argtypes = [C.c_uint64]
        L.ref_double_hash64_u64.argtypes = [C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64)]
        L.ref_double_hash64_u128.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64)]
        for f in (L.ref_kmer_view_u64, L.ref_kmer_view_u128):
            f.restype = C.c_size_t
            f.argtypes = [C.c_char_p, C.c_size_t, C.c_uint, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
        L.ref_mmer_stream.argtypes = [C.c_char_p, C.c_size_t, C.c_uint, C.c_uint, C.c_uint64, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_minimizer_view_literal_count.restype = C.c_size_t
        L.ref_minimizer_view_literal_count.argtypes = [C.c_char_p, C.c_size_t, C.c_uint, C.c_uint, C.c_uint64, C.c_int]
        L.ref_generate.argtypes = [C.c_uint64, C.c_size_t, C.c_void_p]
        L.ref_jaccard_sorted.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p]
        L.ref_ordered_unique.restype = C.c_size_t
        L.ref_ordered_unique.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
        L.ref_packed_vector_words.restype = C.c_size_t
        L.ref_packed_vector_words.argtypes = [C.c_void_p, C.c_size_t, C.c_uint, C.c_void_p, C.c_size_t]
        L.ref_kseq_parse.restype = C.c_size_t
        L.ref_kseq_parse.argtypes = [C.c_char_p, C.c_size_t, C.c_void_p, C.c_void_p]
        L.ref_mutate.restype = C.c_size_t
        L.ref_mutate.argtypes = [C.c_char_p, C.c_size_t, C.c_double, C.c_double, C.c_double, C.c_uint64, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
        L.ref_nt4_table.argtypes = [C.c_void_p]
        L.ref_super_kmers_sums.restype = C.c_size_t
        L.ref_super_kmers_sums.argtypes = [C.c_char_p, C.c_size_t, C.c_uint, C.c_uint, C.c_uint64, C.c_int, C.c_void_p]

    def nt4_table(self):
        t = np.zeros(256, np.uint8)
        self.lib.ref_nt4_table(t.ctypes.data)
        return t

    def hash64(self, v, seed):
        return int(self.lib.ref_hash64_u64(v & (2**64 - 1), seed & (2**64 - 1)))

    def double_hash64(self, v, seed):
        out = (C.c_uint64 * 2)()
        self.lib.ref_double_hash64_u64(v, seed, out)
        return int(out[0]), int(out[1])

    def double_hash64_u128(self, lo, hi, seed):
        out = (C.c_uint64 * 2)()
        self.lib.ref_double_hash64_u128(lo, hi, seed, out)
        return int(out[0]), int(out[1])

    def remix(self, z):
        return int(self.lib.ref_remix(z))

    def minimizer_positions(self, values, k, m, has_value=None):
        v = np.ascontiguousarray(values, dtype=np.uint64)
        hv = None if has_value is None else np.ascontiguousarray(has_value, dtype=np.uint8)
        out = np.empty(v.size, dtype=np.uint64)
        self.lib.ref_minimizer_positions.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_uint, C.c_uint, C.c_void_p]
        self.lib.ref_minimizer_positions.restype = None
        self.lib.ref_minimizer_positions(v.ctypes.data, None if hv is None else hv.ctypes.data, v.size, k, m, out.ctypes.data)
        return out

    def hash_sampler(self, values, seed, rate):
        v = np.ascontiguousarray(values, dtype=np.uint64)
        kept = np.zeros(v.size, np.uint64)
        thr = C.c_uint64()
        self.lib.ref_hash_sampler_u64.restype = C.c_size_t
        self.lib.ref_hash_sampler_u64.argtypes = [C.c_void_p, C.c_size_t, C.c_uint64, C.c_double, C.c_void_p, C.POINTER(C.c_uint64)]
        n = self.lib.ref_hash_sampler_u64(v.ctypes.data, v.size, seed & (2**64 - 1), float(rate), kept.ctypes.data, C.byref(thr))
        return kept[:n], int(thr.value)

    def syncmer_stream(self, seq, k, m, soff, eoff, canonical):
        b = _buf(seq)
        cap = len(b) + 1
        lo = np.zeros(cap, np.uint64); pos = np.zeros(cap, np.uint64); idv = np.zeros(cap, np.uint64)
        self.lib.ref_syncmers_u64.restype = C.c_size_t
        self.lib.ref_syncmers_u64.argtypes = [C.c_char_p, C.c_size_t, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_int,
                                              C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
        n = self.lib.ref_syncmers_u64(bytes(b), len(b), k, m, soff, eoff, int(canonical), lo.ctypes.data, pos.ctypes.data, idv.ctypes.data, cap)
        return dict(lo=lo[:n], pos=pos[:n], id=idv[:n])

    def kmer_stream(self, seq, k, canonical, bits=64):
        b = _buf(seq)
        cap = len(b) + 1
        has = np.zeros(cap, np.uint8); pos = np.zeros(cap, np.uint64); idv = np.zeros(cap, np.uint64)
        if bits == 64:
            lo = np.zeros(cap, np.uint64)
            n = self.lib.ref_kmer_view_u64(b, len(b), k, int(canonical), lo.ctypes.data, has.ctypes.data, pos.ctypes.data, idv.ctypes.data, cap)
            hi = np.zeros(cap, np.uint64)
        else:
            pair = np.zeros(2 * cap, np.uint64)
            n = self.lib.ref_kmer_view_u128(b, len(b), k, int(canonical), pair.ctypes.data, has.ctypes.data, pos.ctypes.data, idv.ctypes.data, cap)
            lo, hi = pair[0::2].copy(), pair[1::2].copy()
        return dict(lo=lo[:n], hi=hi[:n], has=has[:n], pos=pos[:n], id=idv[:n])

    def mmer_stream(self, seq, k, m, seed, canonical):
        b = _buf(seq)
        n = len(b)
        val = np.zeros(n, np.uint64); idx = np.zeros(n, np.uint64); h = np.zeros(n, np.uint64)
        bslb = np.zeros(n, np.uint64); ok = np.zeros(n, np.uint8)
        self.lib.ref_mmer_stream(b, n, k, m, seed, int(canonical), val.ctypes.data, idx.ctypes.data, h.ctypes.data, bslb.ctypes.data, ok.ctypes.data)
        return dict(value=val, idx=idx, hash=h, bslb=bslb, valid=ok)

    def super_kmers_sums(self, seq, k, m, seed, canonical):
        """count + wrapping sums (value, mm_start, size, hash): reference arithmetic + restated windowing."""
        b = seq if isinstance(seq, np.ndarray) else np.frombuffer(_buf(seq), np.uint8)
        sums = np.zeros(4, np.uint64)
        n = self.lib.ref_super_kmers_sums(b.ctypes.data_as(C.c_char_p), b.size, k, m, seed & (2**64 - 1), int(canonical), sums.ctypes.data)
        return int(n), sums

    def minimizer_view_literal_count(self, seq, k, m, seed, canonical):
        b = _buf(seq)
        return int(self.lib.ref_minimizer_view_literal_count(b, len(b), k, m, seed, int(canonical)))

    def kmer_values(self, seq, k, canonical, offsets=None):
        """What tests/test_jaccard.cpp:61-66 pushes into its vector: the present values of kmer_view, sequence by sequence."""
        b = np.frombuffer(bytes(_buf(seq)), np.uint8)
        offs = [0, b.size] if offsets is None else [int(x) for x in offsets]
        parts = []
        for a, e in zip(offs[:-1], offs[1:]):
            r = self.kmer_stream(b[a:e], k, canonical)
            parts.append(r["lo"][r["has"] == 1])
        return np.concatenate(parts) if parts else np.zeros(0, np.uint64)

    def ordered_unique(self, sorted_values):
        v = np.ascontiguousarray(sorted_values, np.uint64)
        out = np.empty(max(v.size, 1), np.uint64)
        n = self.lib.ref_ordered_unique(v.ctypes.data, v.size, out.ctypes.data)
        return out[:n].copy()

    def jaccard_sorted(self, a, b):
        """ordered_unique_sampler + algorithm::jaccard over two sorted lists: (intersection, union, size1, size2)."""
        a, b = np.ascontiguousarray(a, np.uint64), np.ascontiguousarray(b, np.uint64)
        out = np.zeros(4, np.uint64)
        self.lib.ref_jaccard_sorted(a.ctypes.data, a.size, b.ctypes.data, b.size, out.ctypes.data)
        return tuple(int(x) for x in out)

    def packed_vector_words(self, values, bitwidth=2):
        """bit::packed::vector<uint64_t>(bitwidth) after push_back of every value: its underlying words."""
        v = np.ascontiguousarray(values, np.uint8)
        out = np.zeros(v.size * bitwidth // 64 + 2, np.uint64)
        nw = self.lib.ref_packed_vector_words(v.ctypes.data, v.size, bitwidth, out.ctypes.data, out.size)
        return out[:nw].copy()

    def kseq_parse(self, text):
        """The reference's kseq.h executed over the text: (bases, offsets)."""
        b = bytes(_buf(text))
        bases = np.empty(len(b) + 1, np.uint8)
        offs = np.empty(len(b) // 2 + 2, np.uint64)
        n_seq = self.lib.ref_kseq_parse(b, len(b), bases.ctypes.data, offs.ctypes.data)
        return bases[:int(offs[n_seq])].copy(), offs[:n_seq + 1].copy()

    def generate(self, seed, length):
        out = np.empty(length, np.uint8)
        self.lib.ref_generate(seed, length, out.ctypes.data)
        return out

    def mutate(self, seq, mutp, indelf, extp, seed):
        b = _buf(seq)
        cap = len(b) * 4 + 64
        out = np.empty(cap, np.uint8)
        counts = np.zeros(3, np.uint64); tot = C.c_int64(0)
        n = self.lib.ref_mutate(b, len(b), mutp, indelf, extp, seed, out.ctypes.data, cap, counts.ctypes.data, C.byref(tot))
        if n > cap:  # long insertions: the call reports the size it needs
            cap = n
            out = np.empty(cap, np.uint8)
            n = self.lib.ref_mutate(b, len(b), mutp, indelf, extp, seed, out.ctypes.data, cap, counts.ctypes.data, C.byref(tot))
        return out[:n].copy(), counts, tot.value
