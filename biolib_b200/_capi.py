"""ctypes binding of the C ABI in include/biolib_gpu.h (biolib_b200/libbiolib_gpu.so).

This is the same thin binding a maintainer of the reference would write for any host language: plain
pointers and sizes. There is no fallback of any kind here -- if the shared library is missing, or no
sm_100 device is usable, calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libbiolib_gpu.so")
CSRC = os.path.join(HERE, "csrc")

BLG_OK, BLG_E_INVALID, BLG_E_UNSUPPORTED, BLG_E_CAPACITY, BLG_E_CUDA, BLG_E_NOMEM = range(6)
F_DEVICE_INPUT, F_DEVICE_OUTPUT, F_EMIT_HASH, F_NO_SIZE, F_EMIT_ID, F_POS32 = 0x1, 0x2, 0x4, 0x8, 0x10, 0x20
F_REF_DROP_LAST = 0x40
F_ASYNC = 0x80
F_RUNS = 0x100
F_TIGHT = 0x200

EXPORTS = [
    "blg_abi_version", "blg_strerror", "blg_create", "blg_destroy", "blg_set_stream", "blg_set_pipeline", "blg_last_error",
    "blg_last_stats", "blg_async_count", "blg_device_info", "blg_host_alloc", "blg_host_free", "blg_kmers", "blg_minimizers",
    "blg_super_kmers", "blg_super_kmers_range", "blg_hash64", "blg_minimizer_positions", "blg_syncmers", "blg_hash_sampled_kmers", "blg_free_kmer_out", "blg_free_sk_out",
    "blg_synth_generate", "blg_synth_mutate", "blg_pack_bases", "blg_super_kmers_packed", "blg_parse_fastx", "blg_free_fastx_out", "blg_super_kmer_payloads", "blg_free_payload_out", "blg_sorted_unique_kmers", "blg_kmer_jaccard",
    "blg_sort_unique_u64", "blg_sorted_sets_jaccard", "blg_lower_bound_u64", "blg_partition_by_top_byte", "blg_kmers_static_scratch",
]


class Params(C.Structure):
    _fields_ = [("k", C.c_uint8), ("m", C.c_uint8), ("canonical", C.c_uint8), ("reserved", C.c_uint8),
                ("flags", C.c_uint32), ("seed", C.c_uint64)]


class KmerOut(C.Structure):
    _fields_ = [("count", C.c_uint64), ("capacity", C.c_uint64), ("value_lo", C.c_void_p), ("value_hi", C.c_void_p),
                ("position", C.c_void_p), ("owner", C.c_void_p),
                ("n_runs", C.c_uint64), ("run_capacity", C.c_uint64), ("run_first_record", C.c_void_p), ("run_start", C.c_void_p)]


class SkOut(C.Structure):
    _fields_ = [("count", C.c_uint64), ("capacity", C.c_uint64), ("value", C.c_void_p), ("mm_start", C.c_void_p),
                ("size", C.c_void_p), ("mm_pos", C.c_void_p), ("hash", C.c_void_p), ("id", C.c_void_p), ("owner", C.c_void_p),
                ("n_runs", C.c_uint64), ("run_capacity", C.c_uint64), ("run_first_record", C.c_void_p), ("run_start", C.c_void_p),
                ("value_lo32", C.c_void_p), ("value_hi8", C.c_void_p), ("size_pos", C.c_void_p)]


class PackedSeq(C.Structure):
    _fields_ = [("codes", C.c_void_p), ("n_bases", C.c_size_t), ("break_start", C.c_void_p), ("break_len", C.c_void_p), ("n_breaks", C.c_size_t)]


class FastxOut(C.Structure):
    _fields_ = [("n_bases", C.c_uint64), ("n_seq", C.c_uint64), ("bases", C.c_void_p), ("seq_offsets", C.c_void_p), ("owner", C.c_void_p)]


class PayloadOut(C.Structure):
    _fields_ = [("count", C.c_uint64), ("n_buckets", C.c_uint32), ("total_bases", C.c_uint64), ("n_words", C.c_size_t),
                ("bucket_offsets", C.c_void_p), ("value", C.c_void_p), ("mm_start", C.c_void_p), ("size", C.c_void_p),
                ("mm_pos", C.c_void_p), ("payload_start", C.c_void_p), ("words", C.c_void_p), ("owner", C.c_void_p)]


class JaccardOut(C.Structure):
    _fields_ = [("intersection", C.c_uint64), ("union_size", C.c_uint64), ("size1", C.c_uint64), ("size2", C.c_uint64)]


class Stats(C.Structure):
    _fields_ = [("n_bases", C.c_uint64), ("n_records", C.c_uint64), ("bytes_h2d", C.c_uint64), ("bytes_d2h", C.c_uint64),
                ("kernel_launches", C.c_uint64), ("kernel_ms", C.c_double), ("total_ms", C.c_double), ("exact_tiles", C.c_uint64)]


def build(force: bool = False) -> str:
    """Compile the CUDA sources for sm_100a into libbiolib_gpu.so (nvcc cross-compiles without a GPU)."""
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h", ".cpp"))]
    srcs.append(os.path.join(HERE, "..", "include", "biolib_gpu.h"))
    stale = force or not os.path.exists(LIB_PATH) or any(
        os.path.exists(s) and os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in srcs)
    if stale:
        subprocess.check_call(["make", "-C", CSRC, "-j4"], stdout=subprocess.DEVNULL)
    return LIB_PATH


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("BLG_LIB_PATH", LIB_PATH)   # development A/B of two builds of the same library
    if not os.path.exists(path):
        raise RuntimeError(f"{path} is missing: run __graft_entry__.build() (there is no CPU fallback)")
    L = C.CDLL(path)
    vp, u64, sz = C.c_void_p, C.c_uint64, C.c_size_t
    L.blg_abi_version.restype = C.c_int
    L.blg_strerror.restype = C.c_char_p
    L.blg_strerror.argtypes = [C.c_int]
    L.blg_create.argtypes = [C.c_int, C.POINTER(vp)]
    L.blg_destroy.argtypes = [vp]
    L.blg_set_stream.argtypes = [vp, vp]
    L.blg_set_pipeline.argtypes = [vp, C.c_int, C.c_int]
    L.blg_last_error.restype = C.c_char_p
    L.blg_last_error.argtypes = [vp]
    L.blg_last_stats.argtypes = [vp, C.POINTER(Stats)]
    L.blg_async_count.argtypes = [vp, C.POINTER(u64)]
    L.blg_device_info.argtypes = [vp, C.POINTER(C.c_int), C.c_char_p, sz]
    L.blg_minimizer_positions.argtypes = [vp, vp, sz, C.c_uint8, C.c_uint8, C.c_uint32, vp]
    L.blg_host_alloc.argtypes = [vp, sz, C.POINTER(vp)]
    L.blg_host_free.argtypes = [vp, vp]
    L.blg_kmers.argtypes = [vp, vp, sz, vp, sz, C.POINTER(Params), C.POINTER(KmerOut)]
    L.blg_syncmers.argtypes = [vp, vp, sz, vp, sz, C.POINTER(Params), C.c_uint16, C.c_uint16, C.POINTER(KmerOut)]
    L.blg_hash_sampled_kmers.argtypes = [vp, vp, sz, vp, sz, C.POINTER(Params), u64, C.POINTER(KmerOut)]
    for f in (L.blg_minimizers, L.blg_super_kmers):
        f.argtypes = [vp, vp, sz, vp, sz, C.POINTER(Params), C.POINTER(SkOut)]
    L.blg_super_kmers_range.argtypes = [vp, vp, sz, vp, sz, u64, u64, C.POINTER(Params), C.POINTER(SkOut)]
    L.blg_hash64.argtypes = [vp, vp, sz, u64, C.c_uint32, vp]
    L.blg_free_kmer_out.argtypes = [vp, C.POINTER(KmerOut)]
    L.blg_free_kmer_out.restype = None
    L.blg_free_sk_out.argtypes = [vp, C.POINTER(SkOut)]
    L.blg_free_sk_out.restype = None
    L.blg_pack_bases.argtypes = [vp, sz, vp, vp, vp, sz, C.POINTER(sz)]
    L.blg_super_kmers_packed.argtypes = [vp, C.POINTER(PackedSeq), vp, sz, u64, u64, C.POINTER(Params), C.POINTER(SkOut)]
    L.blg_parse_fastx.argtypes = [vp, vp, sz, C.c_int, C.c_uint32, C.POINTER(FastxOut)]
    L.blg_free_fastx_out.argtypes = [vp, C.POINTER(FastxOut)]
    L.blg_free_fastx_out.restype = None
    L.blg_super_kmer_payloads.argtypes = [vp, vp, sz, vp, sz, C.POINTER(Params), C.c_uint32, C.POINTER(PayloadOut)]
    L.blg_free_payload_out.argtypes = [vp, C.POINTER(PayloadOut)]
    L.blg_free_payload_out.restype = None
    L.blg_sorted_unique_kmers.argtypes = [vp, vp, sz, vp, sz, C.POINTER(Params), C.POINTER(KmerOut)]
    L.blg_kmer_jaccard.argtypes = [vp, vp, sz, vp, sz, vp, sz, vp, sz, C.POINTER(Params), C.POINTER(JaccardOut)]
    L.blg_sort_unique_u64.argtypes = [vp, vp, sz, C.c_int, C.POINTER(u64)]
    L.blg_sorted_sets_jaccard.argtypes = [vp, vp, sz, vp, sz, C.c_int, C.POINTER(JaccardOut)]
    L.blg_lower_bound_u64.argtypes = [vp, vp, sz, vp, sz, vp]
    L.blg_partition_by_top_byte.argtypes = [vp, vp, sz, C.c_int, vp, vp]
    L.blg_kmers_static_scratch.argtypes = [vp, sz, C.c_uint8, C.c_int, vp, vp, vp, u64, vp, vp, C.POINTER(u64), C.c_char_p, sz]
    L.blg_synth_generate.argtypes = [u64, sz, vp]
    L.blg_synth_mutate.argtypes = [vp, sz, C.c_double, C.c_double, C.c_double, u64, vp, sz, C.POINTER(sz), vp, vp]
    _lib = L
    return L
