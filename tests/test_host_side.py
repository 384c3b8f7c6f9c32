"""CPU: host-side pieces of the product that can be checked without a GPU.

* the SWAR ASCII->2-bit packer compiled for the device (biolib_b200/csrc/nt4_swar.h) is plain C: it is compiled
  here with gcc and checked exhaustively against the reference's table (constants.hpp:12-21 via the oracle);
* the C-ABI library loads and exports every symbol include/biolib_gpu.h declares; struct layouts match the
  ctypes mirror; compute entry points fail loudly (BLG_E_CUDA) without a device -- no CPU fallback;
* the C++ shim include/biolib_gpu.hpp compiles as plain C++17 against the header;
* the synthetic-input generators and the multi-GPU shard plan.
"""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="session")
def built():
    sys.path.insert(0, ROOT)
    import __graft_entry__ as g
    g.build()
    from biolib_b200 import _capi
    return _capi


def test_nt4_swar_exhaustive(tmp_path, oracle):
    src = tmp_path / "swar.c"
    src.write_text(
        '#include <stdio.h>\n#include "nt4_swar.h"\n'
        "int main(void){ for (unsigned b = 0; b < 256; ++b) for (int lane = 0; lane < 4; ++lane) {\n"
        "  /* the byte under test in `lane`, the other three lanes filled with every class of neighbour */\n"
        "  static const unsigned char fill[6] = {'A','t',0x00,0xFF,'N','G'};\n"
        "  for (int f = 0; f < 6; ++f) { unsigned x = 0; for (int q = 0; q < 4; ++q) x |= (unsigned)(q == lane ? b : fill[(f + q) % 6]) << (8*q);\n"
        "    unsigned r = blg_pack4(x); printf(\"%u %d %u %u\\n\", b, lane, (r >> (2*lane)) & 3u, (r >> (8 + 2*lane)) & 3u);\n"
        "    /* the split forms of the fused kernels' fast path must say the same */\n"
        "    unsigned fb = ((blg_bad4(x) >> (8*lane)) & 0xFFu) != 0, fc = (blg_codes4(x) >> (2*lane)) & 3u;\n"
        "    if (fb != ((r >> (8 + 2*lane)) & 1u) || (!fb && fc != ((r >> (2*lane)) & 3u))) { printf(\"SPLIT MISMATCH %u %d\\n\", b, lane); return 1; } } }\n"
        "  unsigned c, fl; blg_pack16(0x54474341u /*ACGT*/, 0x7567634Eu /*Ncgu*/, 0x41414141u, 0x0A3E4155u /*UA>\\n*/, &c, &fl);\n"
        "  printf(\"P16 %u %u\\n\", c, fl); return 0; }\n")
    exe = tmp_path / "swar"
    subprocess.check_call(["gcc", "-O2", "-std=c11", "-I", os.path.join(ROOT, "biolib_b200", "csrc"), str(src), "-o", str(exe)])
    out = subprocess.check_output([str(exe)]).decode().split("\n")
    table = oracle.nt4_table()
    n = 0
    for line in out:
        if not line or line.startswith("P16"):
            continue
        b, lane, code, bad = map(int, line.split())
        want = int(table[b])
        assert bad in (0, 1)
        assert bad == (1 if want == 4 else 0), (b, lane)
        if want < 4:
            assert code == want, (b, lane)
        n += 1
    assert n == 256 * 4 * 6
    p16 = [l for l in out if l.startswith("P16")][0].split()
    codes, flags = int(p16[1]), int(p16[2])
    seq = b"ACGTNcguAAAAUA>\n"
    for q, ch in enumerate(seq):
        want = int(table[ch])
        assert ((flags >> (2 * q)) & 3) == (1 if want == 4 else 0)
        if want < 4:
            assert ((codes >> (2 * q)) & 3) == want


def test_capi_exports_every_declared_symbol(built):
    hdr = open(os.path.join(ROOT, "include", "biolib_gpu.h")).read()
    body = hdr[hdr.index('extern "C"'):]
    declared = set(re.findall(r"\b(blg_[a-z0-9_]+)\s*\(", body))
    assert declared, "no declarations parsed"
    assert declared == set(built.EXPORTS), declared ^ set(built.EXPORTS)
    lib = C.CDLL(built.LIB_PATH)
    for name in declared:
        assert getattr(lib, name) is not None
    assert built.lib().blg_abi_version() >= 1
    for st in range(6):
        assert built.lib().blg_strerror(st)


def test_struct_layouts_match_header(built, tmp_path):
    src = tmp_path / "lay.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "biolib_gpu.h"\n'
                   'int main(void){ printf("%zu %zu %zu %zu %zu %zu %zu\\n", sizeof(blg_params), offsetof(blg_params, flags), offsetof(blg_params, seed),'
                   ' sizeof(blg_kmer_out), sizeof(blg_sk_out), sizeof(blg_stats), offsetof(blg_sk_out, hash)); return 0; }\n')
    exe = tmp_path / "lay"
    subprocess.check_call(["gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = list(map(int, subprocess.check_output([str(exe)]).split()))
    want = [C.sizeof(built.Params), built.Params.flags.offset, built.Params.seed.offset, C.sizeof(built.KmerOut),
            C.sizeof(built.SkOut), C.sizeof(built.Stats), built.SkOut.hash.offset]
    assert got == want


@pytest.mark.skipif(_have_gpu(), reason="only meaningful on a box without a GPU")
def test_no_cpu_fallback_without_device(built):
    import biolib_b200
    with pytest.raises(biolib_b200.BlgError) as e:
        biolib_b200.Context(0)
    assert e.value.status == built.BLG_E_CUDA


def test_cpp_shim_compiles(tmp_path):
    hpp = os.path.join(ROOT, "include", "biolib_gpu.hpp")
    if not os.path.exists(hpp):
        pytest.skip("shim not present")
    src = tmp_path / "use.cpp"
    src.write_text('#include "biolib_gpu.hpp"\n#include <string>\n'
                   "int main(){ std::string s(\"ACGTACGTACGTACGTACGT\");\n"
                   " try { auto v = biolib::gpu::kmer_view_from_string<uint64_t>(s, 5, true); for (auto it = v.cbegin(); it != v.cend(); ++it) { auto r = *it; (void)r; }\n"
                   "       auto mv = biolib::gpu::minimizer_view_from_string<uint64_t, uint64_t, hash::hash64>(s, 7, 4, 42, false); for (auto const& r : mv) (void)r;\n"
                   "       biolib::gpu::super_kmer_view<uint64_t, uint64_t, hash::hash64> sk(s, 7, 4, false); for (auto const& r : sk) (void)r; }\n"
                   " catch (std::exception const&) { return 3; } return 0; }\n")
    obj = tmp_path / "use.o"
    subprocess.check_call(["g++", "-std=c++17", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"), "-c", str(src), "-o", str(obj)])


def test_synth_matches_reference_generator(oracle):
    from biolib_b200 import synth
    for seed in (0, 42, 43, 2**32 + 42):
        assert np.array_equal(synth.generate(seed, 100_000, chunk=30_000), oracle.generate(seed, 100_000))
    g = synth.generate_sharded(42, 2500, 1000)
    assert np.array_equal(g[:1000], oracle.generate(42, 1000)) and np.array_equal(g[1000:2000], oracle.generate(43, 1000))
    assert np.array_equal(g[2000:], oracle.generate(44, 500))
    reads = synth.generate(43, 150 * 4000)
    before = reads.copy()
    synth.inject_n_runs(reads, 150, seed=44)
    changed = np.nonzero(reads != before)[0]
    assert 0 < changed.size < reads.size * 0.05
    assert set(np.unique(reads[changed]).tolist()) <= set(b"Nnacgt")  # lower-cased reads turn N into n (still a break)
    again = synth.inject_n_runs(before.copy(), 150, seed=44)
    assert np.array_equal(again, reads)


def test_synth_mutator_matches_reference(oracle):
    """biolib_b200.synth.mutate (blg_synth_mutate, csrc/synth_host.cpp) against the reference's sequence_mutator executed
    verbatim when oracle/_ref is here, and against the oracle restatement pinned to it (tests/test_oracle_vs_ref.py)."""
    import ctypes as C
    import pyoracle
    from biolib_b200 import synth, _capi
    checkers = [oracle] + ([pyoracle.Ref()] if pyoracle.Ref.available() else [])
    src = synth.generate(5, 6000)
    out = np.empty(100, np.uint8)
    assert _capi.lib().blg_synth_generate(2**32 + 42, 100, out.ctypes.data) == 0 and np.array_equal(out, oracle.generate(42, 100))
    for seed, (mp, idf, ep) in enumerate(((0.1, 0.3, 0.3), (0.5, 0.5, 0.5), (1.0, 1.0, 0.9), (0.02, 0.0, 0.0), (0.0, 0.5, 0.5))):
        got, rep = synth.mutate(src, mp, idf, ep, seed, with_report=True)
        for chk in checkers:
            b, cb, tb = chk.mutate(src, mp, idf, ep, seed)
            assert np.array_equal(got, b), (seed, type(chk).__name__)
            assert [rep["substitutions"], rep["insertions"], rep["deletions"]] == [int(x) for x in cb] and rep["total_indel_len"] == tb
    assert synth.mutate(np.zeros(0, np.uint8), 0.1, 0.3, 0.3, 1).size == 0
    assert synth.mutate("A", 0.1, 0.3, 0.3, 1).size == 0            # the last source base is never emitted
    with pytest.raises(ValueError):
        synth.mutate("ACGNNACGT" * 50, 0.9, 0.0, 0.3, 1)            # the reference indexes bases.at(4) here
    bases, offs = synth.mutated_long_reads(6, lo=500, hi=900)
    assert offs[0] == 0 and offs[-1] == bases.size and len(offs) == 7 and set(np.unique(bases).tolist()) <= set(b"ACGT")
    b2, o2 = synth.mutated_long_reads(6, lo=500, hi=900, threads=1)
    assert np.array_equal(bases, b2) and np.array_equal(offs, o2)


def test_hash_sampler_threshold_rate_one_is_clamped():
    """rate 1.0 is outside the reference's defined domain (uint64 cast of 2^64); we keep everything."""
    from biolib_b200.context import Context
    assert Context.hash_sampler_threshold(1.0) == 2**64 - 1
    assert Context.hash_sampler_threshold(0.0) == 0
    assert Context.hash_sampler_threshold(0.5) == int(0.5 * float(2**64 - 1))
    with pytest.raises(ValueError):
        Context.hash_sampler_threshold(1.5)


def test_run_list_and_tight_records_unpack(oracle):
    """Host side of BLG_F_RUNS / BLG_F_TIGHT: positions rebuilt from the run list and values / sizes rebuilt from the 6-byte records
    equal the oracle's records (the packing is restated here the way include/biolib_gpu.h documents it)."""
    from biolib_b200.context import expand_runs, unpack_tight
    rng = np.random.default_rng(11)
    s = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=30_000)
    s[5000:5003] = ord("N"); s[20_000] = ord("N")
    o = oracle.super_kmers(s, 31, 19, 42, True)
    n = o["n"]
    value, sk_start, size, mm_start = o["value"][:n], o["sk_start"][:n], o["size"][:n].astype(np.uint8), o["mm_start"][:n]
    mm_pos = (mm_start - sk_start).astype(np.uint8)
    # run list: a record starts a run iff it does not start where its predecessor ended
    first = np.flatnonzero(np.concatenate([[True], sk_start[1:] != sk_start[:-1] + size[:-1]])).astype(np.uint64)
    assert first.size == 3
    assert np.array_equal(expand_runs(first, sk_start[first.astype(np.int64)], size, mm_pos), mm_start)
    lo32, hi8, nib = (value & np.uint64(0xFFFFFFFF)).astype(np.uint32), (value >> np.uint64(32)).astype(np.uint8), ((size - 1) | (mm_pos << 4)).astype(np.uint8)
    v2, s2, p2 = unpack_tight(lo32, hi8, nib)
    assert np.array_equal(v2, value) and np.array_equal(s2, size) and np.array_equal(p2, mm_pos)
    v3, _, _ = unpack_tight(lo32, None, nib)
    assert np.array_equal(v3, value & np.uint64(0xFFFFFFFF))


def test_flag_values_match_header(built, tmp_path):
    """The Python mirror's BLG_F_* values and the ABI version are the header's."""
    names = ["DEVICE_INPUT", "DEVICE_OUTPUT", "EMIT_HASH", "NO_SIZE", "EMIT_ID", "POS32", "REF_DROP_LAST", "ASYNC", "RUNS", "TIGHT"]
    src = tmp_path / "flags.c"
    src.write_text('#include <stdio.h>\n#include "biolib_gpu.h"\nint main(void){ printf("%d' + " %u" * len(names) + '\\n", BLG_ABI_VERSION, '
                   + ", ".join("BLG_F_" + n for n in names) + "); return 0; }\n")
    exe = tmp_path / "flags"
    subprocess.check_call(["gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = list(map(int, subprocess.check_output([str(exe)]).split()))
    assert got[0] == built.lib().blg_abi_version() == 3
    assert got[1:] == [getattr(built, "F_" + n) for n in names]


def test_kmer_run_list_expands_to_positions(oracle):
    """Host side of blg_kmers + BLG_F_RUNS: positions rebuilt from the run list equal the oracle's k-mer positions."""
    from biolib_b200.context import expand_kmer_runs
    rng = np.random.default_rng(12)
    s = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=20_000)
    s[3000:3004] = ord("N"); s[9000] = ord("n"); s[9050] = ord("N")
    pos = oracle.kmers(s, 31, True)["pos"]
    first = np.flatnonzero(np.concatenate([[True], np.diff(pos.astype(np.int64)) != 1])).astype(np.uint64)
    assert first.size == 4
    assert np.array_equal(expand_kmer_runs(first, pos[first.astype(np.int64)], pos.size), pos)
    assert expand_kmer_runs(np.zeros(0, np.uint64), np.zeros(0, np.uint64), 0).size == 0
