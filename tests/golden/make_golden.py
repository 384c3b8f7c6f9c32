#!/usr/bin/env python
"""Generates tests/golden/*.json from the REFERENCE'S OWN CODE executed in this container.

Run here (needs /root/reference; the GPU box never runs this):  python tests/golden/make_golden.py

Everything below calls oracle/_ref/libbiolib_ref.so = oracle/ref_tier_a.cpp, a TU with no algorithm of its
own that #includes the reference headers where they lie and drives
  hash::hash64 / double_hash64 / remix           include/hash.hpp:11-85, bundled/MurmurHash3.cpp:263-341
  wrapper::kmer_view iteration                    include/kmer_view.hpp:155-234
  minimizer_view::const_iterator::read_base()     include/minimizer_view.hpp:229-245
  sequence::generator, wrapper::sequence_mutator  include/sequence_generator.hpp, include/sequence_mutator.hpp
The super-k-mer records in `superkmers.json` are NOT reference output (the reference's iterators for that part
are non-functional): they are the brute-force D1-D9 restatement evaluated over the reference-
executed per-base (value, hash) stream stored beside them, and the file says so in its "provenance" key.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import pyoracle  # noqa: E402

ref = pyoracle.Ref()


def dump(name, obj):
    with open(os.path.join(HERE, name), "w") as f:
        json.dump(obj, f, indent=None, separators=(",", ":"))
        f.write("\n")
    print(name, os.path.getsize(os.path.join(HERE, name)), "bytes")


def rand_seq(rng, n, p_n=0.0, lower=0.0):
    s = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=n).copy()
    if p_n:
        s[rng.random(n) < p_n] = ord("N")
    if lower:
        mk = rng.random(n) < lower
        s[mk] |= 0x20
    return s.tobytes().decode()


# ---------------------------------------------------------------------------------------------- nt4 + hash
keys = [0, 1, 0x1234567890ABCDEF, 0xFFFFFFFFFFFFFFFF, 107, 0x5555555555555555]
seeds = [0, 42, 0x10000002A, 0xFFFFFFFF, 0xDEADBEEFCAFE]
hash_kat = {
    "provenance": "reference hash::hash64::hash / double_hash64::hash / remix executed verbatim (oracle/ref_tier_a.cpp)",
    "nt4_table": ref.nt4_table().tolist(),
    "hash64": [[k, s, ref.hash64(k, s)] for k in keys for s in seeds],
    "double_hash64_u64": [[k, s, list(ref.double_hash64(k, s))] for k in keys[:4] for s in (0, 42)],
    "double_hash64_u128": [[lo, hi, s, list(ref.double_hash64_u128(lo, hi, s))]
                           for lo, hi in ((0xFEDCBA9876543210, 0x0123456789ABCDEF), (0, 0), (1, 0), (0, 1)) for s in (0, 42)],
    "remix": [[z, ref.remix(z)] for z in (0, 1, 42, 0xFFFFFFFFFFFFFFFF)],
}
dump("hash_kat.json", hash_kat)

# ---------------------------------------------------------------------------------------------- kmer_view
rng = np.random.default_rng(2024)
kv_cases = []
LONG = "ACGGTCATTGACCGTAGGATCCAGTTACGGATTTGACACCAGTAGGACTAGG"
fixed = [("ACGTACGT", 3), ("ACGT", 3), ("ACG", 3), ("ACGTN", 3), ("ACGTNNNACGT", 3), ("ACGTNA", 3), ("ACNGT", 3),
         ("NNN", 3), ("AC", 3), ("ACGTNN", 3), ("acgtUuNACGTAC", 4), (LONG, 31), (LONG, 15), (LONG, 32), (LONG, 1)]
for s, k in fixed:
    for canon in (0, 1):
        r = ref.kmer_stream(s, k, canon)
        kv_cases.append(dict(seq=s, k=k, canonical=canon, bits=64, value=r["lo"].tolist(), has=r["has"].tolist(),
                             pos=r["pos"].tolist(), id=r["id"].tolist()))
for n, k, pn in ((200, 7, 0.03), (300, 31, 0.01), (300, 16, 0.02), (150, 21, 0.0), (400, 32, 0.01)):
    s = rand_seq(rng, n, pn, 0.05)
    for canon in (0, 1):
        r = ref.kmer_stream(s, k, canon)
        kv_cases.append(dict(seq=s, k=k, canonical=canon, bits=64, value=r["lo"].tolist(), has=r["has"].tolist(),
                             pos=r["pos"].tolist(), id=r["id"].tolist()))
# 128-bit words: the reference is correct only non-canonical (K7)
for n, k in ((200, 33), (220, 63), (200, 64), (180, 47)):
    s = rand_seq(rng, n, 0.01)
    r = ref.kmer_stream(s, k, 0, bits=128)
    kv_cases.append(dict(seq=s, k=k, canonical=0, bits=128, value=r["lo"].tolist(), value_hi=r["hi"].tolist(),
                         has=r["has"].tolist(), pos=r["pos"].tolist(), id=r["id"].tolist()))
g = ref.generate(42, 1_000_000)
r = ref.kmer_stream(g, 31, 1)
assert int(r["has"].min()) == 1
kv_1m = dict(n=int(len(r["lo"])), xor=int(np.bitwise_xor.reduce(r["lo"])), sum=int(r["lo"].sum(dtype=np.uint64)),
             first=int(r["lo"][0]), last=int(r["lo"][-1]), first_pos=int(r["pos"][0]), last_pos=int(r["pos"][-1]))
dump("kmer_view.json", {"provenance": "reference wrapper::kmer_view iterated verbatim (K5/K6/K8 guards in oracle/ref_tier_a.cpp)",
                        "cases": kv_cases, "config1_1mbp_seed42_k31_canonical": kv_1m})

# ---------------------------------------------------------------------------------------------- m-mer stream
mm_cases = []
WORKED = "ACGGTCATTGACNCGTAGGATCCAGTT"
specs = [(WORKED, 7, 4, 42), (WORKED, 7, 4, 0), (LONG, 31, 19, 42), (LONG, 21, 11, 0x10000002A)]
for n, k, m, pn in ((300, 31, 19, 0.01), (300, 21, 11, 0.02), (400, 63, 31, 0.005), (200, 15, 10, 0.03), (200, 9, 5, 0.05),
                    (300, 64, 32, 0.004), (200, 8, 4, 0.02), (200, 12, 6, 0.02)):
    specs.append((rand_seq(rng, n, pn, 0.05), k, m, 42))
specs.append(("AT" * 60 + "N" + "ACGT" * 20 + "TTTTTTTTAAAAAAAA" * 4, 12, 6, 1))   # even m palindromes (D3)
specs.append(("AT" * 60 + "N" + "ACGT" * 20 + "TTTTTTTTAAAAAAAA" * 4, 9, 4, 1))
for s, k, m, seed in specs:
    for canon in (0, 1):
        r = ref.mmer_stream(s, k, m, seed, canon)
        # per input byte i: the m-mer that ENDS at i (0-based start = idx - 1, D7)
        mm_cases.append(dict(seq=s, k=k, m=m, seed=seed, canonical=canon, valid=r["valid"].tolist(),
                             idx=r["idx"].tolist(), value=r["value"].tolist(), hash=r["hash"].tolist()))
dump("mmer_stream.json", {"provenance": "reference minimizer_view::const_iterator::read_base() + hash::hash64::hash executed "
                                        "verbatim per input byte (oracle/ref_tier_a.cpp ref_mmer_stream)", "cases": mm_cases})

# ---------------------------------------------------------------------------------------------- super-k-mers (Tier B over Tier A)
def brute_records(case):
    """D4-D8 by brute force over the reference-executed per-base stream of `case`."""
    s, k, m = case["seq"], case["k"], case["m"]
    n, w = len(s), case["k"] - case["m"] + 1
    H, V, ok = {}, {}, {}
    for i in range(n):
        if case["valid"][i]:
            j = case["idx"][i] - 1
            H[j], V[j], ok[j] = case["hash"][i], case["value"][i], True
    recs, prev = [], None
    for st in range(0, n - k + 1):
        if not all(ok.get(j, False) for j in range(st, st + w)) or any(pyo_nt4[ord(c)] > 3 for c in s[st:st + k]):
            prev = None
            continue
        arg = min(range(st, st + w), key=lambda j: (H[j], j))
        if prev is not None and prev == arg:
            recs[-1]["size"] += 1
        else:
            recs.append(dict(sk_start=st, size=1, mm_start=arg, mm_pos=arg - st, value=V[arg], hash=H[arg]))
        prev = arg
    return recs


pyo_nt4 = ref.nt4_table().tolist()
sk_cases = []
for c in mm_cases:
    recs = brute_records(c)
    sk_cases.append(dict(seq=c["seq"], k=c["k"], m=c["m"], seed=c["seed"], canonical=c["canonical"],
                         **{key: [r[key] for r in recs] for key in ("sk_start", "size", "mm_start", "mm_pos", "value", "hash")}))
dump("superkmers.json", {"provenance": "Tier B: brute-force D1-D9 (SURVEY 8(a)) in pure Python over the reference-executed "
                                       "per-base (value, hash) stream of mmer_stream.json; NOT reference iterator output "
                                       "(minimizer_view yields nothing, super_kmer_view does not compile)", "cases": sk_cases})

# ---------------------------------------------------------------------------------------------- generator / mutator
gen = {"provenance": "reference sequence::generator::nucleic<packed::uniform>::get_sequence and wrapper::sequence_mutator "
                     "(driven as tests/test_random_sequence_generation.cpp:44-48) executed verbatim",
       "generate": [[seed, 200, ref.generate(seed, 200).tobytes().decode()] for seed in (0, 1, 42, 43, 0x10000002A)],
       "generate_1m_seed42_sha": None, "mutate": []}
import hashlib  # noqa: E402
gen["generate_1m_seed42_sha"] = hashlib.sha256(g.tobytes()).hexdigest()
for seed, (mp, idf, ep) in ((7, (0.1, 0.3, 0.3)), (47, (0.10, 0.3, 0.3)), (3, (0.5, 0.5, 0.5)), (9, (0.0, 0.3, 0.3)), (5, (1.0, 0.0, 0.0))):
    src = ref.generate(100 + seed, 400).tobytes().decode()
    out, counts, tot = ref.mutate(src, mp, idf, ep, seed)
    gen["mutate"].append(dict(src=src, mutp=mp, indelf=idf, extp=ep, seed=seed, out=out.tobytes().decode(),
                              counts=counts.tolist(), total_indel_len=int(tot)))
dump("generator.json", gen)


def minpos(ref, path):
    """tests/golden/minpos.json: hash::minimizer_position_extractor (reference include/kmer_view.hpp:266-283) executed
    verbatim over seeded k-mer values, incl. all-A, all-T, (AC)n-like and 1."""
    import json
    import numpy as np
    rng = np.random.default_rng(20261104)
    cases = []
    for k, m in [(31, 19), (21, 11), (15, 10), (32, 32), (32, 1), (7, 4), (5, 5), (32, 16), (17, 16)]:
        v = rng.integers(0, 2**63, size=64, dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, size=64, dtype=np.uint64)
        full = (1 << (2 * k)) - 1 if k < 32 else 2**64 - 1
        if k < 32:
            v &= np.uint64(full)
        v[:4] = [0, full, 0x5555555555555555 & full, 1]
        pos = ref.minimizer_positions(v, k, m)
        cases.append(dict(k=k, m=m, values=[int(x) for x in v], positions=[int(x) for x in pos]))
    nullopt = int(ref.minimizer_positions(np.array([5], np.uint64), 7, 4, has_value=np.array([0], np.uint8))[0])
    json.dump(dict(source="hash::minimizer_position_extractor executed verbatim (oracle/_ref)", nullopt_k7_m4=nullopt, cases=cases),
              open(path, "w"))


def syncmers(ref, path):
    """tests/golden/syncmers.json: sampler::syncmer_sampler<kmer_view::const_iterator, hash::minimizer_position_extractor>
    (reference include/syncmer_sampler.hpp:130-137) executed verbatim over the reference's kmer_view.
    Run: python -c "import sys; sys.path.insert(0,'oracle'); import pyoracle, runpy; ..."  or call from here."""
    import json
    import numpy as np
    rng = np.random.default_rng(20261105)
    cases = []
    for k, m, n, pn in [(31, 19, 600, 0.0), (21, 11, 500, 0.01), (15, 10, 400, 0.02), (7, 4, 200, 0.05), (32, 16, 500, 0.0), (5, 5, 60, 0.0),
                        (9, 2, 300, 0.0)]:
        seq = rand_seq(rng, n, p_n=pn, lower=0.05)
        for canon in (0, 1):
            for soff, eoff in ((0, k - m), (1, 1)):
                r = ref.syncmer_stream(seq, k, m, soff, eoff, canon)
                cases.append(dict(seq=seq, k=k, m=m, soff=soff, eoff=eoff, canonical=canon, value=[int(x) for x in r["lo"]],
                                  position=[int(x) for x in r["pos"]], id=[int(x) for x in r["id"]]))
    json.dump(dict(provenance="reference sampler::syncmer_sampler over wrapper::kmer_view executed verbatim (oracle/ref_tier_a.cpp "
                              "ref_syncmers_u64)", cases=cases), open(path, "w"), separators=(",", ":"))


def hash_sampler(ref, path):
    """tests/golden/hash_sampler.json: sampler::hash_sampler<vector<uint64_t>::const_iterator, hash::hash64> (reference
    include/hash_sampler.hpp:73-78,138-141) executed verbatim: kept values and the constructor's threshold."""
    import json
    import numpy as np
    rng = np.random.default_rng(20261106)
    v = rng.integers(0, 2**62, size=400, dtype=np.uint64)
    cases = []
    for seed, rate in ((42, 0.1), (0, 0.5), (7, 0.01), (2**32 + 42, 0.25), (1, 0.0), (3, 0.999)):
        kept, thr = ref.hash_sampler(v, seed, rate)
        cases.append(dict(seed=seed, rate=rate, threshold=thr, kept=[int(x) for x in kept]))
    json.dump(dict(provenance="reference sampler::hash_sampler executed verbatim (oracle/ref_tier_a.cpp ref_hash_sampler_u64)",
                   values=[int(x) for x in v], cases=cases), open(path, "w"), separators=(",", ":"))


minpos(ref, os.path.join(HERE, "minpos.json"))
hash_sampler(ref, os.path.join(HERE, "hash_sampler.json"))
syncmers(ref, os.path.join(HERE, "syncmers.json"))
