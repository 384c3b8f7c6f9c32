import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import biolib_b200
ctx = biolib_b200.Context(0)
rng = np.random.default_rng(5)
s = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=400_000)
for k, m in ((63, 31), (47, 15), (33, 1), (56, 24), (49, 17), (48, 16), (31, 19), (21, 11), (43, 31), (41, 31)):
    for canon in (False, True):
        r = ctx.super_kmers(s, k, m, 42, canon, want_hash=False)
        print(k, m, canon, "records", r["n"], "exact_tiles", ctx.stats()["exact_tiles"], flush=True)
ctx.close()
