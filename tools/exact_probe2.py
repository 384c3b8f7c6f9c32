import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import biolib_b200
ctx = biolib_b200.Context(0)
rng = np.random.default_rng(5)
s = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=40_000)
r = ctx.super_kmers(s, 63, 31, 42, True, want_hash=False)
print("records", r["n"], "exact_tiles", ctx.stats()["exact_tiles"], flush=True)
ctx.close()
