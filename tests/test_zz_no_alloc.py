"""blg_kmers_static_scratch: the k-mer kernel with its scratch in a static device array (no allocation inside the call).

Collected last on purpose. The new entry point is called through the C ABI, twice in a row on
different inputs (the second call finds the first call's scratch contents), and must give exactly the records of the existing
entry point (blg_kmers with device input / output): integer work, so the tolerance is bit equality, as everywhere in tests/.
"""
import ctypes as C

import numpy as np
import pytest



def rand_seq(rng, n, p_n=0.0, lower=0.0):
    s = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=n)
    if p_n > 0:
        s = s.copy()
        s[rng.random(n) < p_n] = ord("N")
    if lower > 0:
        s = s.copy()
        m = rng.random(n) < lower
        s[m] = s[m] | 0x20
    return np.ascontiguousarray(s, dtype=np.uint8)


def _static_call(L, torch, s, k, canon, with_hi, stream=None):
    n = s.size
    d = torch.from_numpy(s).cuda() if n else torch.zeros(16, dtype=torch.uint8, device="cuda")
    cap = max(n - k + 1, 1)
    lo = torch.zeros(cap, dtype=torch.int64, device="cuda")
    hi = torch.zeros(cap if with_hi else 1, dtype=torch.int64, device="cuda")
    pos = torch.zeros(cap, dtype=torch.int64, device="cuda")
    dcount = torch.zeros(1, dtype=torch.int64, device="cuda")
    count = C.c_uint64(0)
    err = C.create_string_buffer(256)
    torch.cuda.synchronize()
    st = L.blg_kmers_static_scratch(C.c_void_p(d.data_ptr()), n, k, int(canon), C.c_void_p(lo.data_ptr()),
                                    C.c_void_p(hi.data_ptr()) if with_hi else None, C.c_void_p(pos.data_ptr()), cap,
                                    C.c_void_p(stream) if stream else None, C.c_void_p(dcount.data_ptr()), C.byref(count), err, 256)
    assert st == 0, (st, err.value)
    torch.cuda.synchronize()
    c = int(count.value)
    assert int(dcount.item()) == c
    return (lo[:c].cpu().numpy().view(np.uint64), hi[:c].cpu().numpy().view(np.uint64) if with_hi else None,
            pos[:c].cpu().numpy().view(np.uint64))


@pytest.mark.gpu
def test_static_scratch_kmers_match_blg_kmers(gpu):
    import torch
    L = gpu.L
    rng = np.random.default_rng(2026)
    # two calls in a row per (k, canonical): a large input first, then a smaller, dirtier one, then an empty one
    for k, canon in ((31, True), (63, True), (21, False)):
        for n, pn in ((3_000_000, 0.0005), (250_000, 0.01), (k - 1, 0.0)):
            s = rand_seq(rng, n, p_n=pn, lower=0.05)
            want = gpu.kmers(s, k, canon)
            lo, hi, pos = _static_call(L, torch, s, k, canon, k > 32)
            assert len(lo) == len(want["lo"]), (k, n)
            assert np.array_equal(lo, want["lo"]) and np.array_equal(pos, want["pos"]), (k, n)
            if k > 32:
                assert np.array_equal(hi, want["hi"]), (k, n)
    # on a stream of the caller
    stream = torch.cuda.Stream()
    s = rand_seq(rng, 1_000_000, p_n=0.001)
    want = gpu.kmers(s, 31, True)
    lo, _, pos = _static_call(L, torch, s, 31, True, False, stream=stream.cuda_stream)
    assert np.array_equal(lo, want["lo"]) and np.array_equal(pos, want["pos"])


@pytest.mark.gpu
def test_static_scratch_limits(gpu):
    import torch
    L = gpu.L
    err = C.create_string_buffer(256)
    count = C.c_uint64(0)
    d = torch.zeros(64, dtype=torch.uint8, device="cuda")
    out = torch.zeros(64, dtype=torch.int64, device="cuda")
    # beyond the stated maximum: a clear error, nothing launched
    st = L.blg_kmers_static_scratch(C.c_void_p(d.data_ptr()), (1 << 26) + 1, 31, 1, C.c_void_p(out.data_ptr()), None, None, 64, None, None,
                                    C.byref(count), err, 256)
    assert st == 3 and b"2^26" in err.value
    st = L.blg_kmers_static_scratch(C.c_void_p(d.data_ptr()), 64, 0, 1, C.c_void_p(out.data_ptr()), None, None, 64, None, None, C.byref(count), err, 256)
    assert st == 1 and err.value
    # output too small: the required size comes back
    s = np.frombuffer(b"ACGT" * 16, np.uint8).copy()
    d = torch.from_numpy(s).cuda()
    st = L.blg_kmers_static_scratch(C.c_void_p(d.data_ptr()), 64, 31, 1, C.c_void_p(out.data_ptr()), None, None, 5, None, None, C.byref(count), err, 256)
    assert st == 3 and count.value == 34
