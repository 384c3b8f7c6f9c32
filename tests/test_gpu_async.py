"""BLG_F_ASYNC: device-resident calls that only enqueue; the count is collected with blg_async_count."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_async_calls_match_sync(gpu, oracle):
    import torch
    from test_gpu_parity import rand_seq
    rng = np.random.default_rng(3)
    s = rand_seq(rng, 500_000, 0.0005)
    d = torch.from_numpy(s).cuda()
    cap = s.size
    out = [torch.empty(cap, dtype=torch.int64, device="cuda") for _ in range(2)] + [torch.empty(cap, dtype=torch.uint8, device="cuda") for _ in range(2)]
    ptrs = [t.data_ptr() for t in out]
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    gpu.set_stream(stream.cuda_stream)
    try:
        for k, m in ((31, 19), (63, 31), (21, 11), (40, 9)):
            o = oracle.super_kmers(s, k, m, 42, True)
            # several launches in flight (different window ranges), then the full one: the count is the last call's
            for lo, hi in ((0, 1000), (100_000, 300_000), (0, 2**63)):
                r = gpu.super_kmers_device(d.data_ptr(), s.size, k, m, 42, True, *ptrs, cap, win_range=(lo, hi), async_call=True)
                assert r == 0 or not __import__("biolib_b200")._capi.lib().blg_abi_version() < 2   # async (sk4) returns 0; other kernels run synchronously
            n = gpu.async_count() if r == 0 else r
            assert n == o["n"]
            assert np.array_equal(out[0][:n].cpu().numpy().view(np.uint64), o["value"])
            assert np.array_equal(out[1][:n].cpu().numpy().view(np.uint64), o["mm_start"])
            assert np.array_equal(out[2][:n].cpu().numpy(), o["size"])
        # capacity verdict arrives with the count
        r = gpu.super_kmers_device(d.data_ptr(), s.size, 31, 19, 42, True, *ptrs, 10, async_call=True)
        with pytest.raises(Exception):
            gpu.async_count()
    finally:
        gpu.set_stream(None)
        torch.cuda.synchronize()
        torch.cuda.set_stream(torch.cuda.default_stream())
