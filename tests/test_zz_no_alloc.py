"""GPU test of the allocation-free entry point smc_resample_counts_noalloc (include/smcsmc_b200.h): called through the
C ABI with caller-owned device buffers, its offspring counts must equal smc_resample_counts' bit for bit and its ESS
to the last bit (same kernels, same order of operations), on two calls in a row -- the second call finds the static
scratch array holding the first call's partial sums and u-chain.  Collected last on purpose
(file name)."""
import ctypes

import numpy as np
import pytest

import smcsmc_b200 as S
from smcsmc_b200._abi import check, lib

pytestmark = pytest.mark.gpu

MAX_N = 131072   # SMC_NOALLOC_MAX_N


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _noalloc(torch, w, u, parity_scan, stream=None):
    """the caller side: device buffers from torch (the entry point itself allocates nothing)"""
    dev = torch.device("cuda:0")
    d_w = torch.from_numpy(w).to(dev)
    d_ess = torch.full((1,), -1.0, dtype=torch.float64, device=dev)
    d_counts = torch.full((len(w),), -7, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(0)[0]
    check(lib().smc_resample_counts_noalloc(0, _ptr(d_w), len(w), u, int(parity_scan), _ptr(d_ess), _ptr(d_counts),
                                            ctypes.c_void_p(stream.cuda_stream if stream is not None else 0)))
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info(0)[0]
    return d_ess.cpu().numpy()[0], d_counts.cpu().numpy(), free0 - free1


@pytest.mark.parametrize("parity_scan", [1, 0])
def test_noalloc_equals_allocating_entry_point_twice_in_a_row(parity_scan):
    import torch
    rng = np.random.default_rng(2026)
    # two different problems back to back: the second (smaller N, other u) runs over the first one's leftovers
    cases = [(50000, 0.37), (777, 0.91), (MAX_N, 0.0)]
    grew = []
    for N, u in cases:
        w = rng.gamma(0.3, size=N)           # heavy-tailed weights: many zero-offspring and multi-offspring parents
        w[rng.integers(0, N, size=N // 50)] = 0.0
        ess_ref, counts_ref = S.resample_counts(w, u, parity_scan=bool(parity_scan))
        ess, counts, used = _noalloc(torch, w, u, parity_scan)
        assert counts.sum() == N
        assert np.array_equal(counts, counts_ref), "offspring counts differ from smc_resample_counts (bit-exact bar)"
        assert ess == ess_ref, (ess, ess_ref)
        grew.append(used)
    # after the first call (lazy module load / symbol look-up) device memory in use must not grow
    assert all(g <= 0 for g in grew[1:]), grew


def test_noalloc_on_a_side_stream_and_repeated():
    import torch
    rng = np.random.default_rng(7)
    w = rng.random(4096)
    st = torch.cuda.Stream()
    ess_ref, counts_ref = S.resample_counts(w, 0.5, parity_scan=True)
    for _ in range(2):
        ess, counts, _used = _noalloc(torch, w, 0.5, 1, stream=st)
        assert np.array_equal(counts, counts_ref) and ess == ess_ref


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    import torch
    dev = torch.device("cuda:0")
    d_w = torch.ones(8, dtype=torch.float64, device=dev)
    d_ess = torch.zeros(1, dtype=torch.float64, device=dev)
    d_counts = torch.zeros(8, dtype=torch.int32, device=dev)
    rc = lib().smc_resample_counts_noalloc(0, _ptr(d_w), MAX_N + 1, 0.5, 1, _ptr(d_ess), _ptr(d_counts), None)
    assert rc == 1   # SMC_ERR_ARG
    msg = lib().smc_last_error().decode()
    assert "SMC_NOALLOC_MAX_N" in msg and str(MAX_N) in msg
    assert lib().smc_resample_counts_noalloc(0, None, 8, 0.5, 1, _ptr(d_ess), _ptr(d_counts), None) == 1
    assert lib().smc_resample_counts_noalloc(0, _ptr(d_w), 8, 1.0, 1, _ptr(d_ess), _ptr(d_counts), None) == 1
