"""b200_outlier_trimmed_noalloc (no counterpart in the reference): the
allocation-free TrimmedDist filter over a static __device__ scratch array, checked through the C ABI against the
existing allocating entry points b200_outlier_dist / b200_dists_quantile — weights and limit BIT-EXACT (the bar the
other outlier-filter tests use), on two calls in a row (the second must not see the first call's scratch contents).

Named test_zz_* so that pytest collects it last. torch only provides the caller-owned device buffers.
"""
import os

import numpy as np
import pytest

from open3d_slam_private_b200 import capi as B

pytestmark = pytest.mark.gpu


def _dists():
    """squared match distances with invalid (+inf) entries and many ties, like a k = 3 match with a small maxDist"""
    rng = np.random.default_rng(7)
    d = (rng.gamma(2.0, 0.01, size=(40000, 3)) ** 2).astype(np.float32)
    d[rng.random(d.shape) < 0.2] = np.inf
    d[:500] = d[500:1000]                      # exact duplicates around every quantile
    return np.ascontiguousarray(d)


def _run(torch, d_dev, n, ratio, stream=0):
    w = torch.full((n,), -7.0, dtype=torch.float32, device="cuda")          # poisoned: every entry must be written
    out = torch.full((2,), 0xdead, dtype=torch.int32, device="cuda")
    B.outlier_trimmed_noalloc(d_dev.data_ptr(), n, ratio, w.data_ptr(), out.data_ptr(), stream=stream)
    torch.cuda.synchronize()
    o = out.cpu().numpy().view(np.uint32)
    return w.cpu().numpy(), int(o[0]), o[1:2].view(np.float32)[0]


def test_noalloc_trimmed_equals_allocating_entry_point_twice_in_a_row():
    import torch
    B.require_device()
    d = _dists()
    n = d.size
    d_dev = torch.from_numpy(d.reshape(-1)).cuda()
    # two calls in a row with different ratios, then the first again: the shared scratch is cleared by every call
    for ratio in (0.9, 0.35, 0.9, 1.0, 0.0):
        want_w = B.outlier_dist(B.OUTLIER_TRIMMEDDIST, ratio, d).reshape(-1)
        want_lim = B.dists_quantile(d, ratio)
        w, st, lim = _run(torch, d_dev, n, ratio)
        assert st == B.B200_OK
        assert np.float32(lim).view(np.uint32) == np.float32(want_lim).view(np.uint32), ratio
        assert (w.view(np.uint32) == want_w.view(np.uint32)).all(), ratio
    # on a caller's own stream as well
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        w, st, lim = _run(torch, d_dev, n, 0.5, stream=s.cuda_stream)
    assert st == B.B200_OK and (w == B.outlier_dist(B.OUTLIER_TRIMMEDDIST, 0.5, d).reshape(-1)).all()


def test_noalloc_trimmed_no_matches_limits_and_reuse_after_failure():
    import torch
    B.require_device()
    inf = torch.full((1000,), float("inf"), dtype=torch.float32, device="cuda")
    w, st, lim = _run(torch, inf, 1000, 0.9)
    assert st == B.B200_E_NO_MATCHES and (w == -7.0).all()        # like the allocating call: no weights, an error
    with pytest.raises(B.ConvergenceError):
        B.outlier_dist(B.OUTLIER_TRIMMEDDIST, 0.9, np.full((1000, 1), np.inf, np.float32))
    # the failed call's status word does not leak into the next call
    d = _dists()
    d_dev = torch.from_numpy(d.reshape(-1)).cuda()
    w, st, lim = _run(torch, d_dev, d.size, 0.85)
    assert st == B.B200_OK and (w == B.outlier_dist(B.OUTLIER_TRIMMEDDIST, 0.85, d).reshape(-1)).all()
    # sizes beyond the stated maximum and bad arguments are rejected with a clear error, nothing is enqueued
    with pytest.raises(B.B200Error, match="B200_NOALLOC_MAX_N"):
        B.outlier_trimmed_noalloc(d_dev.data_ptr(), B.NOALLOC_MAX_N + 1, 0.9, d_dev.data_ptr())
    with pytest.raises(B.B200Error, match=r"\[0,1\]"):
        B.outlier_trimmed_noalloc(d_dev.data_ptr(), 10, 1.5, d_dev.data_ptr())
    with pytest.raises(B.B200Error):
        B.outlier_trimmed_noalloc(0, 10, 0.5, 0)
    # n = 0: nothing valid -> the device reports "no matches"
    w0 = torch.zeros(1, dtype=torch.float32, device="cuda")
    out = torch.zeros(2, dtype=torch.int32, device="cuda")
    B.outlier_trimmed_noalloc(0, 0, 0.5, 0, out.data_ptr())
    torch.cuda.synchronize()
    assert int(out.cpu()[0]) == B.B200_E_NO_MATCHES
