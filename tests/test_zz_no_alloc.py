"""akm_update_centroids_prealloc: the centroid update for hosts that cannot allocate GPU memory
(a shared GPU with a fixed reservation, or a latency-critical thread: cudaMalloc may fail there and it
stalls the whole device).  The entry point keeps its working memory inside the library, so the test
only has to hand it device buffers (allocated here with torch, which the entry point itself never
touches) and compare with the ordinary akm_update_centroids on the same input -- twice in a row,
because the built-in working memory is shared between calls and must not leak state from one to the next.
This file is collected last on purpose."""
import ctypes

import numpy as np
import pytest

from alink_b200 import native

pytestmark = pytest.mark.gpu


def _prealloc_update(L, metric, d, sum_host, k0, k):
    import torch
    sum_dev = torch.from_numpy(np.ascontiguousarray(sum_host, dtype=np.float64)).cuda()
    c_dev = torch.full((k0, d), float("nan"), dtype=torch.float64, device="cuda")
    k_dev = torch.full((1,), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    rc = L.akm_update_centroids_prealloc(0, None, metric, d, ctypes.c_void_p(sum_dev.data_ptr()), k0, k,
                                         ctypes.c_void_p(c_dev.data_ptr()), ctypes.c_void_p(k_dev.data_ptr()))
    assert rc == 0, L.akm_last_error(None)
    torch.cuda.synchronize()
    return c_dev.cpu().numpy(), int(k_dev.item())


@pytest.mark.parametrize("metric", [native.EUCLIDEAN, native.COSINE])
def test_prealloc_update_equals_regular_update_twice_in_a_row(metric):
    L = native.lib()
    L.akm_update_centroids_prealloc.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                                ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                                ctypes.c_void_p]
    rng = np.random.default_rng(17 + metric)
    d = 19
    # two different problems back to back: the second is smaller and has other empty clusters, so stale
    # entries of the shared working memory would show up as wrong compaction
    for k0, k in [(300, 300), (57, 41)]:
        buf = rng.normal(size=(k0, d + 1))
        buf[:, d] = rng.integers(0, 4, size=k0)
        with native.Handle(d, metric) as h:
            C_ref, k_ref, _ = h.update_centroids(buf.ravel(), k0, k)
        C, k_new = _prealloc_update(L, metric, d, buf.ravel(), k0, k)
        assert k_new == k_ref
        assert (C == C_ref).all()                     # the update step is bit-exact in this test suite


def test_prealloc_update_rejects_oversized_k():
    import torch
    L = native.lib()
    L.akm_update_centroids_prealloc.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                                ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                                ctypes.c_void_p]
    t = torch.zeros(8, dtype=torch.float64, device="cuda")
    k0 = 16384 + 1
    rc = L.akm_update_centroids_prealloc(0, None, native.EUCLIDEAN, 1, ctypes.c_void_p(t.data_ptr()), k0, k0,
                                         ctypes.c_void_p(t.data_ptr()), ctypes.c_void_p(t.data_ptr()))
    assert rc == -1
    assert b"AKM_PREALLOC_MAX_K" in L.akm_last_error(None)
