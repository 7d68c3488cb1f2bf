"""gtsfm_retriever_pairs_noalloc (include/gtsfm_b200.h): the allocation-free variant of the retriever step, whose similarity
matrix lives in a fixed __device__ array of the library instead of a cudaMalloc'ed context buffer.  No counterpart in the
reference, so the check is against the existing entry points
gtsfm_retriever_similarity + gtsfm_retriever_topk on the same inputs: indices equal, scores within the tolerance of
tests/test_retriever.py.  Two calls in a row with different sizes share the array (the second one meets the first one's
contents).  This file is collected last on purpose (pytest -x).
"""
import ctypes

import numpy as np
import pytest

TOL = 1e-5          # tests/test_retriever.py


def _regular(ctx, G, blocksize, k, min_score):
    ctx.retriever_similarity(G, blocksize, fetch=False)
    return ctx.retriever_topk(None, len(G), k, min_score)


def _noalloc(native, torch, G, blocksize, k, min_score):
    L = native.lib()
    n, dim = G.shape
    d_g = torch.from_numpy(G).cuda()
    d_idx = torch.full((n, k), -7, dtype=torch.int32, device="cuda")
    d_score = torch.full((n, k), 123.0, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    ms = float("nan") if min_score is None else float(min_score)
    stream = torch.cuda.current_stream().cuda_stream
    rc = L.gtsfm_retriever_pairs_noalloc(ctypes.c_void_p(d_g.data_ptr()), n, dim, blocksize, k, ms, ctypes.c_void_p(d_idx.data_ptr()),
                                         ctypes.c_void_p(d_score.data_ptr()), ctypes.c_void_p(stream))
    assert rc == native.OK, L.gtsfm_last_error(None).decode()
    torch.cuda.synchronize()
    return d_idx.cpu().numpy(), d_score.cpu().numpy()


@pytest.mark.gpu
def test_noalloc_retriever_equals_regular_entry_points_twice_in_a_row():
    import torch
    import gtsfm_b200
    from gtsfm_b200 import _native as native
    rng = np.random.default_rng(77)
    ctx = gtsfm_b200.Context(0)
    torch.cuda.set_device(0)
    try:
        # call 1: 601 images (ragged 64-wide tiles), duplicates -> ties; call 2: smaller, other row stride, no minimum score
        for n, dim, blocksize, k, min_score in [(601, 512, 50, 20, 0.0), (130, 100, 7, 5, None), (1024, 64, 50, 3, 0.01)]:
            G = rng.standard_normal((n, dim)).astype(np.float32)
            G /= np.linalg.norm(G, axis=1, keepdims=True)
            G[n // 2] = G[3]
            G[n - 2] = G[3]
            want_idx, want_score = _regular(ctx, G, blocksize, k, min_score)
            got_idx, got_score = _noalloc(native, torch, G, blocksize, k, min_score)
            np.testing.assert_array_equal(got_idx, want_idx)
            fin = np.isfinite(want_score)
            assert (np.isfinite(got_score) == fin).all()
            np.testing.assert_allclose(got_score[fin], want_score[fin], atol=TOL, rtol=0)
            assert (got_score[~fin] == want_score[~fin]).all()          # unused slots: -inf in both
            assert (got_idx[3, :2] == [n // 2, n - 2]).all()             # duplicates of image 3: score 1, lower index first
    finally:
        ctx.close()


@pytest.mark.gpu
def test_noalloc_retriever_rejects_sizes_beyond_the_fixed_scratch():
    import torch
    from gtsfm_b200 import _native as native
    L = native.lib()
    torch.cuda.set_device(0)
    n = 1025                                                              # GTSFM_RETRIEVER_NOALLOC_MAX_IMAGES + 1
    d_g = torch.zeros((n, 8), dtype=torch.float32, device="cuda")
    d_idx = torch.zeros((n, 2), dtype=torch.int32, device="cuda")
    d_score = torch.zeros((n, 2), dtype=torch.float32, device="cuda")
    rc = L.gtsfm_retriever_pairs_noalloc(ctypes.c_void_p(d_g.data_ptr()), n, 8, 50, 2, float("nan"), ctypes.c_void_p(d_idx.data_ptr()),
                                         ctypes.c_void_p(d_score.data_ptr()), None)
    assert rc == native.E_ARG
    msg = L.gtsfm_last_error(None).decode()
    assert "1025" in msg and "1024" in msg
    rc = L.gtsfm_retriever_pairs_noalloc(ctypes.c_void_p(d_g.data_ptr()), 4, 8, 50, 5, float("nan"), ctypes.c_void_p(d_idx.data_ptr()),
                                         ctypes.c_void_p(d_score.data_ptr()), None)
    assert rc == native.E_ARG                                             # num_select > n_images
