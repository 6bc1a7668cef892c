"""bb_gram_noalloc: the Gram build for callers that cannot call cudaMalloc (include/bayesbridge_b200.h).

It has no counterpart in the reference; the check is against the
existing entry point bb_gram on the same device buffers: same kernels, same row split, so the two must agree bit for
bit (asserted to the 1e-12 the Gram tests use, and exactly).  Collected last on purpose:
a failure here hides no other test.  The test owns its inputs and outputs (torch device tensors); the
entry point itself allocates nothing, which the free-memory check below watches.
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

MAX_P = 3071


def test_noalloc_refuses_sizes_beyond_the_static_scratch():
    """the size check comes before any CUDA call, so this runs with or without a GPU"""
    import bayesbridge_b200 as bb
    L = bb._lib.load()
    assert L.bb_gram_noalloc(None, None, 100, MAX_P + 1, None, 0) == 1
    msg = bb._lib.last_error()
    assert "bb_gram_noalloc" in msg and str(MAX_P) in msg and "bb_gram" in msg
    assert L.bb_gram_noalloc(None, None, 0, 8, None, 0) == 1
    assert "positive" in bb._lib.last_error()


def _problem(N, P, seed):
    rng = np.random.default_rng(seed)
    X = np.asfortranarray(rng.standard_normal((N, P)) * (1.0 + np.arange(P))[None, :] ** 0.25)
    y = X @ rng.standard_normal(P) + rng.standard_normal(N)
    return X, y


@pytest.mark.gpu
def test_noalloc_gram_matches_bb_gram_on_consecutive_calls():
    import torch
    import bayesbridge_b200 as bb
    L = bb._lib.load()
    assert L.bb_device_count() >= 1, "no CUDA device (there is no CPU fallback)"
    dev = torch.device("cuda", 0)

    def both(N, P, seed):
        X, y = _problem(N, P, seed)
        Xd = torch.from_numpy(X.ravel(order="F").copy()).to(dev)
        yd = torch.from_numpy(y.copy()).to(dev)
        G0 = torch.full(((P + 1) * (P + 1),), float("nan"), dtype=torch.float64, device=dev)
        G1 = torch.full(((P + 1) * (P + 1),), float("nan"), dtype=torch.float64, device=dev)
        torch.cuda.synchronize(dev)
        assert L.bb_gram(Xd.data_ptr(), yd.data_ptr(), N, P, G0.data_ptr(), 0) == 0, bb._lib.last_error()
        free_before = torch.cuda.mem_get_info(dev)[0]
        assert L.bb_gram_noalloc(Xd.data_ptr(), yd.data_ptr(), N, P, G1.data_ptr(), 0) == 0, bb._lib.last_error()
        free_after = torch.cuda.mem_get_info(dev)[0]
        torch.cuda.synchronize(dev)
        g0 = G0.cpu().numpy().reshape(P + 1, P + 1)
        g1 = G1.cpu().numpy().reshape(P + 1, P + 1)
        A = np.column_stack([X, y])
        return g0, g1, A.T @ A, free_before, free_after

    # two calls in a row on the shared scratch: a large multi-split problem first, then a small one whose slots all
    # hold the first call's leftovers, then the large one again
    frees = []
    for k, (N, P, seed) in enumerate([(6000, 300, 1), (442, 64, 2), (6000, 300, 3), (500, 2000, 4)]):
        g0, g1, want, fb, fa = both(N, P, seed)
        assert np.isfinite(g1).all()
        scale = np.abs(want).max()
        assert np.abs(g1 - want).max() <= 1e-12 * scale          # against numpy, the tolerance of the Gram tests
        assert np.abs(g1 - g0).max() <= 1e-12 * scale            # against the existing entry point, same tolerance
        assert np.array_equal(g1, g0), f"call {k}: not bit-identical to bb_gram"
        assert np.array_equal(g1, g1.T)
        frees.append((fb, fa))
    # once the first call has done the per-device set-up, a call leaves the device's free memory where it was
    for fb, fa in frees[1:]:
        assert fa == fb, f"free device memory moved across a bb_gram_noalloc call: {fb} -> {fa}"
