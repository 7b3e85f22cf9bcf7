"""spb_cosine_scores_dev_noalloc: the allocation-free variant of the cosine score entry point
(static __device__ scratch inside the library; include/speechpy_b200.h).  It has no counterpart in
the reference -- it was added for callers that cannot call
cudaMalloc -- so the check is against the existing entry point spb_cosine_scores_dev (and the
oracle), through the C ABI, on two calls in a row that share the static array.

Collected last on purpose: a failure here hides no other test.
"""
import numpy as np
import pytest

from oracle import scoring_oracle as so
from speaker_verification_b200 import _lib

pytestmark = pytest.mark.gpu
REL, FLOOR = 1e-5, 0.02  # the cosine tolerance of tests/test_gpu_scoring.py


def _close(got, ref):
    bound = REL * np.maximum(np.abs(ref), FLOOR)
    err = np.abs(got.astype(np.float64) - ref.astype(np.float64))
    assert np.all(err <= bound), (float(err.max()), float((err / bound).max()))


def _both(ctx, torch, E, T):
    """(existing entry point, allocation-free entry point) on the same device inputs."""
    dE, dT = torch.from_numpy(E).cuda(), torch.from_numpy(T).cuda()
    m, d = E.shape
    n = T.shape[0]
    ref = torch.full((m, n), 7.0, dtype=torch.float32, device="cuda")
    got = torch.full((m, n), -7.0, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    _lib.check(ctx.lib.spb_cosine_scores_dev(ctx.handle, dE.data_ptr(), m, dT.data_ptr(), n, d, ref.data_ptr()),
               ctx.handle)
    _lib.check(ctx.lib.spb_cosine_scores_dev_noalloc(ctx.handle, dE.data_ptr(), m, dT.data_ptr(), n, d,
                                                     got.data_ptr()), ctx.handle)
    ctx.sync()
    return ref.cpu().numpy(), got.cpu().numpy()


def test_noalloc_cosine_matches_existing_entry_point_twice(lib_built):
    import torch
    ctx = _lib.context(0)
    rng = np.random.Generator(np.random.Philox(2024))
    # first call: close to the limit (1 024 + 1 024 rows at d = 512 = exactly 8 MiB of split scratch)
    E1 = rng.standard_normal((1024, 512)).astype(np.float32)
    T1 = rng.standard_normal((1024, 512)).astype(np.float32)
    E1[5] = 0.0  # sklearn: zero rows stay zero
    ref1, got1 = _both(ctx, torch, E1, T1)
    _close(got1, ref1)
    _close(got1, so.cosine_scores(E1, T1))
    assert np.all(got1[5] == 0.0)
    # second call in a row: a smaller, ragged problem (d padded to 128) over the first call's leftovers
    E2 = rng.standard_normal((130, 100)).astype(np.float32)
    T2 = rng.standard_normal((301, 100)).astype(np.float32)
    ref2, got2 = _both(ctx, torch, E2, T2)
    _close(got2, ref2)
    _close(got2, so.cosine_scores(E2, T2))
    # and the first problem again: the array is shared, every call rewrites what it reads
    ref3, got3 = _both(ctx, torch, E1, T1)
    _close(got3, ref3)
    assert np.array_equal(got3, got1)
    # the call itself takes no device memory: buffers made first, free memory read around two calls
    dE, dT = torch.from_numpy(E2).cuda(), torch.from_numpy(T2).cuda()
    out = torch.empty((E2.shape[0], T2.shape[0]), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for _ in range(2):
        _lib.check(ctx.lib.spb_cosine_scores_dev_noalloc(ctx.handle, dE.data_ptr(), E2.shape[0], dT.data_ptr(),
                                                         T2.shape[0], E2.shape[1], out.data_ptr()), ctx.handle)
    ctx.sync()
    assert torch.cuda.mem_get_info()[0] >= free0
    assert np.array_equal(out.cpu().numpy(), got2)


def test_noalloc_cosine_rejects_sizes_beyond_the_static_scratch(lib_built):
    import torch
    ctx = _lib.context(0)
    m = n = 1025
    d = 512
    E = torch.zeros((m, d), device="cuda")
    T = torch.zeros((n, d), device="cuda")
    S = torch.zeros((m, n), device="cuda")
    rc = ctx.lib.spb_cosine_scores_dev_noalloc(ctx.handle, E.data_ptr(), m, T.data_ptr(), n, d, S.data_ptr())
    assert rc == _lib.SPB_ERR_INVALID
    msg = ctx.lib.spb_last_error(ctx.handle).decode()
    assert "static scratch" in msg and str(8 << 20) in msg
    # empty problems are fine and touch nothing
    assert ctx.lib.spb_cosine_scores_dev_noalloc(ctx.handle, E.data_ptr(), 0, T.data_ptr(), n, d, S.data_ptr()) == 0
