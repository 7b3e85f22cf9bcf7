"""Time mfcc + delta (cfg2: 4874 x 8 s) through the tensor-core path and the FFT kernel."""
import ctypes
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from speaker_verification_b200 import _lib, batch  # noqa: E402

rng = np.random.default_rng(0)
U, N = 4874, 128000
big = torch.from_numpy((np.round(rng.normal(0, 0.1, N * 64) * 32768) / 32768).astype(np.float32)).cuda().repeat(U // 64 + 1)[: U * N]
fb = batch.FeatureBatch(np.arange(U, dtype=np.int64) * N, np.full(U, N, dtype=np.int64))
out = torch.empty((fb.total_frames, 13, 3), dtype=torch.float32, device="cuda")
for tc in (True, False):
    _lib.set_option("feature_kernel", 0 if tc else 2)
    for _ in range(3):
        fb.mfcc_delta_dev(big.data_ptr(), big.numel(), out.data_ptr())
    fb.ctx.sync()
    t0 = time.perf_counter()
    for _ in range(10):
        fb.mfcc_delta_dev(big.data_ptr(), big.numel(), out.data_ptr())
    fb.ctx.sync()
    dt = (time.perf_counter() - t0) / 10
    print(f"mfcc+delta {'tc ' if tc else 'fft'}: {dt*1e3:.3f} ms per {U*N/16000/3600:.2f} audio-h -> {U*N/16000/3600/dt:.0f} audio-h/s")
_lib.set_option("feature_kernel", 0)
