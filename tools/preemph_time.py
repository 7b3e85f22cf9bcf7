"""lmfe with and without the fused pre-emphasis prologue (4 874 x 8 s, device-resident) on one B200."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from speaker_verification_b200 import _lib, batch

U, N = 4874, 128000
starts = np.arange(U, dtype=np.int64) * N
lens = np.full(U, N, dtype=np.int64)
d = torch.randn(U * N, device="cuda") * 0.1
for name, pre, opts in (("tensor-core kernel, no pre-emphasis", 0.0, {}), ("FFT kernel, no pre-emphasis", 0.0, dict(feature_kernel=2)),
                        ("tensor-core kernel, pre-emphasis 0.98 in the epilogue", 0.98, {}),
                        ("the same without the float64 guard", 0.98, dict(redo=0)),
                        ("FFT kernel, pre-emphasis 0.98 on the staged tile", 0.98, dict(feature_kernel=2)),
                        ("the same without the float64 guard", 0.98, dict(feature_kernel=2, redo=0))):
    with _lib.options(**opts):
        fb = batch.FeatureBatch(starts, lens, preemphasis=pre)
        out = torch.empty((fb.total_frames, 40), device="cuda")
        for _ in range(3):
            fb.lmfe_dev(d.data_ptr(), d.numel(), out.data_ptr())
        fb.ctx.sync()
        t = time.perf_counter()
        for _ in range(10):
            fb.lmfe_dev(d.data_ptr(), d.numel(), out.data_ptr())
        fb.ctx.sync()
        dt = (time.perf_counter() - t) / 10
        print(f"{name:56s}: {dt * 1e3:7.3f} ms -> {U * 8 / 3600 / dt:8.0f} audio-h/s")
        fb.close()
