"""Log-mel error of the tensor-core and the FFT lmfe kernels on spectra with a large dynamic
range (a strong low tone over a weak noise floor): both carry an absolute error relative to the
frame's strongest bins, the float64 reference does not."""
import ctypes
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from oracle import speechpy_oracle as o  # noqa: E402
from speaker_verification_b200 import _lib, batch  # noqa: E402

raw = ctypes.CDLL(_lib.LIB_PATH)
rng = np.random.default_rng(7)
N = 160000
t = np.arange(N) / 16000.0
for floor_db in (-20, -40, -60, -80):
    x = 0.5 * np.sin(2 * np.pi * 220.0 * t) + 0.5 * 10 ** (floor_db / 20) * rng.standard_normal(N)
    x = (np.round(np.clip(x, -1, 1) * 32767) / 32768).astype(np.float32)
    arena, starts, lens = batch.pack([x])
    fb = batch.FeatureBatch(starts, lens)
    d_in = torch.from_numpy(arena).cuda()
    res = {}
    for name, bit in (("tc", 0), ("fft", 4)):
        out = torch.empty((fb.total_frames, 40), dtype=torch.float32, device="cuda")
        with _lib.options(feature_kernel=2 if bit else 0):
            fb.lmfe_dev(d_in.data_ptr(), d_in.numel(), out.data_ptr())
            fb.ctx.sync()
        res[name] = out.cpu().numpy()
    ref = o.lmfe(x, 16000)
    et, ef = np.abs(res["tc"] - ref), np.abs(res["fft"] - ref)
    print(f"noise floor {floor_db:4d} dB re tone: log-mel range {ref.min():.1f}..{ref.max():.1f}  "
          f"tc max {et.max():.2e} p99 {np.quantile(et, 0.99):.2e}   fft max {ef.max():.2e} p99 {np.quantile(ef, 0.99):.2e}")
    fb.close()
