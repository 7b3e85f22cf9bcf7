"""Accumulation noise of the tensor-core chain: GPU lmfe vs an exact (float64) evaluation of the
same fp16-split products.  SPB_TC_DBG=16: hi.hi term only (10-MMA chain); 0: all three terms."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from oracle import speechpy_oracle as o  # noqa: E402
from speaker_verification_b200 import batch, synth  # noqa: E402

FS = 16000
hh_only = (int(os.environ.get("SPB_TC_DBG", "0")) & 16) != 0
x = synth.audio(908, 320000, np.float32)
arena, starts, lens = batch.pack([x])
fb = batch.FeatureBatch(starts, lens)
out = torch.empty((fb.total_frames, 40), dtype=torch.float32, device="cuda")
a_t = torch.from_numpy(arena).cuda()
fb.lmfe_dev(a_t.data_ptr(), a_t.numel(), out.data_ptr())
fb.ctx.sync()
got = out.cpu().numpy().astype(np.float64)
F = got.shape[0]
n = np.arange(160)
bins = np.arange(5, 128)
ang = 2 * np.pi * (np.outer(n, bins) % 512) / 512
W = np.concatenate([np.cos(ang), -np.sin(ang)], 1)
Wh = W.astype(np.float32).astype(np.float16).astype(np.float64)
Wl = (W - Wh).astype(np.float32).astype(np.float16).astype(np.float64)
xs = (x[: (F + 1) * 160].reshape(F + 1, 160) * np.float32(32768)).astype(np.float32)
xh = xs.astype(np.float16)
xl = (xs - xh.astype(np.float32)).astype(np.float16).astype(np.float64)
xh = xh.astype(np.float64)
A = xh @ Wh if hh_only else xh @ Wh + xl @ Wh + xh @ Wl
Ah = A[:, :123] + 1j * A[:, 123:]
c = np.exp(-2j * np.pi * 160 * bins / 512)
X = Ah[:-1] + c[None, :] * Ah[1:]
P = np.abs(X) ** 2 / (512.0 * 2.0 ** 30)
fb_m = o.filterbanks(40, 257, FS, 0, None)[:, 5:128]
ref = np.log(P @ fb_m.T)
d = np.abs(got - ref)
print(f"hh_only={hh_only} F={F} max {d.max():.3e} p99.9 {np.quantile(d, 0.999):.3e} median {np.median(d):.3e}")
true = o.lmfe(x, FS)
d2 = np.abs(got - true)
print(f"   vs oracle: max {d2.max():.3e} p99.9 {np.quantile(d2, 0.999):.3e} median {np.median(d2):.3e}")
