"""Time the lmfe launch (tensor-core path unless SPB_TC=0) on 4874 x 8 s."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from speaker_verification_b200 import batch  # noqa: E402

rng = np.random.default_rng(0)
U, N = 4874, 128000
big = torch.from_numpy((np.round(rng.normal(0, 0.1, N * 64) * 32768) / 32768).astype(np.float32)).cuda().repeat(U // 64 + 1)[: U * N]
fb3 = batch.FeatureBatch(np.arange(U, dtype=np.int64) * N, np.full(U, N, dtype=np.int64))
out = torch.empty((fb3.total_frames, 40), dtype=torch.float32, device="cuda")
for _ in range(3):
    fb3.lmfe_dev(big.data_ptr(), big.numel(), out.data_ptr())
fb3.ctx.sync()
t0 = time.perf_counter()
for _ in range(10):
    fb3.lmfe_dev(big.data_ptr(), big.numel(), out.data_ptr())
fb3.ctx.sync()
dt = (time.perf_counter() - t0) / 10
import os
print(f"dbg={os.environ.get('SPB_TC_DBG')} tc={os.environ.get('SPB_TC')}: {dt*1e3:.3f} ms -> {U*N/16000/3600/dt:.0f} audio-h/s")
