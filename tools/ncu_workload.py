"""Workload for the ncu --set full captures: one cfg2-sized (4 874 x 8 s) lmfe + cmvnw batch, three
calls, then the cfg5 cosine GEMM three times (kernels are selected with ncu -k / -s / -c)."""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from speaker_verification_b200 import batch, scoring  # noqa: E402

rng = np.random.default_rng(0)
U, N = 4874, 128000
big = torch.from_numpy((np.round(rng.normal(0, 0.1, N * 64) * 32768) / 32768).astype(np.float32)).cuda().repeat(U // 64 + 1)[: U * N]
fb = batch.FeatureBatch(np.arange(U, dtype=np.int64) * N, np.full(U, N, dtype=np.int64))
out = torch.empty((fb.total_frames, 40), dtype=torch.float32, device="cuda")
for _ in range(3):
    fb.lmfe_cmvnw_dev(big.data_ptr(), big.numel(), out.data_ptr())
fb.ctx.sync()
g = torch.Generator(device="cuda")
g.manual_seed(5)
E = torch.randn((4874, 512), generator=g, device="cuda")
T = torch.randn((4874, 512), generator=g, device="cuda")
S = torch.empty((4874, 4874), device="cuda")
for _ in range(6):
    scoring.cosine_scores_dev(E, T, S)
torch.cuda.synchronize()
