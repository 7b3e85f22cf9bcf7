import sys, numpy as np, torch
sys.path.insert(0, ".")
from speaker_verification_b200 import _lib, batch
rng = np.random.default_rng(0)
U, N = 4874, 128000
big = torch.from_numpy((np.round(rng.normal(0, 0.1, N * 64) * 32768) / 32768).astype(np.float32)).cuda().repeat(U // 64 + 1)[: U * N]
fb = batch.FeatureBatch(np.arange(U, dtype=np.int64) * N, np.full(U, N, dtype=np.int64))
out = torch.empty((fb.total_frames, 40), dtype=torch.float32, device="cuda")
for opts in (dict(), dict(cmvnw_kernel=2)):
    with _lib.options(**opts):
        for _ in range(3): fb.lmfe_cmvnw_dev(big.data_ptr(), big.numel(), out.data_ptr())
        fb.ctx.sync(); fb.ctx.timing_enable(True)
        for _ in range(10): fb.lmfe_cmvnw_dev(big.data_ptr(), big.numel(), out.data_ptr())
        print(opts, {k: tuple(round(x, 3) if isinstance(x, float) else x for x in fb.ctx.timing_read(k)) for k in (0, 1, 2, 4, 5)})
        fb.ctx.timing_enable(False)
