"""Time the lmfe launch on 4874 x 8 s (10.83 audio-h) for every kernel variant, CUDA events on the
context's stream.  Usage: python tools/tc_time.py [lmfe|lmfe_cmvnw]"""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from speaker_verification_b200 import _lib, batch  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "lmfe"
rng = np.random.default_rng(0)
U, N = 4874, 128000
big = torch.from_numpy((np.round(rng.normal(0, 0.1, N * 64) * 32768) / 32768).astype(np.float32)).cuda().repeat(U // 64 + 1)[: U * N]
fb = batch.FeatureBatch(np.arange(U, dtype=np.int64) * N, np.full(U, N, dtype=np.int64))
out = torch.empty((fb.total_frames, 40), dtype=torch.float32, device="cuda")
stream = torch.cuda.ExternalStream(fb.ctx.stream())
call = fb.lmfe_dev if what == "lmfe" else fb.lmfe_cmvnw_dev
variants = [
    ("tc2 static (default)", dict()),
    ("tc2 static, no redo", dict(redo=0)),
    ("tc2 runs-of-four epilogue", dict(tc_epilogue=1)),
    ("tc2 generic epilogue", dict(tc_epilogue=0)),
    ("tc1", dict(tc_version=1)),
    ("fft512", dict(feature_kernel=2)),
]
if len(sys.argv) > 2 and sys.argv[2] == "one":
    variants = variants[:1]
elif what != "lmfe":
    variants += [("tc2 + single-pass streaming cmvnw", dict(cmvnw_kernel=3))]
for name, opts in variants:
    with _lib.options(**opts):
        for _ in range(3):
            call(big.data_ptr(), big.numel(), out.data_ptr())
        fb.ctx.sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(10):
            call(big.data_ptr(), big.numel(), out.data_ptr())
        e1.record(stream)
        fb.ctx.sync()
        dt = e0.elapsed_time(e1) / 10 / 1e3
    print(f"{what} {name:34s}: {dt*1e3:7.3f} ms -> {U*N/16000/3600/dt:7.0f} audio-h/s", flush=True)
