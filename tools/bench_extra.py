"""GPU: throughput of the other BASELINE configs through the device-pointer batch API
(CUDA events on the context's stream, 3 warm-ups, 10 timed calls each).
  cfg2  4874 x 8 s   mfcc(13) + delta + delta-delta      (spb_mfcc_delta_batch_dev)
  cfg4  1024 x 60 s  lmfe + cmvnw(301, var)               (spb_lmfe_cmvnw_batch_dev)
  f1    cfg2 corpus  lmfe -> delta cube -> cmvn -> 20 crops of 80 frames (spb_feature_cubes_batch_dev)
Inputs: N(0, 0.1) float32 generated on the device; results are checked for finiteness only here
(parity is the job of tests/)."""
import json
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from speaker_verification_b200 import batch  # noqa: E402

FS = 16000
dev = torch.device("cuda", 0)


def corpus(U, n):
    lengths = np.full(U, n, dtype=np.int64)
    starts = np.arange(U, dtype=np.int64) * n
    arena = torch.empty(U * n, dtype=torch.float32, device=dev).normal_(0.0, 0.1)
    return arena, starts, lengths


def timed(fn, ctx, reps=10):
    stream = torch.cuda.ExternalStream(ctx.stream(), device=dev)
    for _ in range(3):
        fn()
    ctx.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps):
        fn()
    e1.record(stream)
    ctx.sync()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def report(name, ms, audio_h, alg_bytes, extra=None):
    line = {"config": name, "ms_per_call": ms, "audio_hours_per_s": audio_h / (ms / 1e3),
            "algorithmic_GBps": alg_bytes / (ms / 1e3) / 1e9}
    line.update(extra or {})
    print(json.dumps(line), flush=True)


# ---- cfg2
arena, starts, lengths = corpus(4874, 128000)
fb = batch.FeatureBatch(starts, lengths)
out = torch.empty((fb.total_frames, 13, 3), dtype=torch.float32, device=dev)
ms = timed(lambda: fb.mfcc_delta_dev(arena.data_ptr(), arena.numel(), out.data_ptr()), fb.ctx)
assert bool(torch.isfinite(out).all())
ah = lengths.sum() / FS / 3600
report("cfg2 mfcc+delta+delta-delta, 4874 x 8 s", ms, ah, 4 * lengths.sum() + 156 * fb.total_frames)
# ---- f1 on the same corpus
F = np.diff(fb.frame_offsets)
crops = np.stack([np.random.RandomState(u).randint(int(f) - 80, size=20) for u, f in enumerate(F)]).astype(np.int32)
for deriv, norm in ((False, False), (True, True)):
    ch = 3 if deriv else 1
    cubes = torch.empty((len(F), ch, 20, 80, 40), dtype=torch.float32, device=dev)
    ms = timed(lambda: fb.feature_cubes_dev(arena.data_ptr(), arena.numel(), cubes.data_ptr(), crops,
                                            derivative=deriv, normalize=norm), fb.ctx)
    assert bool(torch.isfinite(cubes).all())
    report(f"f1 feature cubes derivative={deriv} normalize={norm}, 4874 x 8 s", ms, ah,
           4 * lengths.sum() + cubes.numel() * 4, {"cubes_per_s": len(F) / (ms / 1e3)})
del out, cubes, arena
fb.close()
# ---- cfg4
arena, starts, lengths = corpus(1024, 960000)
fb = batch.FeatureBatch(starts, lengths)
out = torch.empty((fb.total_frames, 40), dtype=torch.float32, device=dev)
ms = timed(lambda: fb.lmfe_cmvnw_dev(arena.data_ptr(), arena.numel(), out.data_ptr()), fb.ctx)
assert bool(torch.isfinite(out).all())
report("cfg4 lmfe+cmvnw, 1024 x 60 s", ms, lengths.sum() / FS / 3600, 4 * lengths.sum() + 160 * fb.total_frames)
del out, arena
fb.close()
# ---- the repo's own call (load_data.py:64-70 with constants.py): 25 ms frames, NFFT 1024, lmfe only
arena, starts, lengths = corpus(4874, 128000)
fb = batch.FeatureBatch(starts, lengths, frame_length=0.025, frame_stride=0.01, num_filters=40, fft_length=1024)
out = torch.empty((fb.total_frames, 40), dtype=torch.float32, device=dev)
ms = timed(lambda: fb.lmfe_dev(arena.data_ptr(), arena.numel(), out.data_ptr()), fb.ctx)
assert bool(torch.isfinite(out).all())
report("repo call lmfe(25 ms, NFFT 1024), 4874 x 8 s", ms, lengths.sum() / FS / 3600, 4 * lengths.sum() + 160 * fb.total_frames)
