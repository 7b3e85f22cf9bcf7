"""GPU check of the tensor-core feature kernel: lmfe through the tensor-core path vs the CUDA-core
FFT kernel (debug bit 2) vs the float64 oracle, plus per-kernel timing on a cfg2-sized batch."""
import ctypes
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from oracle import speechpy_oracle as o  # noqa: E402
from speaker_verification_b200 import _lib, batch, synth  # noqa: E402

raw = ctypes.CDLL(_lib.LIB_PATH)
FS = 16000


def run(fb, arena_t, tc):
    out = torch.empty((fb.total_frames, 40), dtype=torch.float32, device="cuda")
    raw.spb_debug_force_generic(0 if tc else 4)
    fb.lmfe_dev(arena_t.data_ptr(), arena_t.numel(), out.data_ptr())
    fb.ctx.sync()
    raw.spb_debug_force_generic(0)
    return out.cpu().numpy()


lengths = [48000, 20000, 64321, 5280, 5281, 800, 480, 10240, 320000, 100321]
sigs = [synth.audio(900 + i, n, np.float32) for i, n in enumerate(lengths)]
arena, starts, lens = batch.pack(sigs)
fb = batch.FeatureBatch(starts, lens)
arena_t = torch.from_numpy(arena).cuda()
a = run(fb, arena_t, True)
b = run(fb, arena_t, False)
print("frames", fb.total_frames, "tc vs fft max|d|", float(np.abs(a - b).max()), "nan", int(np.isnan(a).sum()))
worst = 0.0
for s, got, got_b in zip(sigs, fb.split(a), fb.split(b)):
    ref = o.lmfe(s, FS)
    if ref.shape[0] == 0:
        continue
    e = float(np.abs(got - ref).max())
    eb = float(np.abs(got_b - ref).max())
    worst = max(worst, e)
    print(f"  N={len(s):7d} F={ref.shape[0]:5d} tc err {e:.3e}  fft err {eb:.3e}")
    if e > 1e-3:
        d = np.abs(got - ref)
        r, c = np.unravel_index(np.argmax(d), d.shape)
        print("   worst at row", r, "col", c, "got", got[r, c], "ref", ref[r, c])
        print("   per-col max err", np.round(d.max(0), 3))
        print("   rows with err>1e-3:", np.where(d.max(1) > 1e-3)[0][:40])
print("worst tc err", worst)
# fallback: a loud utterance (|x| >= 2) must come out right through the guarded FFT launch
loud = [s * np.float32(50.0) for s in sigs[:3]]
ar2, st2, ln2 = batch.pack(loud)
fb2 = batch.FeatureBatch(st2, ln2)
g = run(fb2, torch.from_numpy(ar2).cuda(), True)
w2 = max(float(np.abs(x - o.lmfe(s, FS)).max()) for s, x in zip(loud, fb2.split(g)))
print("loud (fallback) err", w2)
# timing: 4874 x 8 s
rng = np.random.default_rng(0)
U, N = 4874, 128000
big = torch.from_numpy((np.round(rng.normal(0, 0.1, N * 64) * 32768) / 32768).astype(np.float32)).cuda().repeat(U // 64 + 1)[: U * N]
fb3 = batch.FeatureBatch(np.arange(U, dtype=np.int64) * N, np.full(U, N, dtype=np.int64))
out = torch.empty((fb3.total_frames, 40), dtype=torch.float32, device="cuda")
for tc in (True, False):
    raw.spb_debug_force_generic(0 if tc else 4)
    for _ in range(3):
        fb3.lmfe_dev(big.data_ptr(), big.numel(), out.data_ptr())
    fb3.ctx.sync()
    t0 = time.perf_counter()
    for _ in range(10):
        fb3.lmfe_dev(big.data_ptr(), big.numel(), out.data_ptr())
    fb3.ctx.sync()
    dt = (time.perf_counter() - t0) / 10
    print(f"lmfe {'tc ' if tc else 'fft'}: {dt*1e3:.3f} ms per {U*N/FS/3600:.2f} audio-h -> {U*N/FS/3600/dt:.0f} audio-h/s")
raw.spb_debug_force_generic(0)
