"""Where do the tensor-core and the FFT lmfe kernels differ most on a many-tile batch?"""
import ctypes
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from oracle import speechpy_oracle as o  # noqa: E402
from speaker_verification_b200 import _lib, batch, synth  # noqa: E402

lengths = synth.cfg3_lengths(128, seed=21)
arena, starts, lens = synth.ragged_corpus(lengths, seed=22)
fb = batch.FeatureBatch(starts, lens)
d_in = torch.from_numpy(arena).cuda()


def run(tc):
    out = torch.empty((fb.total_frames, 40), dtype=torch.float32, device="cuda")
    with _lib.options(feature_kernel=0 if tc else 2):
        fb.lmfe_dev(d_in.data_ptr(), d_in.numel(), out.data_ptr())
        fb.ctx.sync()
    return out.cpu().numpy()


tc, tc2, fft = run(True), run(True), run(False)
print("tc run-to-run identical:", np.array_equal(tc, tc2))
ref = np.concatenate([o.lmfe(arena[starts[u]:starts[u] + lens[u]], 16000) for u in range(len(starts))])
et, ef = np.abs(tc - ref), np.abs(fft - ref)
print(f"frames {fb.total_frames}: tc max {et.max():.3e} p99.99 {np.quantile(et, 0.9999):.3e}; fft max {ef.max():.3e} p99.99 {np.quantile(ef, 0.9999):.3e}")
print("count tc err > 5e-5:", int((et > 5e-5).sum()), " fft:", int((ef > 5e-5).sum()))
idx = np.argsort(et.ravel())[::-1][:12]
for i in idx:
    r, c = divmod(int(i), 40)
    u = int(np.searchsorted(fb.frame_offsets, r, side="right") - 1)
    f = r - int(fb.frame_offsets[u])
    print(f"  row {r} utt {u} frame {f} (f%31={f % 31}) col {c}: tc {tc[r, c]:.6f} fft {fft[r, c]:.6f} ref {ref[r, c]:.6f}  ref row min {ref[r].min():.3f}")
