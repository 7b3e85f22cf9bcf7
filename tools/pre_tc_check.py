import sys; sys.path.insert(0, ".")
import numpy as np, torch
from oracle import speechpy_oracle as o
from speaker_verification_b200 import _lib, batch, synth
lengths = synth.cfg3_lengths(96, seed=41)
arena, starts, lens = synth.ragged_corpus(lengths, seed=42)
for u in range(len(starts)): arena[starts[u]+lens[u]-1] = 0.7
fb = batch.FeatureBatch(starts, lens, preemphasis=0.97)
d = torch.from_numpy(arena).cuda()
out = torch.full((fb.total_frames, 40), float("nan"), device="cuda")
def run(**opts):
    with _lib.options(**opts):
        n0 = fb.ctx.launch_count(); fb.lmfe_dev(d.data_ptr(), d.numel(), out.data_ptr()); fb.ctx.sync()
        return out.cpu().numpy().copy(), fb.ctx.launch_count() - n0
tc, l1 = run(); fft, l2 = run(feature_kernel=2); raw, _ = run(redo=0)
ref = np.concatenate([o.lmfe(o.preemphasis(arena[starts[u]:starts[u]+lens[u]].astype(np.float64), 1, 0.97), 16000) for u in range(len(starts))])
print("launches", l1, l2, "tc err", np.abs(tc-ref).max(), "fft err", np.abs(fft-ref).max(), "raw tc err", np.abs(raw-ref).max(), "redone", (raw != tc).mean())
e = np.abs(tc-ref); print("argmax", np.unravel_index(e.argmax(), e.shape), "q9999", np.quantile(e, 0.9999))
