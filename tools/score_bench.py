"""GPU: time the cosine GEMM (kernel timing slot 3) at config 5 and report error statistics."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from oracle import scoring_oracle as so
from speaker_verification_b200 import scoring, _lib

m = n = 4874; d = 512
E, T = so.synth_embeddings(m, n, d, seed=5)
dE, dT = torch.from_numpy(E).cuda(), torch.from_numpy(T).cuda()
S = torch.empty((m, n), dtype=torch.float32, device="cuda")
ctx = _lib.context(0)
flops = 2.0 * m * n * d
K = 20
for rep in range(2):
    for mc in (0, 1):
        with _lib.options(gemm_multicast=mc):
            for _ in range(3):
                scoring.cosine_scores_dev(dE, dT, S)
            ctx.timing_enable(True)
            for _ in range(K):
                scoring.cosine_scores_dev(dE, dT, S)
            ms, cnt = ctx.timing_read(3)
            ctx.timing_enable(False)
        print(f"gemm_multicast={mc}: {ms / cnt * 1e3:.1f} us/launch", flush=True)
for _ in range(3):
    scoring.cosine_scores_dev(dE, dT, S)
ctx.timing_enable(True)
for _ in range(K):
    scoring.cosine_scores_dev(dE, dT, S)
ms, cnt = ctx.timing_read(3)
ctx.timing_enable(False)
per = ms / cnt
print(f"cosine_gemm_kernel: {per*1e3:.1f} us/launch over {cnt} launches; useful {flops/per/1e9:.1f} TFLOP/s "
      f"(3xTF32 issued {3*flops/per/1e9:.1f} TFLOP/s); output write {m*n*4/per/1e6:.1f} GB/s")
rows = np.r_[0:256, 4700:4874]
ref = so.cosine_scores(E[rows], T)
got = S[torch.from_numpy(rows).cuda()].cpu().numpy()
err = np.abs(got - ref)
for lo, hi in [(0, 0.02), (0.02, 0.05), (0.05, 0.1), (0.1, 1.0)]:
    sel = (np.abs(ref) >= lo) & (np.abs(ref) < hi)
    if sel.any():
        print(f"|ref| in [{lo},{hi}): n={sel.sum()} max abs err {err[sel].max():.3e} rms {np.sqrt((err[sel]**2).mean()):.3e} "
              f"mean signed*sign(ref) {((got-ref)*np.sign(ref))[sel].mean():.3e}")
# reference-style fp32 sklearn for context
from sklearn.metrics.pairwise import cosine_similarity
sk = cosine_similarity(E[rows], T)
print("sklearn fp32 vs f64 oracle: max abs err %.3e" % np.abs(sk - ref).max())
