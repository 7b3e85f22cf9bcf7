"""GPU debug: error statistics of the tcgen05 cosine GEMM against the float64 oracle."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
from oracle import scoring_oracle as so
from speaker_verification_b200 import scoring

for (m, n, d) in [(97, 113, 64), (128, 256, 32), (130, 300, 100), (1, 5, 128), (1000, 1200, 512)]:
    E, T = so.synth_embeddings(m, n, d, seed=5)
    S = scoring.cosine_scores(E, T)
    ref = so.cosine_scores(E, T)
    err = np.abs(S - ref)
    bound = 1e-5 * np.maximum(np.abs(ref), 0.02)
    print(m, n, d, "max abs err %.3e" % err.max(), "max err/bound %.3f" % (err / bound).max(),
          "mean signed err %.3e" % (S - ref).mean(), "nan", int(np.isnan(S).sum()), flush=True)
    if err.max() > 1e-3:
        bad = np.argwhere(err > 1e-3)
        print("  bad count", len(bad), "first", bad[:5].tolist(), "rows bad", np.unique(bad[:, 0])[:10], "cols bad", np.unique(bad[:, 1])[:10])
