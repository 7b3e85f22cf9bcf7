"""CPU oracle for the scoring path of evaluation.py -- TEST INFRASTRUCTURE ONLY.

Follows ``/root/reference/evaluation.py``:

* cosine score        evaluation.py:76-77  (sklearn ``cosine_similarity`` on 1xd rows)
* argmax assignment   evaluation.py:79
* EER / AUC           evaluation.py:47-52  (``roc_curve`` + ``roc_auc_score`` + ``brentq``
                      on ``interp1d(fpr, tpr)``)
* k-fold averaging    evaluation.py:11-33

The arithmetic itself lives in third-party packages that are not vendored under
``/root/reference`` and are not pinned by it (no requirements file at the repo
root): scikit-learn (``sklearn.metrics.roc_curve``, ``roc_auc_score``,
``sklearn.metrics.pairwise.cosine_similarity``) and scipy
(``scipy.optimize.brentq``, ``scipy.interpolate.interp1d``).  The versions in
this image are scikit-learn 1.9.0 / scipy 1.18.1; ``get_eer_auc`` below calls
them exactly like the reference does, and ``*_restated`` re-derive the same
quantities in plain numpy from the published algorithms so that the two can be
cross-checked (tests/test_oracle.py) and so the algorithm the CUDA kernels
implement is written down in one place.

Parity status: pinned against ``evaluation.get_eer_auc`` imported from the
reference (with stub ``librosa`` / ``matplotlib`` modules) by
``oracle/make_golden.py``; goldens in ``tests/golden/scoring_*.npz``.
"""
from __future__ import annotations

import numpy as np


def cosine_scores(E, T):
    """S[i, j] = <E_i, T_j> / (|E_i| |T_j|), float64.

    sklearn ``cosine_similarity`` = ``normalize(X) @ normalize(Y).T`` with
    zero-norm rows left as zeros (evaluation.py:77 calls it one pair at a time).
    Inputs are promoted to float64 to serve as the parity oracle.
    """
    E = np.asarray(E, dtype=np.float64)
    T = np.asarray(T, dtype=np.float64)
    en = np.sqrt(np.einsum('ij,ij->i', E, E))
    tn = np.sqrt(np.einsum('ij,ij->i', T, T))
    en[en == 0.0] = 1.0
    tn[tn == 0.0] = 1.0
    return (E / en[:, None]) @ (T / tn[:, None]).T


def cosine_scores_sklearn(E, T):
    """The literal third-party call of evaluation.py:77, batched."""
    from sklearn.metrics.pairwise import cosine_similarity
    return cosine_similarity(E, T)


def assign_speaker(similarity_vec):
    """One-hot argmax row.  evaluation.py:73,79."""
    out = np.zeros(len(similarity_vec))
    out[np.argmax(similarity_vec)] = 1
    return out


def get_eer_auc(label, distance):
    """evaluation.py:47-52, verbatim call sequence into sklearn / scipy."""
    from scipy.interpolate import interp1d
    from scipy.optimize import brentq
    from sklearn.metrics import roc_auc_score, roc_curve
    fpr, tpr, _ = roc_curve(label, distance, pos_label=1)
    auc = roc_auc_score(label, distance)
    eer = brentq(lambda x: 1. - x - interp1d(fpr, tpr)(x), 0., 1.)
    return eer, auc, fpr, tpr


def k_fold_eer_auc(label, scores, k=1):
    """Mean EER / AUC over k contiguous chunks.  evaluation.py:11-33 (no plots)."""
    step = int(label.shape[0] / float(k))
    eers, aucs = [], []
    for s in range(k):
        e, a, _, _ = get_eer_auc(label[s * step:(s + 1) * step],
                                 scores[s * step:(s + 1) * step])
        eers.append(e)
        aucs.append(a)
    return float(np.mean(eers)), float(np.mean(aucs))


# ----------------------------------------------------------------------------
# numpy restatement of the published algorithms (what the CUDA kernels follow)
# ----------------------------------------------------------------------------
def roc_points_restated(label, score):
    """ROC vertices at distinct thresholds, with the (0, 0) origin prepended.

    sklearn ``_binary_clf_curve``: stable sort by descending score; cumulative
    true/false positive counts evaluated at the last element of every run of
    equal scores.  ``drop_intermediate`` only removes collinear vertices and
    does not change the piecewise-linear curve, so it is not applied here.
    Returns integer arrays (fps, tps) and the totals (n_neg, n_pos).
    """
    label = np.asarray(label)
    score = np.asarray(score)
    order = np.argsort(score, kind='mergesort')[::-1]
    s = score[order]
    y = (label[order] == 1).astype(np.int64)
    last_of_run = np.r_[np.nonzero(np.diff(s))[0], y.size - 1]
    tps = np.cumsum(y)[last_of_run]
    fps = 1 + last_of_run - tps
    tps = np.r_[0, tps]
    fps = np.r_[0, fps]
    return fps, tps, int(fps[-1]), int(tps[-1])


def roc_curve_restated(label, score):
    """``sklearn.metrics.roc_curve(label, score, pos_label=1)`` (evaluation.py:48) restated:
    vertices of ``_binary_clf_curve``; with more than two of them, ``drop_intermediate`` keeps
    the first, the last and every vertex where the second difference of fps or of tps is
    non-zero; then ``(0, 0, inf)`` is prepended.  Returns (fpr, tpr, thresholds) float64."""
    label = np.asarray(label)
    score = np.asarray(score)
    order = np.argsort(score, kind='mergesort')[::-1]
    s = score[order]
    y = (label[order] == 1).astype(np.int64)
    last_of_run = np.r_[np.nonzero(np.diff(s))[0], y.size - 1]
    tps = np.cumsum(y)[last_of_run]
    fps = 1 + last_of_run - tps
    thr = s[last_of_run].astype(np.float64)
    if fps.size > 2:
        keep = np.r_[True, np.logical_or(np.diff(fps, 2), np.diff(tps, 2)), True]
        fps, tps, thr = fps[keep], tps[keep], thr[keep]
    tps = np.r_[0, tps]
    fps = np.r_[0, fps]
    thr = np.r_[np.inf, thr]
    return fps / float(fps[-1]), tps / float(tps[-1]), thr


def eer_auc_restated(label, score):
    """EER = root of 1 - x - tpr(x) on the piecewise-linear ROC; AUC = trapezoid.

    g(v) = 1 - fpr_v - tpr_v is non-increasing along the vertices.  With v the
    first vertex where g <= 0 and u = v - 1: on a vertical segment
    (fpr_u == fpr_v) the root brentq converges to is that fpr; otherwise solve
    the linear equation on the segment.
    """
    fps, tps, n_neg, n_pos = roc_points_restated(label, score)
    fpr = fps / float(n_neg)
    tpr = tps / float(n_pos)
    g = 1.0 - fpr - tpr
    v = int(np.argmax(g <= 0))
    if g[v] == 0:
        eer = fpr[v]
    else:
        u = v - 1
        dx = fpr[v] - fpr[u]
        dy = tpr[v] - tpr[u]
        eer = fpr[u] if dx == 0 else fpr[u] + (1.0 - fpr[u] - tpr[u]) * dx / (dx + dy)
    auc = float(np.sum(np.diff(fpr) * (tpr[1:] + tpr[:-1]) * 0.5))
    return float(eer), auc


# ----------------------------------------------------------------------------
# deterministic synthetic scoring workload
# ----------------------------------------------------------------------------
def synth_embeddings(m, n, d, seed=5, noise=6.0):
    """E ~ N(0,1) (m x d); T_j = E_{j mod m} + noise*N(0,1) so that target
    trials (i == j mod m) score higher than impostors and the EER is neither 0
    nor 0.5.  float32, Philox-seeded so any host regenerates the same arrays."""
    rng = np.random.Generator(np.random.Philox(seed))
    E = rng.standard_normal((m, d)).astype(np.float32)
    T = (E[np.arange(n) % m] + noise * rng.standard_normal((n, d))).astype(np.float32)
    return E, T


def synth_trials(m, n, num_trials, seed=5):
    """num_trials (i, j, label) triples: half target (i == j mod m), half random."""
    rng = np.random.Generator(np.random.Philox(seed + 1000))
    j = rng.integers(0, n, size=num_trials)
    is_target = rng.random(num_trials) < 0.5
    i_rand = rng.integers(0, m, size=num_trials)
    i = np.where(is_target, j % m, i_rand)
    label = (i == (j % m)).astype(np.int32)
    return i.astype(np.int32), j.astype(np.int32), label
