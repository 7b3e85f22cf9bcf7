"""Cosine trial scoring and EER/AUC -- the scoring half of the reference's evaluation.py.

Mirrors
  * ``sklearn.metrics.pairwise.cosine_similarity`` as called at evaluation.py:77
    (``cosine_similarity``), batched over all (enrolment, test) pairs
    (``cosine_scores``) instead of one 1x1 call per pair;
  * ``Evaluation.compute_Similarity`` (evaluation.py:67-84): ``SpeakerScorer``;
  * ``get_eer_auc`` (evaluation.py:47-52) and the k-fold averaging of
    ``get_and_plot_k_eer_auc`` (evaluation.py:11-33, plots left out).

Everything runs in libspeechpy_b200.so (tcgen05 3xTF32 GEMM, device sort + integer ROC
counts); there is no CPU fallback.  Multi-GPU (``score_trials_sharded``): rows of the
score matrix (enrolment embeddings) are sharded over the ranks, the test embeddings
are broadcast once from rank 0, every rank scores the trials whose row it owns and
the per-trial scores are gathered on rank 0, which computes EER/AUC.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib


def _f32c(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def cosine_scores(E, T, out=None, device=None):
    """S[i, j] = <E_i, T_j> / (|E_i| |T_j|) for all pairs; zero-norm rows give 0.

    E (m, d), T (n, d) array-likes -> float32 (m, n).  evaluation.py:76-77 over every
    (speaker model, utterance) pair in one call."""
    E, T = _f32c(E), _f32c(T)
    assert E.ndim == 2 and T.ndim == 2, "Expected 2D arrays"
    if E.shape[1] != T.shape[1]:
        raise ValueError(f"Incompatible dimension for X and Y matrices: X.shape[1] == {E.shape[1]} "
                         f"while Y.shape[1] == {T.shape[1]}")
    m, d = E.shape
    n = T.shape[0]
    if out is None:
        out = np.empty((m, n), dtype=np.float32)
    assert out.shape == (m, n) and out.dtype == np.float32 and out.flags.c_contiguous
    ctx = _lib.context(device)
    _lib.check(ctx.lib.spb_cosine_scores(ctx.handle, _lib.ptr(E), m, _lib.ptr(T), n, d, _lib.ptr(out)),
               ctx.handle)
    return out


def cosine_similarity(X, Y=None):
    """Drop-in for the sklearn call at evaluation.py:77 (any (m, d) x (n, d) shapes)."""
    X = np.atleast_2d(np.asarray(X))
    Y = X if Y is None else np.atleast_2d(np.asarray(Y))
    return cosine_scores(X, Y)


def _labels_scores(label, distance):
    label = np.asarray(label)
    distance = np.asarray(distance)
    if label.shape != distance.shape or label.ndim != 1:
        raise ValueError("label and distance must be 1-D arrays of the same length")
    lab = np.ascontiguousarray(label == 1, dtype=np.int32)  # pos_label=1 (evaluation.py:48)
    if lab.size == 0 or lab.min() == lab.max():
        raise ValueError("Only one class present in y_true. ROC AUC score is not defined in that case.")
    return lab, _f32c(distance)


def roc_curve(label, distance, device=None):
    """``sklearn.metrics.roc_curve(label, distance, pos_label=1)`` as called at
    evaluation.py:48: ``(fpr, tpr, thresholds)`` float64 arrays, collinear vertices dropped,
    first point ``(0, 0, inf)``.  Scores are ranked as float32 (the dtype the scorer emits)."""
    lab, s = _labels_scores(label, distance)
    ctx = _lib.context(device)
    cap = lab.size + 1
    fpr, tpr, thr = (np.empty(cap, dtype=np.float64) for _ in range(3))
    count = C.c_int64()
    _lib.check(ctx.lib.spb_roc_curve(ctx.handle, _lib.ptr(s), _lib.ptr(lab), lab.size, _lib.ptr(fpr),
                                     _lib.ptr(tpr), _lib.ptr(thr), cap, C.byref(count)), ctx.handle)
    k = count.value
    return fpr[:k].copy(), tpr[:k].copy(), thr[:k].copy()


def get_eer_auc(label, distance, device=None, curve=True):
    """evaluation.py:47-52: returns ``(eer, auc, fpr, tpr)`` like the reference.  ``curve=False``
    skips the ROC export (the reference only plots it) and returns ``None`` for fpr / tpr."""
    lab, s = _labels_scores(label, distance)
    ctx = _lib.context(device)
    eer, auc = C.c_double(), C.c_double()
    _lib.check(ctx.lib.spb_eer(ctx.handle, _lib.ptr(s), _lib.ptr(lab), lab.size, C.byref(eer), C.byref(auc)),
               ctx.handle)
    if not curve:
        return eer.value, auc.value, None, None
    fpr, tpr, _ = roc_curve(lab, s, device=device)
    return eer.value, auc.value, fpr, tpr


def k_fold_eer_auc(label, scores, k=1, verbose=True, device=None):
    """Mean EER / AUC over ``k`` contiguous chunks: ``get_and_plot_k_eer_auc``
    (evaluation.py:11-33) without the matplotlib part; prints the same two lines."""
    label = np.asarray(label)
    scores = np.asarray(scores)
    step = int(label.shape[0] / float(k))
    eers, aucs = np.zeros((k, 1)), np.zeros((k, 1))
    for split_num in range(k):
        a, b = split_num * step, (split_num + 1) * step
        eers[split_num], aucs[split_num], _, _ = get_eer_auc(label[a:b], scores[a:b], device=device, curve=False)
    if verbose:
        print("EER=", np.mean(eers) * 100)
        print("AUC=", np.mean(aucs) * 100)
    return float(np.mean(eers)), float(np.mean(aucs))


class SpeakerScorer:
    """``Evaluation`` (evaluation.py:55-84) without the model: enrolled speaker
    embeddings in, similarity vectors and argmax assignment out."""

    def __init__(self, speaker_models, device=None):
        # dict name -> (1, d) or (d,) embedding, in insertion order like self.speaker_models
        self.names = list(speaker_models.keys())
        self.E = _f32c(np.stack([np.asarray(speaker_models[k]).reshape(-1) for k in self.names]))
        self.device = device

    def compute_similarity(self, speaker_features):
        """One utterance embedding (1, d) -> (similarity_vec, assigned_speaker_vec), both
        float64 of length n_speakers, as compute_Similarity returns them."""
        sim, assigned = self.compute_similarity_batch(np.asarray(speaker_features).reshape(1, -1))
        return sim[0], assigned[0]

    compute_Similarity = compute_similarity  # the reference's spelling

    def compute_similarity_batch(self, features):
        """(u, d) utterance embeddings -> (u, n_speakers) similarities and one-hot rows."""
        S = cosine_scores(features, self.E, device=self.device).astype(np.float64)
        assigned = np.zeros_like(S)
        assigned[np.arange(S.shape[0]), np.argmax(S, axis=1)] = 1
        return S, assigned


# ------------------------------------------------------------------------------------
# device-resident API (torch tensors on the context's device)
# ------------------------------------------------------------------------------------
def _check_cuda_f32(t, name):
    assert t.is_cuda and t.dtype.is_floating_point and t.element_size() == 4 and t.is_contiguous(), \
        f"{name} must be a contiguous float32 CUDA tensor"


def cosine_scores_dev(E, T, out=None):
    """torch CUDA tensors in/out; asynchronous GEMM on the context's stream, then synced."""
    import torch
    _check_cuda_f32(E, "E")
    _check_cuda_f32(T, "T")
    m, d = E.shape
    n = T.shape[0]
    if out is None:
        out = torch.empty((m, n), dtype=torch.float32, device=E.device)
    ctx = _lib.context(E.device.index)
    torch.cuda.current_stream(E.device).synchronize()
    _lib.check(ctx.lib.spb_cosine_scores_dev(ctx.handle, E.data_ptr(), m, T.data_ptr(), n, d, out.data_ptr()),
               ctx.handle)
    ctx.sync()
    return out


def trial_scores_dev(S, trial_i, trial_j):
    """out[t] = S[trial_i[t], trial_j[t]] on the device (int32 index tensors)."""
    import torch
    _check_cuda_f32(S, "S")
    assert trial_i.dtype == torch.int32 and trial_j.dtype == torch.int32
    out = torch.empty(trial_i.numel(), dtype=torch.float32, device=S.device)
    ctx = _lib.context(S.device.index)
    torch.cuda.current_stream(S.device).synchronize()
    _lib.check(ctx.lib.spb_trial_scores_dev(ctx.handle, S.data_ptr(), S.shape[0], S.shape[1],
                                            trial_i.data_ptr(), trial_j.data_ptr(), trial_i.numel(),
                                            out.data_ptr()), ctx.handle)
    ctx.sync()
    return out


def argmax_rows_dev(S):
    import torch
    _check_cuda_f32(S, "S")
    out = torch.empty(S.shape[0], dtype=torch.int32, device=S.device)
    ctx = _lib.context(S.device.index)
    torch.cuda.current_stream(S.device).synchronize()
    _lib.check(ctx.lib.spb_argmax_rows_dev(ctx.handle, S.data_ptr(), S.shape[0], S.shape[1], out.data_ptr()),
               ctx.handle)
    ctx.sync()
    return out


def eer_auc_dev(scores, labels):
    """scores float32, labels int32 (0/1) CUDA tensors -> (eer, auc) Python floats."""
    import torch
    _check_cuda_f32(scores, "scores")
    assert labels.dtype == torch.int32 and labels.is_cuda and labels.is_contiguous()
    ctx = _lib.context(scores.device.index)
    torch.cuda.current_stream(scores.device).synchronize()
    eer, auc = C.c_double(), C.c_double()
    _lib.check(ctx.lib.spb_eer_dev(ctx.handle, scores.data_ptr(), labels.data_ptr(), scores.numel(),
                                   C.byref(eer), C.byref(auc)), ctx.handle)
    return eer.value, auc.value


# ------------------------------------------------------------------------------------
# multi-GPU: rows of the score matrix sharded over ranks
# ------------------------------------------------------------------------------------
def row_partition(m: int, world_size: int):
    """Contiguous row blocks [bounds[r], bounds[r+1]) of near-equal size (609/610 rows at
    m = 4874, world_size = 8)."""
    return [(m * r) // world_size for r in range(world_size + 1)]


def _cuda_block_scores(E_block, T, ti_local, tj):
    S = cosine_scores_dev(E_block, T)
    return trial_scores_dev(S, ti_local, tj), S


class ShardedTrials:
    """The per-rank part of a trial list for row-sharded scoring, computed once: which trials this
    rank's row block [row0, row0 + rows) owns and their local indices.  ``score`` is the per-call
    work: [broadcast of T] -> normalise + GEMM of the block -> gather -> reduce to rank 0 -> EER."""

    def __init__(self, row0, rows, trial_i, trial_j, labels, group=None):
        import torch
        self.row0, self.rows, self.group = int(row0), int(rows), group
        own = (trial_i >= row0) & (trial_i < row0 + rows)
        self.idx = torch.nonzero(own).flatten()
        self.ti_local = (trial_i[self.idx] - row0).to(torch.int32).contiguous()
        self.tj_local = trial_j[self.idx].to(torch.int32).contiguous()
        self.num_trials = int(trial_i.numel())
        self.labels = labels.to(torch.int32).contiguous()

    def score(self, E_block, T, t_resident=False, block_scores=None):
        import torch
        import torch.distributed as dist
        group = self.group
        on = dist.is_initialized()
        rank = dist.get_rank(group) if on else 0
        world = dist.get_world_size(group) if on else 1
        root = dist.get_global_rank(group, 0) if (on and group is not None) else 0
        if world > 1 and not t_resident:
            dist.broadcast(T, src=root, group=group)
        full = torch.zeros(self.num_trials, dtype=torch.float32, device=E_block.device)
        if self.rows > 0 and self.idx.numel() > 0:
            fn = block_scores or _cuda_block_scores
            picked, _ = fn(E_block, T, self.ti_local, self.tj_local)
            full[self.idx] = picked
        if world > 1:
            dist.reduce(full, dst=root, op=dist.ReduceOp.SUM, group=group)
        if rank != 0:
            return None, None, None
        if block_scores is not None and not full.is_cuda:
            return None, None, full  # CPU test of the exchange: the caller checks the scores
        eer, auc = eer_auc_dev(full, self.labels)
        return eer, auc, full


def score_trials_sharded(E_block, row0, m, T, trial_i, trial_j, labels, group=None, block_scores=None,
                         t_resident=False):
    """Score a trial list with the enrolment rows sharded over the ranks of ``group``.

    E_block : this rank's rows [row0, row0 + E_block.shape[0]) of the (m, d) enrolment matrix
    T       : (n, d) test embeddings; only rank 0's contents are used (broadcast to the rest)
    trial_i, trial_j, labels : int32 tensors, identical on every rank
    Returns ``(eer, auc, scores)`` on rank 0 (``scores`` = per-trial cosine, float32) and
    ``(None, None, None)`` elsewhere.

    Exchange steps: ONE broadcast of T (NCCL over NVLink on GPUs) and one small
    reduce-to-rank-0 of the per-trial score vector, in which every trial is non-zero on
    exactly the rank that owns its row (x + 0 is exact, so the reduction is a gather).
    ``t_resident=True`` skips the broadcast: every rank already holds T (broadcast once, outside).
    ``block_scores`` replaces the CUDA block computation in CPU (gloo) tests of this
    exchange logic; the product path never passes it.
    """
    plan = ShardedTrials(row0, E_block.shape[0], trial_i, trial_j, labels, group)
    return plan.score(E_block, T, t_resident=t_resident, block_scores=block_scores)
