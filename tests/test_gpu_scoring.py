"""Scoring path on the GPU (tcgen05 cosine GEMM, trial gather, argmax, EER/AUC) against the
oracle and the goldens generated from the reference's evaluation.get_eer_auc.

Tolerances (BASELINE.json north_star): cosine scores within 1e-5 relative -- written as
|got - ref| <= 1e-5 * max(|ref|, FLOOR) with FLOOR = 0.02 because scores of random
embeddings cluster around 0 (sigma ~ 0.044 at d = 512) where a pure relative error is
ill-conditioned -- and EER identical to 4 decimals.
"""
import numpy as np
import pytest

from conftest import golden
from oracle import scoring_oracle as so
from speaker_verification_b200 import scoring

pytestmark = pytest.mark.gpu
REL, FLOOR = 1e-5, 0.02


def assert_scores(got, ref):
    assert got.shape == ref.shape and got.dtype == np.float32
    bound = REL * np.maximum(np.abs(ref), FLOOR)
    err = np.abs(got.astype(np.float64) - ref)
    assert np.all(err <= bound), (float(err.max()), float((err / bound).max()))


def test_cosine_small_against_reference_golden(lib_built):
    g = golden("scoring.npz")
    E, T = so.synth_embeddings(97, 113, 64, seed=5)
    assert_scores(scoring.cosine_scores(E, T), g["cos_small"])


@pytest.mark.parametrize("m,n,d", [(1, 5, 128), (5, 1, 128), (130, 300, 100), (257, 513, 36), (128, 256, 512),
                                   (129, 257, 32), (3, 2, 1)])
def test_cosine_shapes(lib_built, m, n, d):
    rng = np.random.Generator(np.random.Philox(m * 1000 + n * 10 + d))
    E = rng.standard_normal((m, d)).astype(np.float32)
    T = rng.standard_normal((n, d)).astype(np.float32)
    assert_scores(scoring.cosine_scores(E, T), so.cosine_scores(E, T))


def test_cosine_cluster_multicast_equals_plain_kernel(lib_built):
    """The default 2-CTA cluster kernel (T tiles multicast, each CTA loads half of them) and the
    one-CTA-per-tile kernel (option gemm_multicast=0) run the same MMAs in the same order: bit-identical scores, odd numbers of row tiles included
    (the phantom tile of the last pair)."""
    from speaker_verification_b200 import _lib
    rng = np.random.Generator(np.random.Philox(4242))
    for m, n, d in [(4874, 1200, 512), (129, 700, 64), (640, 333, 100), (256, 256, 16), (1000, 4874, 512)]:
        E = rng.standard_normal((m, d)).astype(np.float32)
        T = rng.standard_normal((n, d)).astype(np.float32)
        a = scoring.cosine_scores(E, T)
        with _lib.options(gemm_multicast=0):
            b = scoring.cosine_scores(E, T)
        assert np.array_equal(a, b), (m, n, d)
        assert_scores(a[:64], so.cosine_scores(E[:64], T))
        assert_scores(a[-64:], so.cosine_scores(E[-64:], T))


def test_cosine_zero_rows_and_scale(lib_built):
    rng = np.random.Generator(np.random.Philox(77))
    E = rng.standard_normal((40, 128)).astype(np.float32)
    T = rng.standard_normal((50, 128)).astype(np.float32)
    E[3] = 0.0
    T[7] = 0.0
    E[5] *= 1e-12   # tiny norm, still a direction
    T[9] *= 1e12
    got = scoring.cosine_scores(E, T)
    assert np.all(got[3] == 0.0) and np.all(got[:, 7] == 0.0)  # sklearn: zero rows stay zero
    assert_scores(got, so.cosine_scores(E, T))
    assert np.max(np.abs(got)) <= 1.0 + 1e-6
    same = scoring.cosine_similarity(E[:4])  # Y=None: self-similarity
    assert np.allclose(np.diag(same)[[0, 1, 2]], 1.0, atol=2e-6) and same[3, 3] == 0.0


def test_empty_and_mismatched(lib_built):
    assert scoring.cosine_scores(np.zeros((0, 16), np.float32), np.ones((3, 16), np.float32)).shape == (0, 3)
    with pytest.raises(ValueError):
        scoring.cosine_scores(np.ones((2, 16), np.float32), np.ones((3, 8), np.float32))


@pytest.mark.parametrize("name", ["ties", "tiny", "perfect"])
def test_eer_auc_goldens(lib_built, name):
    g = golden("scoring.npz")
    lab, s = g[f"label_{name}"], g[f"score_{name}"]
    eer, auc, _, _ = scoring.get_eer_auc(lab, s)
    assert abs(eer - float(g[f"eer_{name}"])) < 1e-9, (eer, float(g[f"eer_{name}"]))
    assert abs(auc - float(g[f"auc_{name}"])) < 1e-12


@pytest.mark.parametrize("name", ["ties", "tiny", "perfect"])
def test_roc_curve_goldens(lib_built, name):
    """get_eer_auc returns (eer, auc, fpr, tpr) like evaluation.py:47-52; the curve must be the
    reference's point for point (integer counts -> bit-exact float64 ratios)."""
    g = golden("scoring.npz")
    lab, s = g[f"label_{name}"], g[f"score_{name}"]
    _, _, fpr, tpr = scoring.get_eer_auc(lab, s)
    assert fpr.dtype == np.float64 and np.array_equal(fpr, g[f"fpr_{name}"]) and np.array_equal(tpr, g[f"tpr_{name}"])
    f2, t2, thr = scoring.roc_curve(lab, s)
    fo, to, tho = so.roc_curve_restated(lab, s.astype(np.float32))
    assert np.array_equal(f2, fo) and np.array_equal(t2, to) and np.array_equal(thr, tho)


def test_roc_curve_large_with_ties_and_cfg5(lib_built):
    from sklearn.metrics import roc_curve as sk_roc_curve
    g = golden("scoring.npz")
    rng = np.random.Generator(np.random.Philox(9))
    s = np.round(rng.standard_normal(200_000), 2).astype(np.float32)
    lab = (rng.random(200_000) < 0.1 + 0.3 * (s > 0.3)).astype(np.int32)
    fpr, tpr, thr = scoring.roc_curve(lab, s)
    f_sk, t_sk, th_sk = sk_roc_curve(lab, s, pos_label=1)
    assert np.array_equal(fpr, f_sk) and np.array_equal(tpr, t_sk)
    assert np.isinf(thr[0]) and np.array_equal(thr[1:], th_sk[1:].astype(np.float64))
    # config 5 trial scores (float32 of the float64 golden scores)
    ti, tj, lab5 = so.synth_trials(4874, 4874, 37720, seed=5)
    f5, t5, _ = scoring.roc_curve(lab5, g["cfg5_trial_scores_f64"].astype(np.float32))
    assert np.array_equal(f5, g["fpr_cfg5"]) and np.array_equal(t5, g["tpr_cfg5"])
    # two points only / all scores equal
    f, t, th = scoring.roc_curve(np.array([1, 0, 1, 0, 0]), np.zeros(5))
    assert f.tolist() == [0.0, 1.0] and t.tolist() == [0.0, 1.0] and th[1] == 0.0


def test_eer_edge_cases(lib_built):
    with pytest.raises(ValueError):
        scoring.get_eer_auc(np.ones(5, np.int32), np.arange(5.0))
    # all scores equal: one ROC segment (0,0)-(1,1): EER 0.5, AUC 0.5
    eer, auc, _, _ = scoring.get_eer_auc(np.array([1, 0, 1, 0, 0]), np.zeros(5))
    assert abs(eer - 0.5) < 1e-12 and abs(auc - 0.5) < 1e-12
    # random large case with many ties vs the sklearn/scipy oracle
    rng = np.random.Generator(np.random.Philox(9))
    s = np.round(rng.standard_normal(200_000), 2).astype(np.float32)
    lab = (rng.random(200_000) < 0.1 + 0.3 * (s > 0.3)).astype(np.int32)
    eer, auc, _, _ = scoring.get_eer_auc(lab, s)
    e_ref, a_ref, _, _ = so.get_eer_auc(lab, s.astype(np.float64))
    assert abs(eer - e_ref) < 1e-9 and abs(auc - a_ref) < 1e-12
    # k-fold averaging (evaluation.py:11-33)
    ek, ak = scoring.k_fold_eer_auc(lab, s, k=4, verbose=False)
    e4, a4 = so.k_fold_eer_auc(lab, s.astype(np.float64), k=4)
    assert abs(ek - e4) < 1e-9 and abs(ak - a4) < 1e-12


def test_cfg5_full_matrix_trials_and_eer(lib_built):
    """BASELINE config 5: 4874 x 4874 x 512 score matrix, 37,720 trials."""
    g = golden("scoring.npz")
    E, T = so.synth_embeddings(4874, 4874, 512, seed=5)
    ti, tj, lab = so.synth_trials(4874, 4874, 37720, seed=5)
    S = scoring.cosine_scores(E, T)
    ref_trials = g["cfg5_trial_scores_f64"]
    assert_scores(S[ti, tj], ref_trials)
    # a full block of rows (covers partial tiles on both edges) against the float64 oracle
    rows = np.r_[0:130, 2400:2530, 4740:4874]
    assert_scores(S[rows], so.cosine_scores(E[rows], T))
    eer, auc, _, _ = scoring.get_eer_auc(lab, S[ti, tj])
    assert round(eer, 4) == round(float(g["eer_cfg5"]), 4), (eer, float(g["eer_cfg5"]))
    assert abs(auc - float(g["auc_cfg5"])) < 1e-5
    # oracle EER of the GPU scores (separates GEMM error from EER error)
    e_or, _, _, _ = so.get_eer_auc(lab, S[ti, tj].astype(np.float64))
    assert abs(eer - e_or) < 1e-9


def test_device_api_gather_argmax_and_sharded_single_rank(lib_built):
    import torch
    E, T = so.synth_embeddings(300, 421, 128, seed=11)
    ti, tj, lab = so.synth_trials(300, 421, 5000, seed=11)
    dE, dT = torch.from_numpy(E).cuda(), torch.from_numpy(T).cuda()
    S = scoring.cosine_scores_dev(dE, dT)
    ref = so.cosine_scores(E, T)
    assert_scores(S.cpu().numpy(), ref)
    picked = scoring.trial_scores_dev(S, torch.from_numpy(ti).cuda(), torch.from_numpy(tj).cuda())
    assert np.array_equal(picked.cpu().numpy(), S.cpu().numpy()[ti, tj])
    am = scoring.argmax_rows_dev(S).cpu().numpy()
    assert np.array_equal(am, np.argmax(S.cpu().numpy(), axis=1))
    with pytest.raises(AssertionError):
        scoring.trial_scores_dev(S, torch.tensor([300], dtype=torch.int32).cuda(),
                                 torch.tensor([0], dtype=torch.int32).cuda())
    eer, auc, full = scoring.score_trials_sharded(dE, 0, 300, dT, torch.from_numpy(ti).cuda(),
                                                  torch.from_numpy(tj).cuda(), torch.from_numpy(lab).cuda())
    e_ref, a_ref, _, _ = so.get_eer_auc(lab, full.cpu().numpy().astype(np.float64))
    assert abs(eer - e_ref) < 1e-9 and abs(auc - a_ref) < 1e-12


def test_speaker_scorer_matches_compute_similarity(lib_built):
    """Evaluation.compute_Similarity (evaluation.py:67-84) for d = 128 speaker models."""
    rng = np.random.Generator(np.random.Philox(21))
    models = {f"id1{k:04d}": rng.standard_normal((1, 128)).astype(np.float32) for k in range(40)}
    scorer = scoring.SpeakerScorer(models)
    utt = (models["id10007"] + 0.5 * rng.standard_normal((1, 128))).astype(np.float32)
    sim, assigned = scorer.compute_Similarity(utt)
    ref = so.cosine_scores(np.concatenate(list(models.values())), utt)[:, 0]
    assert sim.shape == (40,) and np.max(np.abs(sim - ref)) < 1e-6
    assert np.array_equal(assigned, so.assign_speaker(ref)) and assigned[7] == 1
