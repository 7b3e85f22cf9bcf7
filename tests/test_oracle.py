"""The oracle against the committed reference outputs (no GPU, no /root/reference)."""
import numpy as np
import pytest

from conftest import golden
from oracle import scoring_oracle as so
from oracle import speechpy_oracle as o
from speaker_verification_b200 import synth

FS = 16000


def test_frame_geometry_table():
    g = golden("frame_geometry.npz")
    for L, S, n, f_nopad, f_pad in g["table"]:
        fl, fst = L / FS, S / FS
        for zp, want in ((False, f_nopad), (True, f_pad)):
            sig = np.arange(n, dtype=np.float64)
            if want < 0:
                with pytest.raises(ValueError):
                    o.stack_frames(sig, FS, fl, fst, zero_padding=zp)
            else:
                fr = o.stack_frames(sig, FS, fl, fst, zero_padding=zp)
                assert fr.shape == (want, L)
                if want:
                    # signal = arange  =>  frame values are the gather indices (bit-exact)
                    assert fr[-1, 0] == (want - 1) * S and fr[-1, -1] == (want - 1) * S + L - 1


def test_windowed_frames_and_preemphasis():
    g = golden("stack_frames_hamming.npz")
    sig = synth.audio(int(g["seed"]), int(g["n"]), np.float64)
    fr = o.stack_frames(sig, FS, 0.020, 0.010, filter=np.hamming, zero_padding=True)
    assert np.array_equal(fr, g["frames"])
    assert np.array_equal(o.preemphasis(sig, 1, 0.98), g["preemphasis"])


def test_filterbanks_bit_exact():
    g = golden("filterbanks.npz")
    for i in range(int(g["count"])):
        m, k, f, lo, hi = (int(v) for v in g[f"params_{i}"])
        lo = None if lo < 0 else lo
        hi = None if hi < 0 else hi
        assert np.array_equal(o.filterbank_edges(m, k, f, lo, hi), g[f"edges_{i}"])
        assert np.array_equal(o.filterbanks(m, k, f, lo, hi), g[f"bank_{i}"])
    # the edges SURVEY.md App. A.3 lists for the default configuration
    want = [4, 5, 6, 7, 8, 9, 10, 12, 13, 14, 16, 17, 19, 20, 22, 24, 26, 28, 30, 32, 35, 37, 40,
            42, 45, 49, 52, 55, 59, 63, 67, 71, 75, 80, 85, 90, 96, 102, 108, 114, 121, 128]
    assert list(o.filterbank_edges(40, 257, 16000, 0, None)) == want


@pytest.mark.parametrize("name,kw", [
    ("cfg1.npz", {}),
    ("repo_call_nfft1024.npz", dict(frame_length=0.025, frame_stride=0.01, num_filters=40, fft_length=1024)),
])
def test_feature_chain(name, kw):
    g = golden(name)
    x = synth.audio(int(g["seed"]), int(g["n"]), np.float32)
    lm = o.lmfe(x, FS, **kw)
    assert np.array_equal(lm, g["lmfe"])
    assert np.array_equal(o.mfe(x, FS, **kw)[1], g["energy"])
    c = o.mfcc(x, FS, num_cepstral=13, **kw)
    assert np.array_equal(c, g["mfcc"])
    assert np.array_equal(o.mfcc(x, FS, num_cepstral=13, dc_elimination=False, **kw), g["mfcc_nodc"])
    assert np.array_equal(o.extract_derivative_feature(c), g["cube"])
    assert np.array_equal(o.cmvnw(lm, 301, True), g["cmvnw"])
    assert np.array_equal(o.cmvnw(lm, 301, False), g["cmvnw_mean"])
    assert np.array_equal(o.cmvn(lm, True), g["cmvn"])
    assert np.array_equal(o.cmvn(lm, False), g["cmvn_mean"])
    assert np.max(np.abs(o.cmvnw_vectorised(lm, 301, True) - g["cmvnw"])) < 5e-5


def test_power_rows_cfg1():
    g = golden("cfg1.npz")
    x = synth.audio(1, 48000, np.float32).astype(np.float64)
    fr = o.stack_frames(x, FS, 0.020, 0.010, zero_padding=False)
    ps = o.power_spectrum(fr, 512)
    assert np.array_equal(ps[g["power_rows"]], g["power"])
    assert np.array_equal(o.log_power_spectrum(fr, 512, True)[[0, 297]], g["logpower_norm_rows"])


def test_int16_and_silence():
    g = golden("int16_input.npz")
    x = np.round(synth.audio(8, 16000, np.float64) * 32768 * 0.5).astype(np.int16)
    assert np.array_equal(o.lmfe(x, FS), g["lmfe"])
    assert np.array_equal(o.mfcc(x, FS), g["mfcc"])
    s = golden("silence.npz")
    assert np.array_equal(o.lmfe(np.zeros(4000), FS), s["lmfe_silence"])
    assert np.all(s["lmfe_silence"] == -36.04365338911715)
    half = synth.audio(9, 4000, np.float64)
    half[1000:2500] = 0.0
    assert np.array_equal(o.lmfe(half, FS), s["lmfe_half"])
    assert np.array_equal(o.mfcc(half, FS), s["mfcc_half"])


def test_empty_results():
    for n in (320, 400, 479):
        x = synth.audio(10, n, np.float64)
        assert o.lmfe(x, FS).shape == (0, 40)
        assert o.mfcc(x, FS).shape == (0, 13)
    with pytest.raises(ValueError):
        o.lmfe(synth.audio(10, 100, np.float64), FS)


def test_cmvnw_row_counts():
    g = golden("cmvnw_rows.npz")
    for rows in (1, 2, 5, 149, 150, 151, 299, 300, 301, 302, 400, 1000):
        v = g[f"in_rows{rows}"]
        for vn in (0, 1):
            assert np.array_equal(o.cmvnw(v, 301, bool(vn)), g[f"w301_rows{rows}_vn{vn}"])
        assert np.array_equal(o.cmvnw(v, 11, True), g[f"w11_rows{rows}_vn1"])
    with pytest.raises(AssertionError):
        o.cmvnw(g["in_rows5"], 300, True)
    with pytest.raises(AssertionError):
        o.cmvnw(g["in_rows5"], 301.0, True)


def test_derivative():
    g = golden("derivative.npz")
    assert np.array_equal(o.extract_derivative_feature(g["x"]), g["cube"])
    assert np.array_equal(o.extract_derivative_feature(g["x1"]), g["cube1"])
    x = g["x"]
    d = o.derivative_extraction(x, 2)
    c1 = np.minimum(np.arange(x.shape[1]) + 1, x.shape[1] - 1)
    c2 = np.minimum(np.arange(x.shape[1]) + 2, x.shape[1] - 1)
    assert np.allclose(d, (x[:, c1] + 2 * x[:, c2]) / 10)


def test_scoring_oracle():
    g = golden("scoring.npz")
    E, T = so.synth_embeddings(97, 113, 64, seed=5)
    assert np.max(np.abs(so.cosine_scores(E, T) - g["cos_small"])) < 1e-14
    for name in ("ties", "tiny", "perfect"):
        lab, s = g[f"label_{name}"], g[f"score_{name}"]
        eer, auc, _, _ = so.get_eer_auc(lab, s)
        assert eer == float(g[f"eer_{name}"]) and auc == float(g[f"auc_{name}"])
        e2, a2 = so.eer_auc_restated(lab, s)
        assert abs(e2 - eer) < 1e-9 and abs(a2 - auc) < 1e-12
        # ROC curve export: restatement == the curve the reference's get_eer_auc returned
        f, t, thr = so.roc_curve_restated(lab, s.astype(np.float32))
        assert np.array_equal(f, g[f"fpr_{name}"]) and np.array_equal(t, g[f"tpr_{name}"])
        assert np.isinf(thr[0]) and np.all(np.diff(thr) < 0)
    ti, tj, lab = so.synth_trials(4874, 4874, 37720, seed=5)
    s = g["cfg5_trial_scores_f64"]
    eer, auc, _, _ = so.get_eer_auc(lab, s)
    assert eer == float(g["eer_cfg5"]) and auc == float(g["auc_cfg5"])
    e2, a2 = so.eer_auc_restated(lab, s)
    assert abs(e2 - eer) < 1e-9 and abs(a2 - auc) < 1e-12
    f, t, _ = so.roc_curve_restated(lab, s.astype(np.float32))
    assert np.array_equal(f, g["fpr_cfg5"]) and np.array_equal(t, g["tpr_cfg5"])
    assert 0.005 < eer < 0.45  # the synthetic trial set is neither trivial nor random


def test_cmvn_transform_and_feature_cubes_oracle():
    """SURVEY 8f-1 oracle (utils.py:325-397) against the cubes the reference's transforms produced."""
    g = golden("feature_cubes.npz")
    from speaker_verification_b200 import synth
    lm = o.lmfe(synth.audio(int(g["seed"]), int(g["n"]), np.float64), 16000)
    for deriv in (False, True):
        for norm in (False, True):
            tag = f"d{int(deriv)}n{int(norm)}"
            cube = o.feature_cube(o.cmvn_transform(lm.copy(), deriv, norm), g[f"idx_{tag}"], 80)
            assert cube.shape == (3 if deriv else 1, 20, 80, 40) and cube.dtype == np.float32
            assert np.array_equal(cube[:, :3], g[f"cube_{tag}"])


def test_staged_reference_equals_oracle_port():
    """oracle/_ref (the reference's own files, staged by oracle/make_ref.py; what bench.py's CPU legs
    time as kind "reference") and the oracle port agree on the benchmarked chain."""
    import pytest
    from oracle import ref_loader
    ref = ref_loader.load()
    if ref is None:
        pytest.skip("oracle/_ref not staged (needs /root/reference: python oracle/make_ref.py)")
    from speaker_verification_b200 import synth
    x = synth.audio(1, 48000, np.float32)
    lm_ref = ref.feature.lmfe(x, 16000, frame_length=0.020, frame_stride=0.01, num_filters=40, fft_length=512)
    lm = o.lmfe(x, 16000)
    assert lm_ref.shape == lm.shape and np.max(np.abs(lm_ref - lm)) < 1e-12
    a = ref.processing.cmvnw(lm_ref, 301, True)
    b = o.cmvnw(lm, 301, True)
    assert a.dtype == b.dtype == np.float32 and np.array_equal(a, b)
