#!/usr/bin/env python
"""Pin the oracle against the real reference and write the golden vectors.

Run in the build container (where ``/root/reference`` exists):

    python oracle/make_golden.py

It imports the UNMODIFIED reference (read-only, no bytecode written):

* ``speech_feature_extraction.speechpy.{processing,feature,functions}`` -- needs
  the one-line shim ``np.lib.pad = np.pad`` because numpy >= 2 dropped
  ``np.lib.pad`` (sfe/speechpy/processing.py:223,301,314);
* ``evaluation.get_eer_auc`` -- ``evaluation.py`` star-imports ``utils`` /
  ``model`` which need ``librosa`` and ``matplotlib`` (absent here); empty stub
  modules are placed in ``sys.modules`` first.

For every function on the hot path it (1) runs reference and oracle on the same
seeded inputs and asserts agreement (bit-exact for integer results, <= 1e-12
otherwise), and (2) stores the REFERENCE outputs under ``tests/golden/`` as
small ``.npz`` files.  Inputs are stored as seeds (regenerated through
``speaker_verification_b200.synth``) unless they are tiny.  The GPU box has no
``/root/reference``; tests there read only the committed fixtures.

Versions used to generate the committed fixtures are recorded in
``tests/golden/VERSIONS.json``.
"""
from __future__ import annotations

import json
import os
import sys
import types

sys.dont_write_bytecode = True
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = os.environ.get("SPB_REFERENCE", "/root/reference")

import numpy as np  # noqa: E402

np.lib.pad = np.pad  # the shim (see module docstring)


def import_reference():
    sys.path.insert(0, REF)
    import speech_feature_extraction.speechpy.feature as rfeature
    import speech_feature_extraction.speechpy.functions as rfunctions
    import speech_feature_extraction.speechpy.processing as rprocessing
    for name in ("librosa", "matplotlib", "matplotlib.pyplot", "torchvision",
                 "torchvision.transforms", "tqdm"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                mod = types.ModuleType(name)
                if name == "tqdm":
                    mod.tqdm = lambda x, *a, **k: x
                sys.modules[name] = mod
    import matplotlib  # noqa: F401  (stub or real)
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    import evaluation as revaluation
    sys.path.remove(REF)
    return rprocessing, rfeature, rfunctions, revaluation


def same(a, b, tol=0.0, what=""):
    a = np.asarray(a)
    b = np.asarray(b)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    assert a.dtype == b.dtype, (what, a.dtype, b.dtype)
    if a.size == 0:
        return
    if tol == 0.0:
        assert np.array_equal(a, b), what
    else:
        err = np.max(np.abs(a.astype(np.float64) - b.astype(np.float64)))
        assert err <= tol, (what, err)


def main():
    from oracle import scoring_oracle as so
    from oracle import speechpy_oracle as o
    from speaker_verification_b200 import synth

    rp, rf, rfn, rev = import_reference()
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    fs = 16000

    # ------------------------------------------------------------------ A.1
    # frame geometry over edge lengths; -1 == ValueError
    lengths = [100, 319, 320, 321, 479, 480, 481, 640, 641, 799, 800, 1000,
               48000, 48001, 48159, 48160, 128000, 960000]
    geoms = [(0.020, 0.010), (0.020, 0.020), (0.025, 0.010)]
    table = []
    for fl, fst in geoms:
        for n in lengths:
            sig = np.arange(n, dtype=np.float64)
            row = [int(round(fs * fl)), int(round(fs * fst)), n]
            for zp in (False, True):
                try:
                    fr = rp.stack_frames(sig, fs, fl, fst, zero_padding=zp)
                    nf = fr.shape[0]
                    # signal = arange => frame values ARE the gather indices
                    idx = o.frame_indices(*o.frame_geometry(n, fs, fl, fst, zp))
                    same(fr, idx.astype(np.float64), what="indices")
                    same(fr, o.stack_frames(sig, fs, fl, fst, zero_padding=zp),
                         what="stack_frames")
                except ValueError:
                    nf = -1
                    try:
                        o.stack_frames(sig, fs, fl, fst, zero_padding=zp)
                        raise AssertionError("oracle did not raise")
                    except ValueError:
                        pass
                row.append(nf)
            table.append(row)
    np.savez_compressed(os.path.join(out_dir, "frame_geometry.npz"),
                        columns=np.array(["L", "S", "N", "F_nopad", "F_pad"]),
                        table=np.array(table, dtype=np.int64))

    # windowed stack_frames (hamming) on a seeded signal: store a few rows
    sig = synth.audio(11, 4000, np.float64)
    fr_ref = rp.stack_frames(sig, fs, 0.020, 0.010, filter=np.hamming, zero_padding=True)
    same(fr_ref, o.stack_frames(sig, fs, 0.020, 0.010, filter=np.hamming, zero_padding=True),
         what="windowed frames")
    pre_ref = rp.preemphasis(sig, shift=1, cof=0.98)
    same(pre_ref, o.preemphasis(sig, shift=1, cof=0.98), what="preemphasis")
    np.savez_compressed(os.path.join(out_dir, "stack_frames_hamming.npz"),
                        seed=11, n=4000, frames=fr_ref, preemphasis=pre_ref)

    # ------------------------------------------------------------------ A.3
    grid = [(40, 257, 16000, 0, None), (40, 513, 16000, 0, None),
            (40, 257, 16000, None, None), (40, 257, 16000, 300, 8000),
            (20, 257, 16000, 0, 4000), (26, 257, 8000, 0, None),
            (40, 257, 16000, 100, 7000), (64, 513, 16000, 50, 7600),
            (23, 129, 8000, 64, 3800), (80, 1025, 44100, 20, 20000),
            (40, 129, 16000, 0, None), (13, 257, 16000, 1000, 2000),
            # narrow band, many filters: repeated edges -> degenerate triangles
            (60, 129, 16000, 1000, 1500), (30, 65, 16000, 200, 900)]
    fb_store = {}
    for i, (m, k, f, lo, hi) in enumerate(grid):
        bank_ref = rf.filterbanks(m, k, f, lo, hi)
        bank_or = o.filterbanks(m, k, f, lo, hi)
        same(bank_ref, bank_or, what=f"filterbank {grid[i]}")
        edges = o.filterbank_edges(m, k, f, lo, hi)
        fb_store[f"params_{i}"] = np.array(
            [m, k, f, -1 if lo is None else lo, -1 if hi is None else hi], dtype=np.int64)
        fb_store[f"edges_{i}"] = edges.astype(np.int64)
        fb_store[f"bank_{i}"] = bank_ref
    fb_store["count"] = np.int64(len(grid))
    for fn in ("frequency_to_mel", "mel_to_frequency"):
        x = np.linspace(0.0, 8000.0, 33)
        same(getattr(rfn, fn)(x), getattr(o, fn)(x), what=fn)
    zh = np.array([0.0, 1e-300, -0.0, 3.0])
    same(rfn.zero_handling(zh), o.zero_handling(zh), what="zero_handling")
    tri_x = np.linspace(3, 9, 7)
    same(rfn.triangle(tri_x, 3, 5, 9), o.triangle(tri_x, 3, 5, 9), what="triangle")
    np.savez_compressed(os.path.join(out_dir, "filterbanks.npz"), **fb_store)

    # --------------------------------------------------- cfg1: 3 s, defaults
    def feature_block(tag, seed, n, dtype, kw):
        x = synth.audio(seed, n, dtype) if dtype != np.int16 else \
            np.round(synth.audio(seed, n, np.float64) * 32768 * 0.5).astype(np.int16)
        feats_ref, en_ref = rf.mfe(x, fs, **kw)
        feats_or, en_or = o.mfe(x, fs, **kw)
        same(feats_ref, feats_or, 0.0, tag + " mfe")
        same(en_ref, en_or, 0.0, tag + " energy")
        l_ref = rf.lmfe(x, fs, **kw)
        same(l_ref, o.lmfe(x, fs, **kw), 0.0, tag + " lmfe")
        c_ref = rf.mfcc(x, fs, num_cepstral=13, **kw)
        same(c_ref, o.mfcc(x, fs, num_cepstral=13, **kw), 0.0, tag + " mfcc")
        c_nodc = rf.mfcc(x, fs, num_cepstral=13, dc_elimination=False, **kw)
        same(c_nodc, o.mfcc(x, fs, num_cepstral=13, dc_elimination=False, **kw), 0.0,
             tag + " mfcc nodc")
        cube_ref = rf.extract_derivative_feature(c_ref)
        same(cube_ref, o.extract_derivative_feature(c_ref), 0.0, tag + " cube")
        w_ref = rp.cmvnw(l_ref, win_size=301, variance_normalization=True)
        same(w_ref, o.cmvnw(l_ref, win_size=301, variance_normalization=True), 0.0,
             tag + " cmvnw")
        same(w_ref, o.cmvnw_vectorised(l_ref, 301, True), 5e-5, tag + " cmvnw vec")
        wm_ref = rp.cmvnw(l_ref, win_size=301, variance_normalization=False)
        same(wm_ref, o.cmvnw(l_ref, win_size=301, variance_normalization=False), 0.0,
             tag + " cmvnw mean")
        g_ref = rp.cmvn(l_ref, variance_normalization=True)
        same(g_ref, o.cmvn(l_ref, variance_normalization=True), 0.0, tag + " cmvn")
        gm_ref = rp.cmvn(l_ref, variance_normalization=False)
        same(gm_ref, o.cmvn(l_ref, variance_normalization=False), 0.0, tag + " cmvn mean")
        return dict(lmfe=l_ref, energy=en_ref, mfcc=c_ref, mfcc_nodc=c_nodc,
                    cube=cube_ref, cmvnw=w_ref, cmvnw_mean=wm_ref, cmvn=g_ref,
                    cmvn_mean=gm_ref, x=x)

    d = feature_block("cfg1", 1, 48000, np.float32, {})
    x = d.pop("x").astype(np.float64)
    frames = rp.stack_frames(x, fs, 0.020, 0.010, zero_padding=False)
    same(frames, o.stack_frames(x, fs, 0.020, 0.010, zero_padding=False), 0.0, "cfg1 frames")
    ps = rp.power_spectrum(frames, 512)
    same(ps, o.power_spectrum(frames, 512), 0.0, "cfg1 power")
    same(rp.fft_spectrum(frames, 512), o.fft_spectrum(frames, 512), 0.0, "cfg1 fft")
    same(rp.log_power_spectrum(frames, 512, True), o.log_power_spectrum(frames, 512, True),
         0.0, "cfg1 logpower")
    np.savez_compressed(os.path.join(out_dir, "cfg1.npz"), seed=1, n=48000,
                        power_rows=np.array([0, 1, 149, 297]),
                        power=ps[[0, 1, 149, 297]],
                        logpower_norm_rows=rp.log_power_spectrum(frames, 512, True)[[0, 297]],
                        **d)

    # repo's own call shape: 25 ms / 10 ms / NFFT 1024 (constants.py:16-19, load_data.py:64-70)
    d = feature_block("nfft1024", 7, 32000, np.float32,
                      dict(frame_length=0.025, frame_stride=0.01, num_filters=40,
                           fft_length=1024))
    d.pop("x")
    np.savez_compressed(os.path.join(out_dir, "repo_call_nfft1024.npz"), seed=7, n=32000, **d)

    # int16 PCM input and float64 input
    d = feature_block("int16", 8, 16000, np.int16, {})
    d.pop("x")
    np.savez_compressed(os.path.join(out_dir, "int16_input.npz"), seed=8, n=16000,
                        lmfe=d["lmfe"], mfcc=d["mfcc"], energy=d["energy"])

    # digital silence and a signal with an all-zero stretch (zero_handling path)
    z = np.zeros(4000)
    l_sil = rf.lmfe(z, fs)
    same(l_sil, o.lmfe(z, fs), 0.0, "silence")
    c_sil = rf.mfcc(z, fs)
    same(c_sil, o.mfcc(z, fs), 0.0, "silence mfcc")
    half = synth.audio(9, 4000, np.float64)
    half[1000:2500] = 0.0
    l_half = rf.lmfe(half, fs)
    same(l_half, o.lmfe(half, fs), 0.0, "half silence")
    np.savez_compressed(os.path.join(out_dir, "silence.npz"), lmfe_silence=l_sil,
                        mfcc_silence=c_sil, seed=9, n=4000, zero_from=1000, zero_to=2500,
                        lmfe_half=l_half, mfcc_half=rf.mfcc(half, fs))

    # empty results: N in [L, L+S) -> (0, M) / (0, C)   (feature.py:144-145)
    for n in (320, 400, 479):
        xe = synth.audio(10, n, np.float64)
        assert rf.lmfe(xe, fs).shape == o.lmfe(xe, fs).shape == (0, 40)
        assert rf.mfcc(xe, fs).shape == o.mfcc(xe, fs).shape == (0, 13)

    # ------------------------------------------------ cmvnw / cmvn row counts
    store = {}
    rng = np.random.Generator(np.random.Philox(21))
    for rows in (1, 2, 5, 149, 150, 151, 299, 300, 301, 302, 400, 1000):
        v = rng.standard_normal((rows, 6)) * 3.0 + rng.standard_normal((1, 6)) * 5.0
        for vn in (False, True):
            r = rp.cmvnw(v, win_size=301, variance_normalization=vn)
            same(r, o.cmvnw(v, win_size=301, variance_normalization=vn), 0.0, "cmvnw rows")
            store[f"w301_rows{rows}_vn{int(vn)}"] = r
        r = rp.cmvnw(v, win_size=11, variance_normalization=True)
        same(r, o.cmvnw(v, win_size=11, variance_normalization=True), 0.0, "cmvnw w11")
        store[f"w11_rows{rows}_vn1"] = r
        store[f"in_rows{rows}"] = v
    np.savez_compressed(os.path.join(out_dir, "cmvnw_rows.npz"), **store)

    # cfg4 (60 s): keep every 61st row of the 5998x40 result
    x = synth.audio(4, 960000, np.float32)
    l_ref = rf.lmfe(x, fs)
    same(l_ref, o.lmfe(x, fs), 0.0, "cfg4 lmfe")
    w_ref = rp.cmvnw(l_ref, 301, True)
    same(w_ref, o.cmvnw(l_ref, 301, True), 0.0, "cfg4 cmvnw")
    _, e_ref = rf.mfe(x, fs)
    rows = np.arange(0, l_ref.shape[0], 61)
    np.savez_compressed(os.path.join(out_dir, "cfg4_rows.npz"), seed=4, n=960000,
                        rows=rows, lmfe=l_ref[rows], cmvnw=w_ref[rows], energy=e_ref[rows],
                        shape=np.array(l_ref.shape))

    # ------------------------------------------------------------ derivative
    v = rng.standard_normal((7, 5))
    for dw in (1, 2, 3):
        same(rp.derivative_extraction(v, dw), o.derivative_extraction(v, dw), 0.0, "delta")
    v1 = rng.standard_normal((4, 1))
    same(rf.extract_derivative_feature(v1), o.extract_derivative_feature(v1), 0.0, "cube 1col")
    np.savez_compressed(os.path.join(out_dir, "derivative.npz"), x=v,
                        cube=rf.extract_derivative_feature(v), x1=v1,
                        cube1=rf.extract_derivative_feature(v1))

    # ------------------------------------------- 8f-1: CMVN / FeatureCube transforms (utils.py:325-397)
    rutils = sys.modules["utils"]  # the reference's utils, imported by `from utils import *` in evaluation.py
    sig_c = synth.audio(21, 40000, np.float64)
    lm = rf.lmfe(sig_c, 16000, frame_length=0.020, frame_stride=0.010, num_filters=40, fft_length=512)
    cubes = {"seed": 21, "n": 40000}
    for deriv in (False, True):
        for norm in (False, True):
            rutils.c.DERIVATIVE, rutils.c.NORMALIZE = deriv, norm
            smp = rutils.CMVN()({"feature": lm.copy(), "label": 3})
            same(smp["feature"], o.cmvn_transform(lm.copy(), deriv, norm), 0.0, "CMVN transform")
            np.random.seed(77)
            cube_cls = rutils.FeatureCube3C((80, 40, 20, 3)) if deriv else rutils.FeatureCube((80, 40, 20))
            ref_cube = cube_cls({"feature": smp["feature"], "label": 3})["feature"]
            np.random.seed(77)
            idx = np.random.randint(lm.shape[0] - 80, size=20)
            same(ref_cube, o.feature_cube(smp["feature"], idx, 80), 0.0, "FeatureCube")
            cubes[f"idx_d{int(deriv)}n{int(norm)}"] = idx
            cubes[f"cube_d{int(deriv)}n{int(norm)}"] = ref_cube[:, :3]  # first 3 crops (small file)
    rutils.c.DERIVATIVE, rutils.c.NORMALIZE = False, False
    np.savez_compressed(os.path.join(out_dir, "feature_cubes.npz"), **cubes)

    # --------------------------------------------------------------- scoring
    sc = {}
    E, T = so.synth_embeddings(97, 113, 64, seed=5)
    from sklearn.metrics.pairwise import cosine_similarity
    s_sk = cosine_similarity(E.astype(np.float64), T.astype(np.float64))
    same(s_sk, so.cosine_scores(E, T), 1e-14, "cosine")
    sc["cos_small"] = s_sk
    cases = {}
    ti, tj, lab = so.synth_trials(97, 113, 2000, seed=5)
    cases["small"] = (lab, s_sk[ti, tj])
    Eb, Tb = so.synth_embeddings(4874, 4874, 512, seed=5)
    Sb = so.cosine_scores(Eb, Tb)
    ti, tj, lab = so.synth_trials(4874, 4874, 37720, seed=5)
    cases["cfg5"] = (lab, Sb[ti, tj])
    sc["cfg5_trial_scores_f64"] = Sb[ti, tj]
    r2 = np.random.Generator(np.random.Philox(33))
    tied = np.round(r2.standard_normal(5000), 1)
    cases["ties"] = ((r2.random(5000) < 0.3 + 0.2 * (tied > 0)).astype(np.int32), tied)
    cases["tiny"] = (np.array([1, 0, 1, 0], dtype=np.int32), np.array([0.7, 0.3, 0.2, 0.8]))
    cases["perfect"] = (np.array([1, 1, 0, 0], dtype=np.int32), np.array([0.9, 0.8, 0.2, 0.1]))
    for name, (lab, s) in cases.items():
        eer, auc, fpr, tpr = rev.get_eer_auc(lab, s)
        e2, a2, _, _ = so.get_eer_auc(lab, s)
        assert eer == e2 and auc == a2, name
        e3, a3 = so.eer_auc_restated(lab, s)
        assert abs(e3 - eer) < 1e-9 and abs(a3 - auc) < 1e-12, (name, e3, eer, a3, auc)
        sc[f"eer_{name}"] = np.float64(eer)
        sc[f"auc_{name}"] = np.float64(auc)
        # ROC curve as the reference's get_eer_auc returns it, on the float32 scores the scorer emits
        from sklearn.metrics import roc_curve as sk_roc_curve
        s32 = np.asarray(s, dtype=np.float32)
        _, _, fpr32, tpr32 = rev.get_eer_auc(lab, s32)
        _, _, thr32 = sk_roc_curve(lab, s32, pos_label=1)
        f3, t3, th3 = so.roc_curve_restated(lab, s32)
        assert np.array_equal(f3, fpr32) and np.array_equal(t3, tpr32), name
        assert np.array_equal(th3[1:], thr32[1:].astype(np.float64)) and np.isinf(thr32[0]), name
        f4, t4, _ = so.roc_curve_restated(lab, s)
        assert np.array_equal(f4, fpr) and np.array_equal(t4, tpr), name
        sc[f"fpr_{name}"] = fpr32
        sc[f"tpr_{name}"] = tpr32
        if name in ("ties", "tiny", "perfect"):
            sc[f"label_{name}"] = lab
            sc[f"score_{name}"] = s
    np.savez_compressed(os.path.join(out_dir, "scoring.npz"), **sc)

    import scipy
    import sklearn
    with open(os.path.join(out_dir, "VERSIONS.json"), "w") as fh:
        json.dump({"python": sys.version.split()[0], "numpy": np.__version__,
                   "scipy": scipy.__version__, "sklearn": sklearn.__version__,
                   "reference": REF, "shim": "np.lib.pad = np.pad"}, fh, indent=1)
    total = sum(os.path.getsize(os.path.join(out_dir, f)) for f in os.listdir(out_dir))
    print("oracle == reference on all cases; goldens written:", total, "bytes")


if __name__ == "__main__":
    main()
