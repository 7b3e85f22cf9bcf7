"""Parity of the CUDA path (through the drop-in speechpy API -> ctypes -> C-ABI) against
the reference's committed outputs and the oracle.  Tolerances are BASELINE.json's:
frame indices / bin edges bit-exact, log-mel / MFCC <= 1e-4 abs, cmvnw <= 1e-4 abs."""
import ctypes

import numpy as np
import pytest

from conftest import golden
from oracle import speechpy_oracle as o
from speaker_verification_b200 import synth

pytestmark = pytest.mark.gpu
FS = 16000


def _lib_path():
    from speaker_verification_b200 import _lib
    _lib.load()
    return _lib.LIB_PATH

TOL = 1e-4


def maxabs(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    return float(np.max(np.abs(a - b))) if a.size else 0.0


def test_stack_frames_bit_exact(sp):
    g = golden("frame_geometry.npz")
    for L, S, n, f_nopad, f_pad in g["table"]:
        if n > 200000:
            continue
        sig = np.arange(n, dtype=np.float64)
        for zp, want in ((False, f_nopad), (True, f_pad)):
            if want < 0:
                with pytest.raises(ValueError):
                    sp.processing.stack_frames(sig, FS, L / FS, S / FS, zero_padding=zp)
                continue
            fr = sp.processing.stack_frames(sig, FS, L / FS, S / FS, zero_padding=zp)
            assert fr.dtype == np.float64 and fr.shape == (want, L)
            assert np.array_equal(fr, o.stack_frames(sig, FS, L / FS, S / FS, zero_padding=zp))
    h = golden("stack_frames_hamming.npz")
    sig = synth.audio(int(h["seed"]), int(h["n"]), np.float64)
    fr = sp.processing.stack_frames(sig, FS, 0.020, 0.010, filter=np.hamming, zero_padding=True)
    assert np.array_equal(fr, h["frames"])  # float64 product of the same two operands


def test_power_spectrum(sp):
    g = golden("cfg1.npz")
    x = synth.audio(1, 48000, np.float32).astype(np.float64)
    fr = sp.processing.stack_frames(x, FS, 0.020, 0.010, zero_padding=False)
    ps = sp.processing.power_spectrum(fr, 512)
    assert ps.dtype == np.float64 and ps.shape == (298, 257)
    ref = o.power_spectrum(fr, 512)
    assert np.max(np.abs(ps - ref)) <= 2e-6 * np.max(ref)
    assert np.max(np.abs(ps[g["power_rows"]] - g["power"])) <= 2e-6 * np.max(ref)
    mag = sp.processing.fft_spectrum(fr, 512)
    assert np.max(np.abs(mag - o.fft_spectrum(fr, 512))) <= 2e-6 * np.max(np.sqrt(ref * 512))
    for nfft in (64, 256, 1024, 2048):
        for L in (50, 320, 400):
            fr2 = np.ascontiguousarray(fr[:37, :L])
            r = o.power_spectrum(fr2, nfft)
            p = sp.processing.power_spectrum(fr2, nfft)
            assert p.shape == r.shape
            assert np.max(np.abs(p - r)) <= 3e-6 * np.max(r), (nfft, L)
    lp = sp.processing.log_power_spectrum(fr, 512, True)
    assert maxabs(lp[[0, 297]], g["logpower_norm_rows"]) < 1e-3
    # any other fft_points (numpy's rfft(n=) takes them all): direct DFT in float64 on the device
    for nfft, L in ((500, 320), (400, 400), (300, 320), (16, 320), (8192, 320), (1, 50), (7, 5), (6000, 100)):
        fr2 = np.ascontiguousarray(fr[:23, :L]) if L <= 320 else np.ascontiguousarray(np.tile(fr[:23], (1, 2))[:, :L])
        for fn, ofn in ((sp.processing.power_spectrum, o.power_spectrum), (sp.processing.fft_spectrum, o.fft_spectrum)):
            r = ofn(fr2, nfft)
            p = fn(fr2, nfft)
            assert p.shape == r.shape and p.dtype == np.float64
            assert np.max(np.abs(p - r)) <= 1e-11 * max(np.max(np.abs(r)), 1e-300), (nfft, L)
        lp = sp.processing.log_power_spectrum(fr2, nfft, True)
        ref_p = o.power_spectrum(fr2, nfft)
        ref_lp = 10 * np.log10(np.maximum(ref_p, 1e-20))
        assert np.max(np.abs(lp - (ref_lp - ref_lp.max()))) < 1e-9, (nfft, L)


def test_preemphasis_and_log_power_spectrum_on_device(sp):
    """SURVEY 8f-2 (processing.py:45-58, 177-198): both run on the device; preemphasis is
    bit-exact in numpy's dtype (two separately rounded operations), for any shift."""
    g = golden("stack_frames_hamming.npz")
    sig = synth.audio(11, 4000, np.float64)
    assert np.array_equal(sp.processing.preemphasis(sig, 1, 0.98), g["preemphasis"])
    rng = np.random.Generator(np.random.Philox(21))
    for x in (rng.standard_normal(10007), rng.standard_normal(513).astype(np.float32),
              rng.integers(-32768, 32767, 9999).astype(np.int16), rng.integers(-10**6, 10**6, 777).astype(np.int32),
              rng.standard_normal((7, 33))):
        for shift, cof in ((1, 0.98), (3, 0.5), (-2, 0.97), (x.size + 5, 0.9), (0, 1.0)):
            ref = o.preemphasis(x, shift, cof)
            got = sp.processing.preemphasis(x, shift, cof)
            assert got.dtype == ref.dtype and got.shape == ref.shape and np.array_equal(got, ref), (x.dtype, shift)
    assert sp.processing.preemphasis(np.zeros(0)).shape == (0,)
    # log power spectrum: float32 FFT, float64 log epilogue
    fr = o.stack_frames(synth.audio(1, 48000, np.float64), FS, 0.020, 0.010, zero_padding=False)
    for normalize in (False, True):
        ref = o.log_power_spectrum(fr, 512, normalize)
        got = sp.processing.log_power_spectrum(fr, 512, normalize)
        assert got.shape == ref.shape and got.dtype == np.float64
        p_ref = o.power_spectrum(fr, 512)
        strong = p_ref >= 1e-4 * p_ref.max(axis=1, keepdims=True)
        assert np.max(np.abs(got - ref)[strong]) < 1e-3          # dB, bins within 40 dB of the row maximum
        assert np.max(np.abs(got - ref)) < 0.05                  # every bin
    assert np.max(sp.processing.log_power_spectrum(fr, 512, True)) == 0.0
    silent = sp.processing.log_power_spectrum(np.zeros((3, 320)), 512, False)
    assert np.all(silent == -200.0)                              # floor 1e-20 -> -200 dB


def test_generic_kernel_matches_fast(sp):
    """Route the same call through the generic radix-2 kernel: two independent FFTs."""
    from speaker_verification_b200 import _lib
    lib = _lib.load()
    x = synth.audio(12, 20000, np.float32)
    fast = sp.feature.lmfe(x, FS)
    with _lib.options(feature_kernel=1):
        gen = sp.feature.lmfe(x, FS)
        mf_gen = sp.feature.mfcc(x, FS)
    assert maxabs(fast, gen) < 2e-5
    assert maxabs(fast, o.lmfe(x, FS)) < TOL
    assert maxabs(mf_gen, o.mfcc(x, FS)) < TOL


@pytest.mark.parametrize("fl,fs_,nf", [(0.025, 0.010, 40), (0.032, 0.0125, 40), (0.0250625, 0.0100625, 24), (0.064, 0.016, 40)])
def test_fast1024_kernel_matches_generic_and_oracle(sp, fl, fs_, nf):
    """NFFT 1024 (the reference's own call, load_data.py:64-70 + constants.py) runs the warp-per-frame
    register FFT kernel; the generic radix-2 kernel is an independent second implementation.
    Cases: L = 400 / S = 160 (pruned inputs), L = 512 / S = 200 (unpruned), odd L = 401 and odd
    S = 161 (unaligned frame starts), L = 1024 (no zero padding)."""
    x = synth.audio(31, 50000, np.float32)
    kw = dict(frame_length=fl, frame_stride=fs_, num_filters=nf, fft_length=1024)
    fast_mfe, fast_en = sp.feature.mfe(x, FS, **kw)
    fast_l = sp.feature.lmfe(x, FS, **kw)
    fast_c = sp.feature.mfcc(x, FS, frame_length=fl, frame_stride=fs_, num_filters=nf, fft_length=1024)
    from speaker_verification_b200 import _lib
    with _lib.options(feature_kernel=1):
        gen_mfe, gen_en = sp.feature.mfe(x, FS, **kw)
        gen_l = sp.feature.lmfe(x, FS, **kw)
    assert fast_l.shape == gen_l.shape and fast_l.shape[0] > 100
    assert maxabs(fast_l, gen_l) < 2e-5
    assert np.max(np.abs(fast_mfe - gen_mfe) / np.maximum(gen_mfe, 1e-30)) < 2e-5
    assert np.max(np.abs(fast_en - gen_en) / gen_en) < 2e-5
    ref_mfe, ref_en = o.mfe(x, FS, **kw)
    assert maxabs(fast_l, o.lmfe(x, FS, **kw)) < TOL
    assert np.max(np.abs(fast_en - ref_en) / ref_en) < 1e-5
    assert maxabs(fast_c, o.mfcc(x, FS, frame_length=fl, frame_stride=fs_, num_filters=nf, fft_length=1024)) < TOL
    # ragged batch through the same kernel (TMA-staged tiles, partial last tiles)
    from speaker_verification_b200 import batch
    sigs = [synth.audio(70 + i, n, np.float32) for i, n in enumerate((40000, 1700, 16000, 99999))]
    arena, starts, lens = batch.pack(sigs)
    fb = batch.FeatureBatch(starts, lens, **kw)
    for sgn, got in zip(sigs, fb.split(fb.lmfe_cmvnw(arena))):
        lm = o.lmfe(sgn, FS, **kw)
        if lm.shape[0] > 1:
            assert maxabs(got, o.cmvnw_vectorised(lm, 301, True)) < TOL
    fb.close()


@pytest.mark.parametrize("name,kw", [
    ("cfg1.npz", {}),
    ("repo_call_nfft1024.npz", dict(frame_length=0.025, frame_stride=0.01, num_filters=40, fft_length=1024)),
])
def test_feature_chain_against_reference_goldens(sp, name, kw):
    g = golden(name)
    x = synth.audio(int(g["seed"]), int(g["n"]), np.float32)
    feats, energy = sp.feature.mfe(x, FS, **kw)
    assert feats.dtype == np.float64 and energy.dtype == np.float64
    assert maxabs(np.log(feats), g["lmfe"]) < TOL
    assert maxabs(np.log(energy), np.log(g["energy"])) < TOL
    lm = sp.feature.lmfe(x, FS, **kw)
    assert lm.dtype == np.float64 and maxabs(lm, g["lmfe"]) < TOL
    c = sp.feature.mfcc(x, FS, num_cepstral=13, **kw)
    assert c.dtype == np.float64 and maxabs(c, g["mfcc"]) < TOL
    assert maxabs(sp.feature.mfcc(x, FS, num_cepstral=13, dc_elimination=False, **kw), g["mfcc_nodc"]) < TOL
    cube = sp.feature.extract_derivative_feature(g["mfcc"])
    assert cube.dtype == np.float64 and maxabs(cube, g["cube"]) < 1e-12
    w = sp.processing.cmvnw(g["lmfe"], win_size=301, variance_normalization=True)
    assert w.dtype == np.float32 and maxabs(w, g["cmvnw"]) < TOL
    wm = sp.processing.cmvnw(g["lmfe"], win_size=301, variance_normalization=False)
    assert wm.dtype == np.float32 and maxabs(wm, g["cmvnw_mean"]) < TOL
    # the full chain on GPU outputs (what config 1 measures)
    assert maxabs(sp.processing.cmvnw(lm, 301, True), g["cmvnw"]) < TOL
    cm = sp.processing.cmvn(g["lmfe"], variance_normalization=True)
    assert cm.dtype == np.float64 and maxabs(cm, g["cmvn"]) < 1e-10
    assert maxabs(sp.processing.cmvn(g["lmfe"], False), g["cmvn_mean"]) < 1e-12


def test_input_dtypes_silence_and_empty(sp):
    g = golden("int16_input.npz")
    x = np.round(synth.audio(8, 16000, np.float64) * 32768 * 0.5).astype(np.int16)
    assert maxabs(sp.feature.lmfe(x, FS), g["lmfe"]) < TOL
    assert maxabs(sp.feature.mfcc(x, FS), g["mfcc"]) < TOL
    x64 = synth.audio(8, 16000, np.float64)
    assert maxabs(sp.feature.lmfe(x64, FS), o.lmfe(x64, FS)) < TOL
    s = golden("silence.npz")
    sil = sp.feature.lmfe(np.zeros(4000), FS)
    assert maxabs(sil, s["lmfe_silence"]) < 1e-5  # exact zeros -> eps -> -36.0436...
    assert maxabs(sp.feature.mfcc(np.zeros(4000), FS), s["mfcc_silence"]) < TOL
    half = synth.audio(9, 4000, np.float64)
    half[1000:2500] = 0.0
    assert maxabs(sp.feature.lmfe(half, FS), s["lmfe_half"]) < TOL
    assert maxabs(sp.feature.mfcc(half, FS), s["mfcc_half"]) < TOL
    for n in (320, 400, 479):
        xe = synth.audio(10, n, np.float64)
        assert sp.feature.lmfe(xe, FS).shape == (0, 40)
        assert sp.feature.mfcc(xe, FS).shape == (0, 13)
        f, e = sp.feature.mfe(xe, FS)
        assert f.shape == (0, 40) and e.shape == (0,)
    with pytest.raises(ValueError):
        sp.feature.lmfe(synth.audio(10, 100, np.float64), FS)
    with pytest.raises(AssertionError):
        sp.feature.lmfe(synth.audio(10, 4000, np.float64), FS, high_frequency=9000)


@pytest.mark.parametrize("n", [480, 481, 640, 5000, 48001, 48159, 48160])
def test_edge_lengths(sp, n):
    x = synth.audio(100 + n, n, np.float32)
    ref = o.lmfe(x, FS)
    got = sp.feature.lmfe(x, FS)
    assert got.shape == ref.shape and maxabs(got, ref) < TOL
    refc = o.mfcc(x, FS)
    assert maxabs(sp.feature.mfcc(x, FS), refc) < TOL


def test_other_filterbank_shapes(sp):
    x = synth.audio(77, 24000, np.float32)
    for kw in (dict(num_filters=20, high_frequency=4000), dict(num_filters=26, low_frequency=100, high_frequency=7000),
               dict(num_filters=64, fft_length=1024, frame_length=0.025, low_frequency=50, high_frequency=7600),
               dict(frame_length=0.032, frame_stride=0.016), dict(frame_length=0.0205, frame_stride=0.0101)):
        assert maxabs(sp.feature.lmfe(x, FS, **kw), o.lmfe(x, FS, **kw)) < TOL, kw
        kc = dict(kw)
        assert maxabs(sp.feature.mfcc(x, FS, num_cepstral=12, **kc), o.mfcc(x, FS, num_cepstral=12, **kc)) < TOL, kw


def test_cmvnw_large_offset_and_ramp(sp):
    """processing.cmvnw takes float64 and subtracts the window mean in float64 (processing.py:302-307).
    Inputs whose offset dwarfs their variation (1e4 +- 1e-2, a ramp of 1e6 with noise on it) lose
    everything if they are narrowed to float32 before the mean is subtracted."""
    rng = np.random.Generator(np.random.Philox(91))
    noise = rng.standard_normal((1200, 40)) * 1e-2
    ramp = np.linspace(0.0, 1e6, 1200)[:, None] + rng.standard_normal((1200, 40))
    for v in (1e4 + noise, -3e5 + noise * 10, ramp):
        # variance mode on the ramp is left out: its mean-subtracted rows span 5 decades inside one window
        # (6e4 at the reflected edges, 1 inside) and stage 2's one-pass float32 variance (documented in
        # include/speechpy_b200.h) is then less accurate than numpy's two-pass float32 std
        for vn in ((False,) if v is ramp else (True, False)):
            got = sp.processing.cmvnw(v, 301, vn)
            ref = o.cmvnw_vectorised(v, 301, vn)
            scale = max(1.0, float(np.abs(ref).max()))
            assert got.dtype == np.float32
            assert np.max(np.abs(got - ref)) < 1e-4 * scale, (float(np.abs(got - ref).max()), scale)


def test_cmvnw_row_counts(sp):
    g = golden("cmvnw_rows.npz")
    for rows in (1, 2, 5, 149, 150, 151, 299, 300, 301, 302, 400, 1000):
        v = g[f"in_rows{rows}"]
        for vn in (0, 1):
            got = sp.processing.cmvnw(v, 301, bool(vn))
            assert got.dtype == np.float32
            assert maxabs(got, g[f"w301_rows{rows}_vn{vn}"]) < TOL, (rows, vn)
        assert maxabs(sp.processing.cmvnw(v, 11, True), g[f"w11_rows{rows}_vn1"]) < TOL, rows
    wide = np.random.default_rng(3).standard_normal((700, 100))
    assert maxabs(sp.processing.cmvnw(wide, 301, True), o.cmvnw_vectorised(wide, 301, True)) < TOL
    odd = np.random.default_rng(4).standard_normal((64, 13)) * 4 + 2
    assert maxabs(sp.processing.cmvnw(odd, 301, True), o.cmvnw(odd, 301, True)) < TOL


def test_cfg4_long_form(sp):
    g = golden("cfg4_rows.npz")
    x = synth.audio(4, 960000, np.float32)
    lm = sp.feature.lmfe(x, FS)
    assert lm.shape == tuple(g["shape"])
    rows = g["rows"]
    assert maxabs(lm[rows], g["lmfe"]) < TOL
    w = sp.processing.cmvnw(lm, 301, True)
    assert maxabs(w[rows], g["cmvnw"]) < TOL
    _, e = sp.feature.mfe(x, FS)
    assert maxabs(np.log(e[rows]), np.log(g["energy"])) < TOL


def test_derivative_and_cube(sp):
    g = golden("derivative.npz")
    assert maxabs(sp.feature.extract_derivative_feature(g["x"]), g["cube"]) < 1e-14
    assert maxabs(sp.feature.extract_derivative_feature(g["x1"]), g["cube1"]) < 1e-14
    for dw in (1, 2, 3):
        assert maxabs(sp.processing.derivative_extraction(g["x"], dw), o.derivative_extraction(g["x"], dw)) < 1e-14


def test_reference_test_suite_semantics(sp):
    """speech_feature_extraction/tests/test_speechpy.py:56-106 against the replacement."""
    rng = np.random.default_rng(0)
    signal = rng.normal(0, 0.1, 1000000)
    pre = sp.processing.preemphasis(signal, cof=0.98)
    assert pre.ndim == 1 and pre.shape == signal.shape
    frames = sp.processing.stack_frames(signal, sampling_frequency=FS, frame_length=0.02, frame_stride=0.02,
                                        filter=lambda x: np.ones((x,)), zero_padding=True)
    assert frames.shape[0] == int(np.ceil((signal.shape[0] - 320) / 320))
    fv = rng.random((50, 100))
    nf = sp.processing.cmvn(fv, variance_normalization=True)
    assert nf.shape == fv.shape
    assert np.allclose(np.mean(nf, axis=0), np.zeros((1, nf.shape[1])))
    assert np.allclose(np.std(nf, axis=0), np.ones((1, nf.shape[1])))
    mf = sp.feature.mfcc(signal, sampling_frequency=FS, frame_length=0.020, num_cepstral=13, frame_stride=0.01,
                         num_filters=40, fft_length=512, low_frequency=0, high_frequency=None)
    assert mf.shape[1] == 13


def test_reference_test_file_runs_unchanged_against_the_dropin(sp):
    """sfe/tests/test_speechpy.py ("must pass unchanged"), staged verbatim under
    oracle/_ref/tests by oracle/make_ref.py, run by pytest in a fresh process in which `speechpy` is the
    GPU-backed drop-in (install_dropin() before the file's own `from speechpy import ...`)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    test_file = os.path.join(root, "oracle", "_ref", "tests", "test_speechpy.py")
    if not os.path.exists(test_file):
        pytest.skip("oracle/_ref not staged (python oracle/make_ref.py where /root/reference exists)")
    code = ("import sys; sys.path.insert(0, %r); import speaker_verification_b200 as p; p.install_dropin(); "
            "import speechpy, pytest; assert 'speaker_verification_b200' in speechpy.__file__; "
            "sys.exit(pytest.main(['-q', '-p', 'no:cacheprovider', %r]))" % (root, test_file))
    r = subprocess.run([sys.executable, "-c", code], cwd=os.path.dirname(test_file), capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert " passed" in r.stdout and "failed" not in r.stdout
