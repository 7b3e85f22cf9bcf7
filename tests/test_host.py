"""CPU-side checks: the library builds, exports every symbol the header declares,
fails loudly without a GPU, and the host logic (filterbank, geometry, sharding,
import shims) matches the reference's behaviour.  No compute calls."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT, golden


def header_symbols():
    text = open(os.path.join(ROOT, "include", "speechpy_b200.h")).read()
    return sorted(set(re.findall(r"SPB_API\s+(?:const\s+char\s*\*|int)\s*(spb_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol(lib_built):
    lib = ctypes.CDLL(lib_built)
    names = header_symbols()
    assert len(names) >= 35
    for n in names:
        assert hasattr(lib, n), n
    from speaker_verification_b200 import _lib
    assert sorted(_lib.PROTOTYPES) == names  # the ctypes table mirrors the header one to one
    assert lib.spb_version() == 100
    # ... and nothing undeclared is exported: no test hooks or debug switches outside the header
    import subprocess
    out = subprocess.run(["nm", "-D", "--defined-only", lib_built], capture_output=True, text=True, check=True).stdout
    exported = sorted(line.split()[-1] for line in out.splitlines() if line.split()[-1].startswith("spb_"))
    assert exported == names
    blob = open(lib_built, "rb").read()
    assert b"SPB_TC_DBG" not in blob and b"spb_debug" not in blob


def test_options_api(lib_built):
    """spb_set_option / spb_get_option (declared kernel-selection knobs): round trip, range checks."""
    from speaker_verification_b200 import _lib
    assert _lib.get_option("feature_kernel") == 0 and _lib.get_option("redo") == 1
    with _lib.options(feature_kernel=2, tau_tc=1e-6):
        assert _lib.get_option("feature_kernel") == 2 and _lib.get_option("tau_tc") == 1e-6
    assert _lib.get_option("feature_kernel") == 0
    for bad in (("feature_kernel", 7), ("tc_version", 3), ("no_such_option", 1), ("tau_fft", -1.0)):
        with pytest.raises(AssertionError):
            _lib.set_option(*bad)


def test_host_integer_math_matches_reference_table(lib_built):
    from speaker_verification_b200 import _lib
    lib = _lib.load()
    g = golden("frame_geometry.npz")
    for L, S, n, f_nopad, f_pad in g["table"]:
        for zp, want in ((0, f_nopad), (1, f_pad)):
            F = _lib.c_i64()
            rc = lib.spb_num_frames(int(n), int(L), int(S), zp, ctypes.byref(F))
            if want < 0:
                assert rc == _lib.SPB_ERR_NEGATIVE_DIM
                assert b"negative dimensions" in lib.spb_last_error(None)
            else:
                assert rc == 0 and F.value == want


def test_no_cpu_fallback(lib_built):
    """Without a CUDA device the product path must raise, not compute."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    code = ("import sys; sys.path.insert(0, %r)\n"
            "import numpy as np, speaker_verification_b200 as p; p.install_dropin()\n"
            "import speechpy\n"
            "try:\n"
            "    speechpy.feature.lmfe(np.zeros(4000), 16000)\n"
            "except RuntimeError as e:\n"
            "    print('RAISED', e)\n" % ROOT)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True).stdout
    assert "RAISED" in out and "no CPU fallback" in out


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "speaker_verification_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("# oracle", ""), os.path.join(dirpath, f)


def test_filterbank_host_bit_exact():
    import speaker_verification_b200 as pkg
    pkg.install_dropin()
    from speechpy import feature
    g = golden("filterbanks.npz")
    for i in range(int(g["count"])):
        m, k, f, lo, hi = (int(v) for v in g[f"params_{i}"])
        lo = None if lo < 0 else lo
        hi = None if hi < 0 else hi
        bank = feature.filterbanks(m, k, f, lo, hi)
        assert bank.dtype == np.float64 and np.array_equal(bank, g[f"bank_{i}"])
    with pytest.raises(AssertionError):
        feature.filterbanks(40, 257, 16000, 0, 9000)
    with pytest.raises(AssertionError):
        feature.filterbanks(40, 257, 16000, -5, 8000)


def test_import_spellings_and_api_surface():
    """Every import spelling the reference uses and every name
    speech_feature_extraction/tests/test_speechpy.py:24-51 checks."""
    import speaker_verification_b200 as pkg
    pkg.install_dropin()
    import speech_feature_extraction.speechpy as speech2
    import speech_feature_extraction.speechpy.feature as speech
    import speechpy
    from speechpy import feature, functions, processing
    assert speech.lmfe is feature.lmfe or speech.lmfe.__name__ == "lmfe"
    assert hasattr(speech2.feature, "extract_derivative_feature") and hasattr(speech2.processing, "cmvn")
    for n in ("preemphasis", "stack_frames", "fft_spectrum", "power_spectrum", "log_power_spectrum",
              "derivative_extraction", "cmvn", "cmvnw"):
        assert hasattr(processing, n) and hasattr(speechpy.processing, n)
    for n in ("filterbanks", "mfcc", "mfe", "lmfe", "extract_derivative_feature"):
        assert hasattr(feature, n)
    for n in ("frequency_to_mel", "mel_to_frequency", "triangle", "zero_handling"):
        assert hasattr(functions, n)
    import inspect
    sig = inspect.signature(feature.mfcc)
    assert list(sig.parameters) == ["signal", "sampling_frequency", "frame_length", "frame_stride",
                                    "num_cepstral", "num_filters", "fft_length", "low_frequency",
                                    "high_frequency", "dc_elimination"]
    d = {k: v.default for k, v in sig.parameters.items()}
    assert (d["frame_length"], d["frame_stride"], d["num_cepstral"], d["num_filters"], d["fft_length"],
            d["low_frequency"], d["high_frequency"], d["dc_elimination"]) == (0.020, 0.01, 13, 40, 512, 0, None, True)
    sig = inspect.signature(processing.stack_frames)
    assert sig.parameters["frame_stride"].default == 0.020 and sig.parameters["zero_padding"].default is True
    assert inspect.signature(processing.cmvnw).parameters["win_size"].default == 301
    # argument validation that needs no device
    with pytest.raises(AssertionError):
        processing.stack_frames(np.zeros((4, 4)), 16000)
    with pytest.raises(AssertionError):
        processing.cmvnw(np.zeros((4, 4)), win_size=300)
    with pytest.raises(AssertionError):
        processing.cmvnw(np.zeros((4, 4)), win_size=301.0)
    with pytest.raises(ValueError):
        processing.stack_frames(np.zeros(100), 16000, 0.02, 0.01)


def test_pack_and_shard_balance():
    from speaker_verification_b200 import batch, synth
    lengths = synth.cfg3_lengths(20000, seed=3)
    assert lengths.min() >= 64000 and lengths.max() <= 320000
    for ws in (2, 4, 8):
        shards = batch.shard_indices_fast(lengths, ws)
        allidx = np.sort(np.concatenate(shards))
        assert np.array_equal(allidx, np.arange(len(lengths)))
        loads = np.array([lengths[s].sum() for s in shards])
        assert (loads.max() - loads.min()) <= lengths.max()
    shards = batch.shard_indices(lengths[:500], 4)
    loads = np.array([lengths[:500][s].sum() for s in shards])
    assert (loads.max() - loads.min()) <= lengths.max()
    sigs = [np.arange(n, dtype=np.float32) for n in (5, 8, 3)]
    arena, starts, lens = batch.pack(sigs)  # starts on multiples of 8 samples (32 bytes: whole sectors)
    assert list(starts) == [0, 8, 16] and list(lens) == [5, 8, 3] and arena.shape[0] == 24
    assert np.array_equal(arena[8:16], sigs[1]) and arena[5] == 0


def test_host_fft_emulation(tmp_path):
    """The __host__ __device__ FFT building blocks, emulated lane by lane on the CPU."""
    exe = str(tmp_path / "host_fft_check")
    subprocess.check_call(["nvcc", "-O1", "-Wno-deprecated-gpu-targets", "-I",
                           os.path.join(ROOT, "speaker_verification_b200", "csrc"), "-o", exe,
                           os.path.join(ROOT, "tests", "host", "host_fft_check.cu")])
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0 and "OK" in out.stdout, out.stdout


def test_fast1024_index_algebra():
    """numpy model of features_fast1024's data movement (csrc/feat_kernels.cu): lane l holds
    z[32 j + l]; FFT16 over j, twiddle W512^(l k1), exchange to lanes (k1, b) holding l = 2a + b,
    FFT16 over a, radix-2 across the half-warps, untangle of bins 0..256 with partners fetched
    from lane ((16 - k1) & 15) + 16 (element 15 - q; lane k1 = 0 uses the previous round's value)."""
    rng = np.random.default_rng(0)
    L = 400
    x = rng.standard_normal(L)
    xx = np.zeros(1024); xx[:L] = x
    z = xx[0::2] + 1j * xx[1::2]                      # 512 complex
    W = lambda N, e: np.exp(-2j * np.pi * e / N)
    # stage 1: lane l holds v[j] = z[32 j + l]
    V = np.zeros((32, 16), complex)                   # [l][k1]
    for l in range(32):
        v = np.array([z[32 * j + l] for j in range(16)])
        V[l] = np.fft.fft(v)                          # over j -> k1
        V[l] *= W(512, l * np.arange(16))
    # exchange: lane m=(k1,b) gets y[a] = V[2a+b][k1]
    Zs = np.zeros((32, 16), complex)                  # lane (k1 + 16 s) element q -> Z[k1 + 16 q + 256 s]
    Y = np.zeros((32, 16), complex)
    for m in range(32):
        k1, b = m & 15, m >> 4
        y = np.array([V[2 * a + b][k1] for a in range(16)])
        Y[m] = np.fft.fft(y)
        if b == 1:
            Y[m] *= W(32, np.arange(16))
    for m in range(32):
        k1, b = m & 15, m >> 4
        other = Y[m ^ 16]
        Zs[m] = Y[m] + other if b == 0 else other - Y[m]
    Zref = np.fft.fft(z)
    for m in range(32):
        k1, s = m & 15, m >> 4
        for q in range(16):
            assert abs(Zs[m][q] - Zref[k1 + 16 * q + 256 * s]) < 1e-9
    # untangle on lanes s = 0
    Xref = np.fft.rfft(xx)
    P = np.zeros(257)
    recv_prev = None
    for q in range(16):
        recv = [Zs[((16 - (m & 15)) & 15) + 16][15 - q] for m in range(32)]   # shfl of element 15-q from lane partner
        for m in range(16):
            k1 = m
            k = k1 + 16 * q
            if k1 == 0:
                zp = Zs[0][0] if q == 0 else recv_prev[0]
            else:
                zp = recv[m]
            zk = Zs[m][q]
            A = complex(zk.real + zp.real, zk.imag - zp.imag)
            B = complex(zk.imag + zp.imag, zp.real - zk.real)
            w = W(1024, k)
            X2 = A + B * w
            assert abs(X2 / 2 - Xref[k]) < 1e-9, (k, X2 / 2, Xref[k])
            P[k] = 0.25 * abs(X2) ** 2
        recv_prev = recv
    z256 = Zs[16][0]
    P[256] = z256.real ** 2 + z256.imag ** 2
    assert np.allclose(P, np.abs(Xref[:257]) ** 2)
    x0 = Zs[0][0].real + Zs[0][0].imag; xn = Zs[0][0].real - Zs[0][0].imag
    assert abs(x0 - Xref[0].real) < 1e-9 and abs(xn - Xref[512].real) < 1e-9



def test_bench_reference_arm_runs_on_cpu_and_prints_the_contract_line():
    """`bench.py --impl reference` (the driver's CPU arm) needs no GPU: one JSON line with the metric, the
    strong-scaling config of the other arm, kind "reference" when oracle/_ref is staged ("port" otherwise), and
    BLAS threads pinned before numpy is imported."""
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1",
                        "--warmup", "0", "--cpu-utts", "8"], capture_output=True, text=True, timeout=600, cwd=root)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    from oracle import ref_loader
    assert d["impl"] == "reference" and d["unit"] == "audio-hours/s" and d["value"] > 0
    assert d["cpu_baseline"]["kind"] == ("reference" if ref_loader.available() else "port")
    assert d["cpu_baseline"]["cores"] >= 1 and d["e2e"]["h2d_bytes_per_step"] == 0
    assert d["config"]["total_utterances"] == 148642 and d["scaling"] == "strong"
    src = open(os.path.join(root, "bench.py")).read()
    assert src.index('os.environ[_k] = "1"') < src.index("import numpy")
