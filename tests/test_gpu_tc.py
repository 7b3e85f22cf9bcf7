"""Tensor-core lmfe kernels (csrc/feat_tc2.cuh, feat_tc_kernels.cu) against the CUDA-core FFT kernel
(spb_set_option("feature_kernel", 2) routes the same call through it) and the float64 oracle."""
import numpy as np
import pytest

from oracle import speechpy_oracle as o
from speaker_verification_b200 import _lib, batch, synth

pytestmark = pytest.mark.gpu
FS = 16000
TOL = 1e-4  # log-mel tolerance of the path (BASELINE.json north_star)


def _lmfe_dev(fb, arena, tc=True, **opts):
    import torch
    d_in = torch.from_numpy(arena).cuda()
    d_out = torch.full((max(fb.total_frames, 1), fb.num_filters), 777.0, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    with _lib.options(feature_kernel=0 if tc else 2, **opts):
        n0 = fb.ctx.launch_count()
        fb.lmfe_dev(d_in.data_ptr(), d_in.numel(), d_out.data_ptr())
        fb.ctx.sync()
        launches = fb.ctx.launch_count() - n0
    return d_out.cpu().numpy()[:fb.total_frames], launches


def _frames_to_samples(f):
    return 320 + 160 * f + 7  # floor((N - 320) / 160) == f


def test_tc_matches_fft_kernel_and_oracle_on_group_boundaries(sp):
    """Groups are 32 hops -> 31 frames, tiles four groups: every boundary case of the table."""
    frames = [1, 2, 30, 31, 32, 61, 62, 63, 93, 123, 124, 125, 248, 249, 1998, 0, 5]
    sigs = [synth.audio(1200 + i, _frames_to_samples(f), np.float32) for i, f in enumerate(frames)]
    arena, starts, lens = batch.pack(sigs)
    fb = batch.FeatureBatch(starts, lens)
    assert fb.total_frames == sum(frames)
    tc, n_tc = _lmfe_dev(fb, arena, True)
    fft, n_fft = _lmfe_dev(fb, arena, False)
    assert n_tc > n_fft  # tensor-core kernel + the guarded (no-op) FFT launch + its redo pass
    assert not np.any(tc == 777.0)  # every row written
    assert np.max(np.abs(tc - fft)) < 5e-5
    for s, got in zip(sigs, fb.split(tc)):
        ref = o.lmfe(s, FS)
        assert got.shape == ref.shape
        if ref.shape[0]:
            assert np.max(np.abs(got - ref)) < TOL
    fb.close()


def test_tc_many_tiles_per_cta(sp):
    """More tiles than SMs: the persistent loop, ring wrap-around and accumulator ping-pong -- and the
    north_star tolerance on EVERY value: all 4.3 M log-mel values of 128 cfg3-shaped utterances within
    1e-4 of the float64 oracle (feature.py:216-217,256), no energy mask.  The fp32 kernels alone cannot
    do that for the rare filter ~1e-5 below its frame's level; the float64 redo of the flagged filters
    (csrc/redo_kernels.cu) is what closes it, for the tensor-core path and the FFT path alike."""
    lengths = synth.cfg3_lengths(128, seed=21)
    arena, starts, lens = synth.ragged_corpus(lengths, seed=22)
    fb = batch.FeatureBatch(starts, lens)
    assert fb.total_frames > 3 * 148 * 124
    ref = np.concatenate([o.lmfe(arena[starts[u]:starts[u] + lens[u]], FS) for u in range(len(starts))])
    assert ref.shape == (fb.total_frames, 40)
    tc, _ = _lmfe_dev(fb, arena, True)
    fft, _ = _lmfe_dev(fb, arena, False)
    assert np.all(np.isfinite(tc))
    e_tc, e_fft = np.abs(tc - ref), np.abs(fft - ref)
    assert e_tc.max() < TOL, (e_tc.max(), np.unravel_index(e_tc.argmax(), e_tc.shape))
    assert e_fft.max() < TOL, (e_fft.max(), np.unravel_index(e_fft.argmax(), e_fft.shape))
    assert np.quantile(e_tc, 0.9999) < 2e-5
    # the guard is what does it: the same kernels without the redo pass show the tail (documented in DESIGN.md)
    raw_tc, _ = _lmfe_dev(fb, arena, True, redo=0)
    changed = raw_tc != tc
    assert 0 < changed.sum() < 2e-3 * tc.size       # a few values per 10^4 are recomputed
    assert np.max(np.abs(tc[changed] - ref[changed])) < 2e-6  # ... exactly (float32 rounding of the float64 value)
    # every epilogue flavour of the kernel and the previous generation agree
    for opts in (dict(tc_epilogue=0), dict(tc_epilogue=1), dict(tc_version=1)):
        alt, _ = _lmfe_dev(fb, arena, True, **opts)
        e_alt = np.abs(alt - ref)
        if "tc_version" in opts:
            assert np.quantile(e_alt, 0.9999) < 2e-5 and e_alt.max() < 1e-3, opts  # generation 1 has no guard
        else:
            assert e_alt.max() < TOL, opts
            assert np.max(np.abs(alt - tc)) < 2e-6, opts
    # the whole pipeline on top of it: every value of cmvnw(lmfe) for a sample of utterances
    full = fb.lmfe_cmvnw(arena)
    for u in (3, 17, 64, 127):
        s = arena[starts[u]:starts[u] + lens[u]]
        want = o.cmvnw_vectorised(o.lmfe(s, FS), 301, True)
        assert np.max(np.abs(full[fb.frame_offsets[u]:fb.frame_offsets[u + 1]] - want)) < TOL, u
    fb.close()


def test_tc_falls_back_for_samples_it_cannot_split(sp):
    """|x| >= 2, NaN and inf cannot be written as two fp16 halves of x * 2^15: the converter raises
    the flag and the guarded FFT launch redoes the group -> results identical to the FFT path."""
    base = [synth.audio(1300 + i, n, np.float32) for i, n in enumerate((20000, 48000, 5280))]
    cases = {
        "loud": [s * np.float32(50.0) for s in base],
        "int16-valued": [np.round(s * 32768).astype(np.float32) for s in base],
        "nan": [s.copy() for s in base],
        "inf": [s.copy() for s in base],
        "just-below-2": [np.clip(s * 30, -1.9990, 1.9990).astype(np.float32) for s in base],
    }
    cases["nan"][1][777] = np.nan
    cases["inf"][2][100] = np.inf
    for name, sigs in cases.items():
        arena, starts, lens = batch.pack(sigs)
        fb = batch.FeatureBatch(starts, lens)
        tc, _ = _lmfe_dev(fb, arena, True)
        fft, _ = _lmfe_dev(fb, arena, False)
        if name == "just-below-2":
            assert np.max(np.abs(tc - fft)) < 5e-5, name  # still the tensor-core result
        else:
            assert np.array_equal(tc, fft, equal_nan=True), name
        if name in ("loud", "int16-valued", "just-below-2"):
            for s, got in zip(sigs, fb.split(tc)):
                assert np.max(np.abs(got - o.lmfe(s, FS))) < TOL, name
        fb.close()


@pytest.mark.parametrize("kw", [
    dict(num_filters=26),                                   # not a multiple of four: scalar stores
    dict(num_filters=24, low_frequency=100),                # bank starts at another bin: table rotation
    dict(num_filters=40, high_frequency=4000),
    dict(num_filters=8, low_frequency=1000, high_frequency=6000),
    dict(num_filters=1),
])
def test_tc_other_filterbanks(sp, kw):
    sigs = [synth.audio(1400 + i, n, np.float32) for i, n in enumerate((30000, 5281, 64321, 480))]
    arena, starts, lens = batch.pack(sigs)
    fb = batch.FeatureBatch(starts, lens, **kw)
    tc, _ = _lmfe_dev(fb, arena, True)
    fft, _ = _lmfe_dev(fb, arena, False)
    assert np.max(np.abs(tc - fft)) < 5e-5
    okw = {k: v for k, v in kw.items()}
    for s, got in zip(sigs, fb.split(tc)):
        ref = o.lmfe(s, FS, **okw)
        assert got.shape == ref.shape
        assert np.max(np.abs(got - ref)) < TOL
    fb.close()


def test_tc_silence_and_gain(sp):
    """Digital silence hits functions.zero_handling (ln eps); a power-of-two gain shifts every
    log-mel value by 2 ln(gain) exactly in the split (scaling by 2^k is exact in fp16 and fp32)."""
    x = synth.audio(1500, 48000, np.float32)
    quiet = (x * np.float32(2.0 ** -6)).astype(np.float32)
    z = np.zeros(20000, dtype=np.float32)
    arena, starts, lens = batch.pack([x, quiet, z])
    fb = batch.FeatureBatch(starts, lens)
    tc, _ = _lmfe_dev(fb, arena, True)
    a, b, c = fb.split(tc)
    assert np.all(c == np.float32(-36.04365338911715))
    assert np.max(np.abs((a - b) - 12 * np.log(2.0))) < 2e-5
    assert np.max(np.abs(b - o.lmfe(quiet, FS))) < TOL
    fb.close()


def test_tc_plans_alternate(sp):
    """The epilogue tables live in one constant bank per device: alternating plans must re-upload."""
    sigs = [synth.audio(1600 + i, n, np.float32) for i, n in enumerate((30000, 64321))]
    arena, starts, lens = batch.pack(sigs)
    fa = batch.FeatureBatch(starts, lens)
    fb_ = batch.FeatureBatch(starts, lens, num_filters=24, low_frequency=100)
    a0, _ = _lmfe_dev(fa, arena, True)
    b0, _ = _lmfe_dev(fb_, arena, True)
    for _ in range(3):
        a1, _ = _lmfe_dev(fa, arena, True)
        b1, _ = _lmfe_dev(fb_, arena, True)
        assert np.array_equal(a0, a1) and np.array_equal(b0, b1)
    assert np.max(np.abs(fa.split(a0)[0] - o.lmfe(sigs[0], FS))) < TOL
    assert np.max(np.abs(fb_.split(b0)[0] - o.lmfe(sigs[0], FS, num_filters=24, low_frequency=100))) < TOL
    fa.close()
    fb_.close()


def _mfcc_delta_dev(fb, arena, tc=True):
    import torch
    d_in = torch.from_numpy(arena).cuda()
    d_out = torch.full((max(fb.total_frames, 1), fb.num_cepstral, 3), 777.0, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    with _lib.options(feature_kernel=0 if tc else 2):
        fb.mfcc_delta_dev(d_in.data_ptr(), d_in.numel(), d_out.data_ptr())
        fb.ctx.sync()
    return d_out.cpu().numpy()[:fb.total_frames]


@pytest.mark.parametrize("kw", [
    dict(),
    dict(dc_elimination=False),
    dict(num_filters=26, num_cepstral=20),
    dict(num_filters=24, low_frequency=100, num_cepstral=24),
])
def test_mfcc_delta_through_tensor_core_lmfe(sp, kw):
    """mfcc + delta cube = tensor-core log-mel rows + mfcc_post_kernel (Parseval frame energy from
    the samples, DCT-II, deltas) against the FFT kernel's fused epilogue and the oracle."""
    frames = [1, 31, 32, 62, 63, 124, 125, 798, 0, 1998]
    sigs = [synth.audio(1700 + i, _frames_to_samples(f), np.float32) for i, f in enumerate(frames)]
    arena, starts, lens = batch.pack(sigs)
    fb = batch.FeatureBatch(starts, lens, **kw)
    tc = _mfcc_delta_dev(fb, arena, True)
    fft = _mfcc_delta_dev(fb, arena, False)
    assert not np.any(tc == 777.0)
    assert np.max(np.abs(tc - fft)) < 5e-5
    host = fb.mfcc_delta(arena)  # host-pointer (chunked) call takes the same kernels
    assert np.array_equal(host, tc)
    okw = {k: v for k, v in kw.items()}
    for s, got in zip(sigs, fb.split(tc)):
        ref = o.extract_derivative_feature(o.mfcc(s, FS, **okw)) if len(s) >= 480 else None
        if ref is not None and ref.shape[0]:
            assert got.shape == ref.shape
            assert np.max(np.abs(got - ref)) < TOL
    fb.close()


def _device_noise(total, seed):
    import torch
    gen = torch.Generator(device="cuda")
    gen.manual_seed(seed)
    arena = torch.empty(total, dtype=torch.float32, device="cuda")
    for s in range(0, total, 1 << 28):
        arena[s:s + (1 << 28)].normal_(0.0, 0.1, generator=gen)
    return arena


def test_full_size_properties_cfg2(sp):
    """BASELINE config 2 at full size (4 874 x 8 s, mfcc 13 + delta cube, generated on the device):
    independence (an utterance alone == inside the batch, bit for bit), oracle spot checks, and gain:
    audio x 2^-3 moves only c0 = ln(energy) (by 2 ln 2^-3; the DCT rows k > 0 sum to zero and the
    reference's deltas never read coefficient 0)."""
    import torch
    U, N = 4874, 128000
    starts = np.arange(U, dtype=np.int64) * N
    lengths = np.full(U, N, dtype=np.int64)
    arena = _device_noise(U * N, 202)
    fb = batch.FeatureBatch(starts, lengths)
    assert fb.total_frames == U * 798
    out = torch.empty((fb.total_frames, 13, 3), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    fb.mfcc_delta_dev(arena.data_ptr(), arena.numel(), out.data_ptr())
    fb.ctx.sync()
    assert bool(torch.isfinite(out).all())
    for u in (0, 2500, U - 1):
        x = arena[starts[u]:starts[u] + N].cpu().numpy()
        one = batch.FeatureBatch(np.zeros(1, dtype=np.int64), np.array([N], dtype=np.int64))
        alone = one.mfcc_delta(x)
        one.close()
        got = out[u * 798:(u + 1) * 798].cpu().numpy()
        assert np.array_equal(got, alone), u
        ref = o.extract_derivative_feature(o.mfcc(x, FS))
        assert np.max(np.abs(got - ref)) < TOL, u
    arena.mul_(0.125)
    out2 = torch.empty_like(out)
    torch.cuda.synchronize()
    fb.mfcc_delta_dev(arena.data_ptr(), arena.numel(), out2.data_ptr())
    fb.ctx.sync()
    d = (out2 - out)
    assert float((d[:, 0, 0] - 2 * np.log(0.125)).abs().max()) < 5e-5
    d[:, 0, 0] = 0
    assert float(d.abs().max()) < 5e-5
    fb.close()


def test_full_size_properties_cfg4(sp):
    """BASELINE config 4 at full size (1 024 x 60 s, lmfe + cmvnw over 5 998-frame utterances: six
    large cmvnw tiles and 194 tensor-core groups each): independence and oracle spot checks."""
    import torch
    U, N = 1024, 960000
    starts = np.arange(U, dtype=np.int64) * N
    lengths = np.full(U, N, dtype=np.int64)
    arena = _device_noise(U * N, 404)
    fb = batch.FeatureBatch(starts, lengths)
    assert fb.total_frames == U * 5998
    out = torch.empty((fb.total_frames, 40), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    fb.lmfe_cmvnw_dev(arena.data_ptr(), arena.numel(), out.data_ptr())
    fb.ctx.sync()
    assert bool(torch.isfinite(out).all())
    for u in (0, 700, U - 1):
        x = arena[starts[u]:starts[u] + N].cpu().numpy()
        alone = batch.lmfe_cmvnw([x])[0]
        got = out[u * 5998:(u + 1) * 5998].cpu().numpy()
        assert np.array_equal(got, alone), u
        ref = o.cmvnw_vectorised(o.lmfe(x, FS), 301, True)
        assert np.max(np.abs(got - ref)) < TOL, u
    fb.close()


@pytest.mark.parametrize("kw", [dict(), dict(fft_length=1024, frame_length=0.025), dict(fft_length=256, frame_length=0.016, frame_stride=0.008)])
def test_mfcc_rows_of_the_fft_kernels_are_guarded(sp, kw):
    """mfcc + deltas straight out of the CUDA-core FFT kernels (feature_kernel=2; every shape the tensor-core kernel does
    not take ends up there): every value of 48 cfg3-shaped utterances within 1e-4 of the float64 reference
    (feature.py:102-153, 261-282).  A row in which some filter lies too far below the frame's level for float32 is
    listed by the kernel and recomputed whole in float64 (redo_kernels.cu: log-mel of all filters, DCT, ln(energy),
    deltas); pre-emphasis makes such rows common (the lowest filters drop 34 dB), so it is switched on here."""
    lengths = synth.cfg3_lengths(48, seed=31)
    arena, starts, lens = synth.ragged_corpus(lengths, seed=32)
    with _lib.options(feature_kernel=2):
        fb = batch.FeatureBatch(starts, lens, preemphasis=0.97, **kw)
        got = fb.mfcc_delta(arena)
        with _lib.options(redo=0):
            raw = fb.mfcc_delta(arena)
    worst = 0.0
    for u, g_ in enumerate(fb.split(got)):
        y = o.preemphasis(arena[starts[u]:starts[u] + lens[u]].astype(np.float64), 1, 0.97)
        ref = o.extract_derivative_feature(o.mfcc(y, FS, **kw))
        worst = max(worst, float(np.max(np.abs(g_ - ref))))
    assert worst < TOL, worst
    rows_changed = np.any(raw != got, axis=(1, 2))
    assert 0 < rows_changed.sum()  # the guard did recompute rows ...
    assert np.max(np.abs(raw - got)) < 5e-3  # ... that were close already
    fb.close()


def test_tc_preemphasis_in_the_epilogue(sp):
    """FeatureBatch(preemphasis=c) on the default shape: the tensor-core kernel applies it in its epilogue, in the
    spectral domain -- Y[k] = (1 - c W^k) X[k] - c x[s-1] + c x[s+319] W^(320 k) per frame, with the phase of the
    folded transform (feat_tc2.cuh, t2_pre_power) -- and the guard's threshold follows |H_k| per filter.  Every
    log-mel value of 96 cfg3-shaped utterances within 1e-4 of lmfe(preemphasis(x)) in float64 (processing.py:45-58,
    feature.py:222-258), the utterances' last samples loud so that a missed wrap (np.roll) shows; the CUDA-core
    FFT kernel (pre-emphasis on the staged tile) agrees; cmvnw on top of it; c < 0 as well."""
    lengths = synth.cfg3_lengths(96, seed=41)
    arena, starts, lens = synth.ragged_corpus(lengths, seed=42)
    for u in range(len(starts)):
        arena[starts[u] + lens[u] - 1] = 0.7
    for cof in (0.97, -0.5):
        fb = batch.FeatureBatch(starts, lens, preemphasis=cof)
        ref = np.concatenate([o.lmfe(o.preemphasis(arena[starts[u]:starts[u] + lens[u]].astype(np.float64), 1, cof), FS)
                              for u in range(len(starts))])
        tc, n_tc = _lmfe_dev(fb, arena, True)
        fft, n_fft = _lmfe_dev(fb, arena, False)
        assert n_tc > n_fft  # the tensor-core path: its kernel, the guarded fallback launch and two redo passes
        e_tc, e_fft = np.abs(tc - ref), np.abs(fft - ref)
        assert e_tc.max() < TOL, (cof, e_tc.max(), np.unravel_index(e_tc.argmax(), e_tc.shape))
        assert e_fft.max() < TOL, (cof, e_fft.max())
        assert np.quantile(e_tc, 0.9999) < 2e-5
        raw, _ = _lmfe_dev(fb, arena, True, redo=0)
        assert (raw != tc).mean() < 2e-3  # the per-filter threshold keeps the float64 redo rare
        if cof > 0:
            full = fb.lmfe_cmvnw(arena)
            for u in (0, 40, 95):
                want = o.cmvnw_vectorised(ref[fb.frame_offsets[u]:fb.frame_offsets[u + 1]], 301, True)
                assert np.max(np.abs(full[fb.frame_offsets[u]:fb.frame_offsets[u + 1]] - want)) < TOL, u
        fb.close()
