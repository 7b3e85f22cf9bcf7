"""Ragged-batch entry points (host and device pointers) against the oracle."""
import numpy as np
import pytest

from oracle import speechpy_oracle as o
from speaker_verification_b200 import batch, synth

pytestmark = pytest.mark.gpu
FS = 16000
TOL = 1e-4


def _signals(seed, lengths):
    return [synth.audio(seed + i, int(n), np.float32) for i, n in enumerate(lengths)]


def test_lmfe_cmvnw_ragged_batch(sp):
    lengths = [64000, 480, 35000, 320, 100321, 8000, 479, 20000, 70000]
    sigs = _signals(300, lengths)
    arena, starts, lens = batch.pack(sigs)
    fb = batch.FeatureBatch(starts, lens)
    out = fb.lmfe_cmvnw(arena)
    assert out.dtype == np.float32 and out.shape == (fb.total_frames, 40)
    parts = fb.split(out)
    for s, got in zip(sigs, parts):
        lm = o.lmfe(s, FS)
        assert got.shape == lm.shape
        if lm.shape[0]:
            ref = o.cmvnw_vectorised(lm, 301, True)
            assert np.max(np.abs(got - ref)) < TOL
    fb.close()


def test_unaligned_starts(sp):
    sigs = _signals(400, [5000, 7001, 6002, 9003])
    lens = np.array([len(s) for s in sigs], dtype=np.int64)
    starts = np.array([1, 5003, 12007, 18010], dtype=np.int64)
    arena = np.zeros(28000, dtype=np.float32)
    for s, st in zip(sigs, starts):
        arena[st:st + len(s)] = s
    fb = batch.FeatureBatch(starts, lens, variance_normalization=False)
    out = fb.split(fb.lmfe_cmvnw(arena))
    for s, got in zip(sigs, out):
        assert np.max(np.abs(got - o.cmvnw(o.lmfe(s, FS), 301, False))) < TOL
    fb.close()


def test_mfcc_delta_batch(sp):
    sigs = _signals(500, [128000, 16000, 640, 48000])
    arena, starts, lens = batch.pack(sigs)
    fb = batch.FeatureBatch(starts, lens)
    out = fb.mfcc_delta(arena)
    assert out.shape == (fb.total_frames, 13, 3)
    for s, got in zip(sigs, fb.split(out)):
        ref = o.extract_derivative_feature(o.mfcc(s, FS))
        assert np.max(np.abs(got - ref)) < TOL
    fb.close()


def test_device_pointer_path_and_launch_count(sp):
    import torch
    from speaker_verification_b200 import _lib
    lengths = synth.cfg3_lengths(64, seed=3)
    arena, starts, lens = synth.ragged_corpus(lengths, seed=33)
    fb = batch.FeatureBatch(starts, lens)
    d_in = torch.from_numpy(arena).cuda()
    d_out = torch.empty((fb.total_frames, 40), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    n0 = fb.ctx.launch_count()
    fb.lmfe_cmvnw_dev(d_in.data_ptr(), d_in.numel(), d_out.data_ptr())
    fb.ctx.sync()
    assert fb.ctx.launch_count() - n0 >= 3
    got = d_out.cpu().numpy()
    host = fb.lmfe_cmvnw(arena)
    assert np.array_equal(got, host)  # same kernels, same order: bit-identical
    for u in (0, 17, 63):
        s = arena[starts[u]:starts[u] + lens[u]]
        ref = o.cmvnw_vectorised(o.lmfe(s, FS), 301, True)
        assert np.max(np.abs(got[fb.frame_offsets[u]:fb.frame_offsets[u + 1]] - ref)) < TOL
    # property at size: every utterance's columns are ~zero-mean locally => global column means small
    assert np.all(np.isfinite(got))
    fb.close()


def test_batch_errors(sp):
    with pytest.raises(ValueError):
        batch.FeatureBatch(np.array([0]), np.array([100]))
    with pytest.raises(AssertionError):
        batch.FeatureBatch(np.array([0]), np.array([4000]), win_size=300)
    fb = batch.FeatureBatch(np.array([0]), np.array([4000]))
    with pytest.raises(AssertionError):
        fb.lmfe_cmvnw(np.zeros(100, dtype=np.float32))
    fb.close()


def test_large_tile_cmvnw_matches_small_tile_kernels(sp):
    """The batch path runs the large-tile cmvnw kernels; bit 1 of the debug hook routes the
    same call through the small-tile kernels (independent tiling, block size, sqrt/div)."""
    from speaker_verification_b200 import _lib
    # multi-tile (> 960 frames), exactly-one-tile, short (< 150 frames: repeated reflection), tiny
    lengths = [320000, 154000, 153920, 100321, 20000, 1000, 480, 800, 235000]
    sigs = _signals(700, lengths)
    arena, starts, lens = batch.pack(sigs)
    for vn in (True, False):
        fb = batch.FeatureBatch(starts, lens, variance_normalization=vn)
        big = fb.lmfe_cmvnw(arena)
        with _lib.options(cmvnw_kernel=1):
            small = fb.lmfe_cmvnw(arena)
        assert np.max(np.abs(big - small)) < 2e-5
        for s, got in zip(sigs, fb.split(big)):
            lm = o.lmfe(s, FS)
            if lm.shape[0]:
                assert np.max(np.abs(got - o.cmvnw_vectorised(lm, 301, vn))) < TOL
        fb.close()
    # public API, other window sizes / widths through both kernel families
    rng = np.random.Generator(np.random.Philox(5))
    for rows, cols, win in [(2500, 40, 301), (700, 24, 101), (64, 40, 31), (5, 8, 301), (1500, 13, 301)]:
        v = rng.standard_normal((rows, cols)) * 3 + 1
        got = sp.processing.cmvnw(v, win, True)
        assert np.max(np.abs(got - o.cmvnw_vectorised(v, win, True))) < TOL, (rows, cols, win)


def test_int16_pcm_arena_matches_astype_float(sp):
    """An int16 array handed to lmfe is widened by signal.astype(float) (feature.py:182); the
    batch call ingests the PCM arena as int16 (sample_scale 1) and must equal that.  With
    sample_scale 1/32768 it must equal the float32 audio librosa.load returns (utils.py:170-173)."""
    rng = np.random.Generator(np.random.Philox(11))
    lengths = [48000, 16001, 480, 320, 99999]
    sigs = [np.clip(np.round(rng.standard_normal(n) * 3000), -32768, 32767).astype(np.int16) for n in lengths]
    arena, starts, lens = batch.pack(sigs, dtype=np.int16)
    assert arena.dtype == np.int16
    fb = batch.FeatureBatch(starts, lens)
    got16 = fb.lmfe_cmvnw(arena)
    got32 = fb.lmfe_cmvnw(arena.astype(np.float32))
    got64 = fb.lmfe_cmvnw(arena.astype(np.float64))
    assert np.array_equal(got16, got32) and np.array_equal(got16, got64)  # widening is exact
    for s, got in zip(sigs, fb.split(got16)):
        lm = o.lmfe(s.astype(float), FS)
        if lm.shape[0] == 1:
            # one frame: every window is 301 copies of the row, the true result is 0.  The
            # reference returns its float64 summation noise (~1e-13 at |lmfe| ~ 16) divided by
            # (0 + 2^-30), i.e. O(1e-4) garbage; the device result is exactly 0.
            assert np.max(np.abs(got)) == 0 and np.max(np.abs(o.cmvnw_vectorised(lm, 301, True))) < 1e-3
        elif lm.shape[0]:
            assert np.max(np.abs(got - o.cmvnw_vectorised(lm, 301, True))) < TOL
    # librosa-style ingest: int16 * 2^-15 on the device == float32 arena scaled on the host
    scaled = arena.astype(np.float32) * np.float32(batch.PCM16_SCALE)
    got_l = fb.lmfe_cmvnw(arena, sample_scale=batch.PCM16_SCALE)
    assert np.array_equal(got_l, fb.lmfe_cmvnw(scaled))
    for s, got in zip(sigs, fb.split(got_l)):
        lm = o.lmfe(s.astype(np.float32) * np.float32(batch.PCM16_SCALE), FS)
        if lm.shape[0] > 1:
            assert np.max(np.abs(got - o.cmvnw_vectorised(lm, 301, True))) < TOL
    mf = fb.mfcc_delta(arena)
    for s, got in zip(sigs, fb.split(mf)):
        if not got.size:
            continue
        ref = o.extract_derivative_feature(o.mfcc(s.astype(float), FS))
        assert np.max(np.abs(got - ref)) < 2e-4  # |values| ~ 20x larger than for N(0, 0.1) audio
    fb.close()


def test_pipelined_host_path_many_chunks(sp):
    """Host-buffer calls overlap H2D / kernels / D2H chunk by chunk (chunks = runs of whole
    utterances of at most 192 Ki frames).  ~80 min of audio = 3 chunks; result must equal the
    device-pointer path bit for bit, for float32 and int16 arenas, and the oracle on samples."""
    import torch
    rng = np.random.Generator(np.random.Philox(12))
    lengths = rng.integers(64000, 320000, size=400)
    pcm = [np.clip(np.round(rng.standard_normal(int(n)) * 2000), -32768, 32767).astype(np.int16) for n in lengths]
    arena16, starts, lens = batch.pack(pcm, dtype=np.int16)
    arena32 = arena16.astype(np.float32)
    fb = batch.FeatureBatch(starts, lens)
    assert fb.total_frames > 2 * 192 * 1024
    host32 = fb.lmfe_cmvnw(arena32)
    host16 = fb.lmfe_cmvnw(arena16)
    d_arena = torch.from_numpy(arena32).cuda()
    d_out = torch.empty((fb.total_frames, 40), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    fb.lmfe_cmvnw_dev(d_arena.data_ptr(), d_arena.numel(), d_out.data_ptr())
    fb.ctx.sync()
    dev = d_out.cpu().numpy()
    assert np.array_equal(host32, dev) and np.array_equal(host16, dev)
    parts = fb.split(host16)
    for u in (0, 97, 250, 399):
        ref = o.cmvnw_vectorised(o.lmfe(pcm[u].astype(float), FS), 301, True)
        assert np.max(np.abs(parts[u] - ref)) < TOL
    # utterances not packed in order (chunk spans overlap): whole-arena copy fallback
    perm = rng.permutation(len(lens))
    fb2 = batch.FeatureBatch(starts[perm], lens[perm])
    out2 = fb2.split(fb2.lmfe_cmvnw(arena16))
    for k in (0, 11, 200):
        assert np.array_equal(out2[k], parts[perm[k]])
    fb.close(); fb2.close()


def test_feature_cubes_match_reference_transforms(sp):
    """utils.CMVN + FeatureCube / FeatureCube3C (utils.py:325-397) on the device.
    Golden = the reference's transforms run on the reference's float64 lmfe with seeded
    np.random crops; tolerance 1e-4 (float32 lmfe; normalised values are O(1))."""
    from conftest import golden
    g = golden("feature_cubes.npz")
    sig = synth.audio(int(g["seed"]), int(g["n"]), np.float64)
    others = _signals(900, [30000, 16160, 52345])
    sigs = [others[0], sig.astype(np.float32), others[1], others[2]]
    arena, starts, lens = batch.pack(sigs)
    fb = batch.FeatureBatch(starts, lens)
    F = np.diff(fb.frame_offsets)
    rs = np.random.RandomState(5)
    for deriv in (False, True):
        for norm in (False, True):
            tag = f"d{int(deriv)}n{int(norm)}"
            crops = np.stack([rs.randint(int(f) - 80, size=20) for f in F]).astype(np.int32)
            crops[1] = g[f"idx_{tag}"]
            out = fb.feature_cubes(arena, crops, derivative=deriv, normalize=norm)
            assert out.shape == (4, 3 if deriv else 1, 20, 80, 40) and out.dtype == np.float32
            assert np.max(np.abs(out[1][:, :3] - g[f"cube_{tag}"])) < TOL, tag
            for u, s in enumerate(sigs):      # every utterance against the oracle transforms
                feat = o.cmvn_transform(o.lmfe(s, FS), deriv, norm)
                assert np.max(np.abs(out[u] - o.feature_cube(feat, crops[u], 80))) < TOL, (tag, u)
    # reference-style RNG use: np.random.seed + one randint call per utterance, in order
    np.random.seed(123)
    auto = fb.draw_crop_starts()
    np.random.seed(123)
    assert np.array_equal(auto, np.stack([np.random.randint(int(f) - 80, size=20) for f in F]))
    # last admissible start is F - 80; one past it is rejected
    edge = np.tile((F - 80).astype(np.int32)[:, None], (1, 2))
    fb.feature_cubes(arena, edge)
    edge[2, 1] += 1
    with pytest.raises(Exception):
        fb.feature_cubes(arena, edge)
    # int16 PCM arena with the librosa scale gives the same cubes as the scaled float arena
    pcm = [synth.pcm16(40 + i, n) for i, n in enumerate([30000, 40000])]
    a16, st, ln = batch.pack(pcm, dtype=np.int16)
    fb2 = batch.FeatureBatch(st, ln)
    cr = fb2.draw_crop_starts(4, 80, rng=np.random.RandomState(1))
    c16 = fb2.feature_cubes(a16, cr, derivative=True, normalize=True, sample_scale=batch.PCM16_SCALE)
    c32 = fb2.feature_cubes(a16.astype(np.float32) * np.float32(batch.PCM16_SCALE), cr, derivative=True, normalize=True)
    assert np.array_equal(c16, c32)
    fb.close(); fb2.close()


def test_cube_collate_is_the_reference_loader_batch(sp):
    """one collate call == the reference's per-utterance chain
    speech.lmfe(25 ms, NFFT 1024) -> CMVN -> FeatureCube[3C] -> ToTensor stacked by the
    DataLoader (load_data.py:61-68, train.py:194-226), with the reference's RNG use."""
    import torch
    lengths = [48000, 20000, 33333, 16400 + 400]
    sigs = _signals(1200, lengths)
    labels = [3, 0, 7, 3]
    for deriv, norm in ((False, False), (True, True)):
        coll = batch.CubeCollate((80, 40, 20), derivative=deriv, normalize=norm)
        np.random.seed(99)
        feats, lab = coll([{"signal": s, "label": l} for s, l in zip(sigs, labels)])
        assert feats.is_cuda and feats.dtype == torch.float32 and lab.dtype == torch.int64
        assert tuple(feats.shape) == (4, 3 if deriv else 1, 20, 80, 40) and lab.tolist() == labels
        np.random.seed(99)
        for u, s in enumerate(sigs):
            lm = o.lmfe(s, FS, frame_length=0.025, frame_stride=0.01, num_filters=40, fft_length=1024)
            idx = np.random.randint(lm.shape[0] - 80, size=20)
            ref = o.feature_cube(o.cmvn_transform(lm, deriv, norm), idx, 80)
            assert np.max(np.abs(feats[u].cpu().numpy() - ref)) < TOL, (deriv, norm, u)
        # numpy flavour, tuple samples, same RNG state -> same cubes
        np.random.seed(99)
        f2, l2 = batch.CubeCollate((80, 40, 20), derivative=deriv, normalize=norm, as_torch=False)(
            list(zip(sigs, labels)))
        assert np.array_equal(f2, feats.cpu().numpy()) and l2.dtype == np.int64
    # int16 PCM with the librosa scale == the scaled float signals
    pcm = [synth.pcm16(70 + i, n) for i, n in enumerate([30000, 25000])]
    rs = np.random.RandomState(4)
    a, _ = batch.CubeCollate(sample_scale=batch.PCM16_SCALE, rng=rs)([(p, 1) for p in pcm])
    rs = np.random.RandomState(4)
    b, _ = batch.CubeCollate(rng=rs)([(p.astype(np.float32) * np.float32(batch.PCM16_SCALE), 1) for p in pcm])
    assert torch.equal(a, b)
    with pytest.raises(ValueError):  # 80 frames or fewer: the reference's randint raises too
        batch.CubeCollate()([(sigs[0][:8000], 0)])


def test_full_size_properties_cfg3_scale(sp):
    """Size-independent properties at benchmark scale (20 000 cfg3-shaped utterances, 45 audio-h,
    10 GB arena generated on the device, several launch groups):
      * independence - an utterance inside the big batch == the same utterance featurised alone,
        bit for bit (first / last utterance of the batch and of launch groups, longest, shortest);
      * gain invariance - scaling the audio by 2^-3 (exact in float32) moves lmfe by a constant
        that cmvnw removes: outputs agree to 2e-5;
      * oracle spot checks on the same utterances (1e-4)."""
    import torch
    lengths = synth.cfg3_lengths(20000, seed=3)
    starts = np.zeros(len(lengths), dtype=np.int64)
    np.cumsum((lengths[:-1] + 3) // 4 * 4, out=starts[1:])
    total = int(starts[-1] + lengths[-1])
    gen = torch.Generator(device="cuda")
    gen.manual_seed(99)
    arena = torch.empty(total, dtype=torch.float32, device="cuda")
    for s in range(0, total, 1 << 28):
        arena[s:s + (1 << 28)].normal_(0.0, 0.1, generator=gen)
    fb = batch.FeatureBatch(starts, lengths)
    assert fb.total_frames > 3 * 4 * 1024 * 1024          # more than three 4 Mi-row launch groups
    out = torch.empty((fb.total_frames, 40), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    fb.lmfe_cmvnw_dev(arena.data_ptr(), total, out.data_ptr())
    fb.ctx.sync()
    assert bool(torch.isfinite(out).all())
    fo = fb.frame_offsets
    group_edge = int(np.searchsorted(fo, 4 * 1024 * 1024))  # utterance straddling the first group boundary
    picks = sorted({0, len(lengths) - 1, group_edge - 1, group_edge, int(np.argmax(lengths)), int(np.argmin(lengths)), 7777})
    for u in picks:
        x = arena[starts[u]:starts[u] + lengths[u]].cpu().numpy()
        alone = batch.lmfe_cmvnw([x])[0]
        got = out[fo[u]:fo[u + 1]].cpu().numpy()
        assert np.array_equal(got, alone), u
        ref = o.cmvnw_vectorised(o.lmfe(x, FS), 301, True)
        assert np.max(np.abs(got - ref)) < TOL, u
    arena.mul_(0.125)
    out2 = torch.empty_like(out)
    torch.cuda.synchronize()
    fb.lmfe_cmvnw_dev(arena.data_ptr(), total, out2.data_ptr())
    fb.ctx.sync()
    assert float((out - out2).abs().max()) < 2e-5
    fb.close()


def test_feature_cubes_many_utterances(sp):
    """More utterances than a CUDA grid dimension (65535): the gather kernel strides over them."""
    import torch
    U, n = 70000, 13440                      # 82 frames each
    starts = np.arange(U, dtype=np.int64) * n
    lens = np.full(U, n, dtype=np.int64)
    gen = torch.Generator(device="cuda")
    gen.manual_seed(3)
    arena = torch.empty(U * n, dtype=torch.float32, device="cuda").normal_(0.0, 0.1, generator=gen)
    fb = batch.FeatureBatch(starts, lens)
    crops = np.random.RandomState(2).randint(82 - 80, size=(U, 2)).astype(np.int32)
    out = torch.empty((U, 3, 2, 80, 40), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    fb.feature_cubes_dev(arena.data_ptr(), arena.numel(), out.data_ptr(), crops, derivative=True, normalize=True)
    fb.ctx.sync()
    for u in (0, 65534, 65535, 65536, U - 1):
        x = arena[u * n:(u + 1) * n].cpu().numpy()
        feat = o.cmvn_transform(o.lmfe(x, FS), True, True)
        assert np.max(np.abs(out[u].cpu().numpy() - o.feature_cube(feat, crops[u], 80))) < TOL, u
    fb.close()


def test_single_pass_cmvnw_matches_two_pass_kernels_and_oracle(sp):
    """Option cmvnw_kernel=3: cmvnw as ONE streaming pass (csrc/cmvn_stream_kernels.cu: rings of lmfe and
    mean-subtracted rows in shared memory, super-steps of 180 rows; reads every row once and writes it
    once, but measured slower than the two-pass kernels, so it is opt-in).  Lengths chosen around every
    boundary of its schedule: < 150 rows (stays with the two-pass kernels), 150 / 151 (minimum), the
    super-step 180, 330 = 180 + 150, ring sizes 525 / 660, many super-steps, and > 3 000 rows (cut into
    segments that re-derive their context).  Checked against the large-tile two-pass kernels (the
    default) and, on every value, against processing.cmvnw in float64."""
    from speaker_verification_b200 import _lib
    rows = [149, 150, 151, 164, 165, 179, 180, 181, 195, 299, 300, 301, 329, 330, 331, 345, 359, 360, 361,
            524, 525, 526, 540, 659, 660, 661, 720, 900, 1998, 3000, 3001, 4801, 5998, 1, 31, 7205,
            2, 29, 30, 59, 60, 61, 74, 75, 76, 929, 930, 931, 899, 901,
            # 931 .. 1 860 rows: with option cmvnw_kernel=4 a cluster of two CTAs per utterance, halves exchanged over DSMEM
            932, 959, 960, 961, 1200, 1499, 1500, 1501, 1830, 1859, 1860, 1861]
    lengths = [320 + 160 * r + (7 * i) % 160 for i, r in enumerate(rows)]
    sigs = _signals(900, lengths)
    # digital silence inside an utterance: constant log-mel rows -> zero variance, the reference gives 0 / (0 + eps) = 0
    sigs[28][40000:140000] = 0.0
    arena, starts, lens = batch.pack(sigs)
    fb = batch.FeatureBatch(starts, lens)
    assert [int(f) for f in np.diff(fb.frame_offsets)] == rows
    # the default: utterances of <= 930 rows resident in one CTA's shared memory (both stages in one pass, stage 1
    # in place: csrc/cmvn_resident_kernels.cu), longer ones two-pass; option 4: up to 1 860 rows in a 2-CTA
    # cluster's shared memory (correct, measured slower than two passes: opt-in), only the longer ones two-pass
    with _lib.options(cmvnw_kernel=2):
        two = fb.lmfe_cmvnw(arena)
    refs = [o.cmvnw_vectorised(o.lmfe(s_, FS), 301, True) for s_ in sigs]
    for opt in (0, 4):
        with _lib.options(cmvnw_kernel=opt):
            res = fb.lmfe_cmvnw(arena)
        assert np.all(np.isfinite(res)), opt
        assert np.max(np.abs(res - two)) < 2e-5, opt
        for ref, got in zip(refs, fb.split(res)):
            if ref.shape[0]:
                assert float(np.max(np.abs(got - ref))) < TOL, (opt, ref.shape)
        assert np.max(np.abs(fb.split(res)[28][400:700])) < 1e-5  # the silent stretch
    with _lib.options(cmvnw_kernel=3):
        one = fb.lmfe_cmvnw(arena)
        # a NaN sample poisons the rows whose windows touch it, as in the two-pass kernels
        bad = arena.copy()
        bad[starts[28] + 200000] = np.nan
        one_nan = fb.split(fb.lmfe_cmvnw(bad))[28]
    assert np.all(np.isfinite(one))
    assert np.max(np.abs(one - two)) < 2e-5
    worst = 0.0
    for ref, got in zip(refs, fb.split(one)):
        worst = max(worst, float(np.max(np.abs(got - ref))))
    assert worst < TOL, worst
    sil = fb.split(one)[28][400:700]
    assert np.max(np.abs(sil)) < 1e-5
    f_nan = 200000 // 160
    nan_rows = np.flatnonzero(np.isnan(one_nan).any(axis=1))
    assert nan_rows.min() >= f_nan - 1 - 300 and nan_rows.max() <= f_nan + 300 + 29
    assert np.isnan(one_nan[f_nan - 290:f_nan + 290]).all()
    fb.close()


def test_cmvnw_random_lengths_resident_and_two_pass(sp):
    """160 utterances of random length (1 .. 1 900 frames: around the resident kernel's 930-row limit, the 2-CTA
    cluster range up to 1 860 of option 4 and every 30-row block phase) in one batch: the default dispatch
    (resident single pass / two-pass) and option 4 (resident / cluster pairs / two-pass) against the two-pass
    kernels alone and, on every value, against processing.cmvnw(feature.lmfe) of the reference."""
    from speaker_verification_b200 import _lib
    rng = np.random.Generator(np.random.Philox(20261105))
    rows = np.concatenate([rng.integers(1, 1101, size=110), rng.integers(900, 961, size=20), rng.integers(931, 1900, size=30)])
    lengths = [int(320 + 160 * r + rng.integers(0, 160)) for r in rows]
    sigs = _signals(5000, lengths)
    arena, starts, lens = batch.pack(sigs)
    fb = batch.FeatureBatch(starts, lens)
    assert [int(f) for f in np.diff(fb.frame_offsets)] == [int(r) for r in rows]
    with _lib.options(cmvnw_kernel=2):
        two = fb.lmfe_cmvnw(arena)
    refs = [o.cmvnw_vectorised(o.lmfe(s_, FS), 301, True) for s_ in sigs]
    for opt in (0, 4):
        with _lib.options(cmvnw_kernel=opt):
            got = fb.lmfe_cmvnw(arena)
        assert np.all(np.isfinite(got)), opt
        assert np.max(np.abs(got - two)) < 2e-5, opt
        for ref, g_, r in zip(refs, fb.split(got), rows):
            if r < 2:
                continue  # a single frame: the reference's own output is rounding noise over 2^-30
            assert np.max(np.abs(g_ - ref)) < TOL, (opt, int(r))
    fb.close()


@pytest.mark.parametrize("kw", [dict(win_size=101), dict(win_size=31, num_filters=24), dict(win_size=301, num_filters=26),
                                dict(win_size=301, variance_normalization=False), dict(win_size=601),
                                dict(win_size=11, variance_normalization=False, num_filters=64, fft_length=1024,
                                     frame_length=0.025)])
def test_batch_cmvnw_other_windows_and_widths(sp, kw):
    """The resident / streaming single-pass kernels take only the default shape (window 301, 40 columns,
    variance normalisation); every other plan must fall through to the two-pass kernels and still match
    processing.cmvnw(feature.lmfe(...)) of the reference."""
    # (no single-frame utterance here: a constant window makes the reference's own output rounding noise of its
    # float64 mean, amplified by 1 / 2^-30 -- 1e-4 at window 601; the ragged-batch test covers F = 1 at window 301)
    lengths = [64000, 20000, 100321, 8000, 800, 154000]
    sigs = _signals(1300, lengths)
    arena, starts, lens = batch.pack(sigs)
    fb = batch.FeatureBatch(starts, lens, **kw)
    out = fb.lmfe_cmvnw(arena)
    lkw = {k: v for k, v in kw.items() if k in ("num_filters", "fft_length", "frame_length")}
    for s_, got in zip(sigs, fb.split(out)):
        lm = o.lmfe(s_, FS, **lkw)
        assert got.shape == lm.shape
        if lm.shape[0]:
            ref = o.cmvnw_vectorised(lm, kw.get("win_size", 301), kw.get("variance_normalization", True))
            assert np.max(np.abs(got - ref)) < TOL, kw
    fb.close()


def test_nan_samples_poison_the_window_like_the_reference(sp):
    """A NaN sample makes its frames NaN and, through the window mean / std, every cmvnw row whose
    window touches them (processing.py:302-322) -- in every column (the block sums convert half of
    the columns with integer instructions, and the variance clamp is a select: both keep NaN).
    Known deviation: the sliding sums are re-anchored every 30 rows (15 in the small-tile kernels),
    so a NaN that has left the window can poison up to 29 further rows; never fewer rows than the
    reference."""
    x = synth.audio(5, 64000, np.float32)
    x[32000] = np.nan
    arena, starts, lens = batch.pack([x])
    lm = o.lmfe(x, FS)
    assert np.isnan(lm).any(axis=1).sum() == 2
    for vn in (True, False):
        ref = o.cmvnw_vectorised(lm, 301, vn)
        fb = batch.FeatureBatch(starts, lens, variance_normalization=vn)
        for got in (fb.lmfe_cmvnw(arena), sp.processing.cmvnw(lm, 301, vn)):  # large- / small-tile kernels
            gn, rn = np.isnan(got), np.isnan(ref)
            assert np.all(gn[rn])                                     # every reference NaN is a NaN here
            extra = (gn & ~rn).any(axis=1)
            assert extra.sum() <= 29 and (not extra.any() or np.flatnonzero(extra).min() > np.flatnonzero(rn.any(axis=1)).max())
            ok = ~gn
            if ok.any():
                assert np.max(np.abs(got[ok] - ref[ok])) < TOL
        fb.close()
    assert np.isnan(o.cmvnw_vectorised(lm, 301, True)).all()          # variance mode: +-300 rows cover this utterance


def _lmfe_dev(fb, arena):
    import torch
    d = torch.from_numpy(arena).cuda()
    out = torch.full((fb.total_frames, fb.num_filters), float("nan"), dtype=torch.float32, device="cuda")
    fb.lmfe_dev(d.data_ptr(), d.numel(), out.data_ptr())
    fb.ctx.sync()
    return out.cpu().numpy()


@pytest.mark.parametrize("kw", [dict(), dict(fft_length=1024, frame_length=0.025), dict(fft_length=256, frame_length=0.016, frame_stride=0.008, num_filters=26)])
def test_preemphasis_fused_into_the_framing(sp, kw):
    """FeatureBatch(preemphasis=0.98): processing.preemphasis(signal, shift=1, cof) (processing.py:45-58) as a
    prologue of the framing kernels -- every utterance's y[i] = x[i] - cof x[i-1], y[0] wrapping to its LAST sample
    (np.roll), formed while the tile is staged in shared memory.  Against the reference chain
    lmfe(preemphasis(x)) / cmvnw / mfcc + deltas in float64, on the three FFT kernels (NFFT 512, 1024, generic),
    with unaligned starts (plain loads instead of the bulk copy), a loud last sample (so that a missed wrap
    shows) and lengths around the 32-frame tile boundary."""
    cof = 0.98
    L = int(round(FS * kw.get("frame_length", 0.020)))
    S = int(round(FS * kw.get("frame_stride", 0.010)))
    lengths = [L, L + S - 1, L + S, L + 31 * S, L + 32 * S, L + 33 * S + 7, 64000, 35001, 20002, 100003]
    sigs = _signals(7100, lengths)
    for s_ in sigs:
        s_[-1] = 0.9  # the sample y[0] wraps to
    refs = [o.preemphasis(s_.astype(np.float64), 1, cof) for s_ in sigs]
    okw = {k: v for k, v in kw.items()}
    for aligned in (True, False):
        if aligned:
            arena, starts, lens = batch.pack(sigs)
        else:
            lens = np.array(lengths, dtype=np.int64)
            starts = np.cumsum([3] + [n + 5 for n in lengths[:-1]]).astype(np.int64)
            arena = np.zeros(int(starts[-1] + lens[-1]) + 8, dtype=np.float32)
            for s_, st in zip(sigs, starts):
                arena[st:st + len(s_)] = s_
        fb = batch.FeatureBatch(starts, lens, preemphasis=cof, **kw)
        plain = batch.FeatureBatch(starts, lens, **kw)
        lm = fb.split(_lmfe_dev(fb, arena))
        cm = fb.split(fb.lmfe_cmvnw(arena))
        mf = fb.split(fb.mfcc_delta(arena))
        lm_plain = fb.split(_lmfe_dev(plain, arena))
        if not kw:
            # the default shape with 16-byte aligned starts runs the tensor-core kernel (pre-emphasis in its epilogue,
            # spectral form); the same call through the CUDA-core FFT kernel must agree with it and with the reference
            from speaker_verification_b200 import _lib
            with _lib.options(feature_kernel=2):
                lm_fft = fb.split(_lmfe_dev(fb, arena))
            for y, a, f_ in zip(refs, lm, lm_fft):
                if a.shape[0]:
                    assert np.max(np.abs(f_ - o.lmfe(y, FS))) < TOL
                    assert np.max(np.abs(a - f_)) < TOL
        for y, a, b_, c_, d_ in zip(refs, lm, cm, mf, lm_plain):
            ref = o.lmfe(y, FS, **okw)
            assert a.shape == ref.shape
            if not ref.shape[0]:
                continue
            assert np.max(np.abs(a - ref)) < TOL, (aligned, len(y))
            assert np.max(np.abs(a - d_)) > 0.1  # it is not the un-emphasised signal
            assert np.max(np.abs(b_ - o.cmvnw_vectorised(ref, 301, True))) < TOL, (aligned, len(y))
            # mfcc rows of the FFT kernels are guarded like the log-mel values: a row with a filter too far below
            # the frame's level for float32 (here: the one-bin filter next to DC that pre-emphasis pushes 34 dB down
            # in the 26-filter bank over 129 bins) is recomputed in float64 (csrc/redo_kernels.cu)
            assert np.max(np.abs(c_ - o.extract_derivative_feature(o.mfcc(y, FS, **okw)))) < TOL, (aligned, len(y))
        fb.close()
        plain.close()


def test_launch_groups_and_pipeline_chunks_do_not_change_results(sp, monkeypatch):
    """The device-pointer calls walk a batch in launch groups (SPB_GROUP_ROWS rows, default 32 Mi) and the host-pointer
    calls in pipeline chunks (SPB_CHUNK_ROWS, default 192 Ki): sizing knobs read when a batch descriptor is created,
    which only cut the batch at utterance boundaries.  The same ragged batch with tiny groups / chunks (dozens of
    launches, every group with its own resident / two-pass / tensor-core tables) gives bit-identical rows, through
    both kinds of call, and matches the reference."""
    import torch
    lengths = synth.cfg3_lengths(40, seed=51)
    arena, starts, lens = synth.ragged_corpus(lengths, seed=52)
    d = torch.from_numpy(arena).cuda()

    def run():
        fb = batch.FeatureBatch(starts, lens)
        out = torch.full((fb.total_frames, 40), float("nan"), dtype=torch.float32, device="cuda")
        n0 = fb.ctx.launch_count()
        fb.lmfe_cmvnw_dev(d.data_ptr(), d.numel(), out.data_ptr())
        fb.ctx.sync()
        launches = fb.ctx.launch_count() - n0
        host = fb.lmfe_cmvnw(arena)
        mf = fb.mfcc_delta(arena)
        res = (out.cpu().numpy(), host, mf, launches, fb.frame_offsets.copy())
        fb.close()
        return res

    dev0, host0, mf0, l0, offs = run()
    monkeypatch.setenv("SPB_GROUP_ROWS", "3000")
    monkeypatch.setenv("SPB_CHUNK_ROWS", "2000")
    dev1, host1, mf1, l1, _ = run()
    assert l1 > 4 * l0  # many groups
    assert np.array_equal(dev0, dev1) and np.array_equal(host0, host1) and np.array_equal(mf0, mf1)
    assert np.array_equal(dev0, host0)
    for u in (0, 13, 39):
        s_ = arena[starts[u]:starts[u] + lens[u]]
        ref = o.cmvnw_vectorised(o.lmfe(s_, FS), 301, True)
        assert np.max(np.abs(dev1[offs[u]:offs[u + 1]] - ref)) < TOL
