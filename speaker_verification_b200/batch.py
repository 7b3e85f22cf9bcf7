"""Ragged utterance batches: packing, length bucketing, sharding, fused pipelines.

This is the batched form of what ``load_data.AudioDataset.__getitem__`` does one
utterance at a time (load_data.py:50-70: ``speech.lmfe`` per file) plus the
``processing.cmvnw`` post-processing of the speechpy examples
(speech_feature_extraction/example/test_package.py:36-37).  One call featurises a
whole ragged batch on the GPU.

Multi-GPU: utterances are independent, so ranks take disjoint shards balanced by
total samples and no collective is needed.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .speechpy import feature as _feature
from .speechpy import processing as _processing


PCM16_SCALE = 1.0 / 32768.0  # librosa / soundfile scaling of 16-bit PCM to [-1, 1)

_ARENA_DTYPES = {np.dtype(np.float32): _lib.F32, np.dtype(np.float64): _lib.F64,
                 np.dtype(np.int16): _lib.I16, np.dtype(np.int32): _lib.I32}


def _as_arena(arena):
    """Contiguous arena in a dtype the library ingests + its spb_dtype code.  int16 PCM stays
    int16: it crosses PCIe at 2 bytes per sample and is widened (and scaled) on the device."""
    arena = np.ascontiguousarray(arena)
    if arena.dtype not in _ARENA_DTYPES:
        arena = arena.astype(np.float32)
    return arena, _ARENA_DTYPES[arena.dtype]


def pack(signals, align: int = 8, dtype=np.float32):
    """Concatenate 1-D signals into one arena of ``dtype`` (float32, or int16 for PCM).

    Returns (arena, starts, lengths).  Every utterance starts on a multiple of
    ``align`` samples: 4 (16 bytes) is what the tensor-core / TMA paths need, the default 8 (32 bytes) makes the
    64-byte runs their converters fetch cover whole memory sectors (7 % faster lmfe on ragged corpora).
    """
    lengths = np.array([len(s) for s in signals], dtype=np.int64)
    padded = (lengths + align - 1) // align * align
    starts = np.zeros(len(signals), dtype=np.int64)
    if len(signals) > 1:
        np.cumsum(padded[:-1], out=starts[1:])
    arena = np.zeros(int(padded.sum()), dtype=dtype)
    for s, st, n in zip(signals, starts, lengths):
        arena[st:st + n] = s
    return arena, starts, lengths


def bucket_order(lengths, bucket_samples: int = 8000):
    """Indices that visit utterances bucket by bucket (0.5 s buckets at 16 kHz by
    default), longest bucket first; stable inside a bucket."""
    lengths = np.asarray(lengths, dtype=np.int64)
    buckets = lengths // int(bucket_samples)
    return np.argsort(-buckets, kind="stable")


def shard_indices(lengths, world_size: int, bucket_samples: int = 8000):
    """Deal utterances to ``world_size`` ranks so that total samples per rank are
    balanced: walk the length-bucketed order and always give the next utterance to
    the currently lightest rank.  Returns a list of index arrays (one per rank), each
    in bucketed order.  Deterministic; no communication needed to agree on it."""
    lengths = np.asarray(lengths, dtype=np.int64)
    order = bucket_order(lengths, bucket_samples)
    if world_size == 1:
        return [order]
    loads = np.zeros(world_size, dtype=np.int64)
    owner = np.empty(len(order), dtype=np.int64)
    # round-robin inside runs of equal bucket keeps this O(U) and within one
    # utterance of optimal per bucket; ties broken by rank id
    for pos, u in enumerate(order):
        r = int(np.argmin(loads))
        owner[pos] = r
        loads[r] += lengths[u]
    return [order[owner == r] for r in range(world_size)]


def shard_indices_fast(lengths, world_size: int, bucket_samples: int = 8000):
    """Vectorised variant of :func:`shard_indices` for very large corpora: bucketed
    order dealt in a serpentine (boustrophedon) pattern.  Imbalance is bounded by one
    utterance per bucket pass."""
    lengths = np.asarray(lengths, dtype=np.int64)
    order = np.argsort(-lengths, kind="stable")
    pos = np.arange(len(order))
    rnd, slot = np.divmod(pos, world_size)
    rank = np.where(rnd % 2 == 0, slot, world_size - 1 - slot)
    return [order[rank == r] for r in range(world_size)]


class FeatureBatch:
    """Device descriptors of a ragged batch for one feature configuration.

    Parameters mirror ``speechpy.feature.lmfe`` / ``mfcc`` (feature.py:102-112,222-223) and
    ``processing.cmvnw`` (processing.py:274).  ``preemphasis`` != 0: every utterance is run through
    ``processing.preemphasis(signal, shift=1, cof=preemphasis)`` (processing.py:45-58) first, fused into the
    framing kernel (the emphasised signal is never written to memory).
    """

    def __init__(self, starts, lengths, sampling_frequency=16000, frame_length=0.020,
                 frame_stride=0.01, num_filters=40, fft_length=512, low_frequency=0,
                 high_frequency=None, num_cepstral=13, dc_elimination=True, win_size=301,
                 variance_normalization=True, device=None, preemphasis=0.0):
        assert isinstance(win_size, int), "Size must be of type 'int'!"
        assert win_size % 2 == 1, "Windows size must be odd!"
        self.ctx = _lib.context(device)
        self.starts = np.ascontiguousarray(starts, dtype=np.int64)
        self.lengths = np.ascontiguousarray(lengths, dtype=np.int64)
        L, S = _processing._geometry(0, sampling_frequency, frame_length, frame_stride)
        high_frequency = high_frequency or sampling_frequency / 2
        K = int(fft_length) // 2 + 1
        bank = _feature.filterbanks(num_filters, K, sampling_frequency, low_frequency, high_frequency)
        self.num_filters = int(num_filters)
        self.num_cepstral = min(int(num_cepstral), int(num_filters))
        self.plan = self.ctx.plan(L, int(S), int(fft_length), self.num_filters, self.num_cepstral,
                                  1 if dc_elimination else 0, bank, int(win_size),
                                  1 if variance_normalization else 0, preemphasis=float(preemphasis))
        h = _lib.c_vp()
        _lib.check(self.ctx.lib.spb_batch_create(self.ctx.handle, self.plan, _lib.ptr(self.starts),
                                                 _lib.ptr(self.lengths), len(self.starts), C.byref(h)),
                   self.ctx.handle)
        self.handle = h
        tf = _lib.c_i64()
        _lib.check(self.ctx.lib.spb_batch_total_frames(h, C.byref(tf)))
        self.total_frames = tf.value
        self.frame_offsets = np.empty(len(self.starts) + 1, dtype=np.int64)
        _lib.check(self.ctx.lib.spb_batch_frame_offsets(h, _lib.ptr(self.frame_offsets)))
        self.min_arena = int((self.starts + self.lengths).max()) if len(self.starts) else 0

    def close(self):
        if self.handle:
            self.ctx.lib.spb_batch_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- host (numpy) entry points -------------------------------------------------
    def lmfe_cmvnw(self, arena: np.ndarray, out: np.ndarray | None = None,
                   sample_scale: float = 1.0) -> np.ndarray:
        """lmfe + cmvnw of every utterance: (total_frames, num_filters) float32; rows of
        utterance u are frame_offsets[u]:frame_offsets[u+1].

        ``arena`` may be float32, float64, int16 or int32; the device computes
        ``float(sample) * sample_scale``.  ``sample_scale=1`` is ``signal.astype(float)``
        (feature.py:182); an int16 PCM arena with ``sample_scale=PCM16_SCALE`` (1/32768) is the
        float32 audio ``librosa.load`` returns in the reference's ``load_wav`` (utils.py:170-173)."""
        arena, code = _as_arena(arena)
        if out is None:
            out = np.empty((self.total_frames, self.num_filters), dtype=np.float32)
        _lib.check(self.ctx.lib.spb_lmfe_cmvnw_batch_dtype(self.ctx.handle, self.handle, _lib.ptr(arena), code,
                                                           float(sample_scale), arena.shape[0], _lib.ptr(out)),
                   self.ctx.handle)
        return out

    def mfcc_delta(self, arena: np.ndarray, out: np.ndarray | None = None,
                   sample_scale: float = 1.0) -> np.ndarray:
        """mfcc + extract_derivative_feature: (total_frames, num_cepstral, 3) float32."""
        arena, code = _as_arena(arena)
        if out is None:
            out = np.empty((self.total_frames, self.num_cepstral, 3), dtype=np.float32)
        _lib.check(self.ctx.lib.spb_mfcc_delta_batch_dtype(self.ctx.handle, self.handle, _lib.ptr(arena), code,
                                                           float(sample_scale), arena.shape[0], _lib.ptr(out)),
                   self.ctx.handle)
        return out

    # ---- training cubes (utils.CMVN + utils.FeatureCube / FeatureCube3C, utils.py:325-397) ----
    def draw_crop_starts(self, num_crops: int = 20, crop_frames: int = 80, rng=None) -> np.ndarray:
        """(U, num_crops) int32 crop starts, drawn like ``FeatureCube.__call__`` does per
        utterance: ``np.random.randint(F - crop_frames, size=num_crops)`` (utils.py:373), in
        utterance order, from the global numpy RNG unless ``rng`` (a RandomState) is given."""
        rnd = np.random if rng is None else rng
        F = np.diff(self.frame_offsets)
        return np.stack([rnd.randint(int(f) - crop_frames, size=num_crops) for f in F]).astype(np.int32)

    def feature_cubes(self, arena: np.ndarray, crop_starts=None, num_crops: int = 20, crop_frames: int = 80,
                      derivative: bool = False, normalize: bool = False, sample_scale: float = 1.0,
                      out: np.ndarray | None = None) -> np.ndarray:
        """lmfe -> [Delta, Delta-Delta] -> [per-utterance cmvn] -> random crops, all on the
        device: (U, channels, num_crops, crop_frames, num_filters) float32, channels = 3 if
        ``derivative`` else 1; slice ``[u]`` is what the reference's transform pipeline
        ``Compose([CMVN(), FeatureCube(...)])`` (train.py:194-210) yields for utterance u."""
        arena, code = _as_arena(arena)
        if crop_starts is None:
            crop_starts = self.draw_crop_starts(num_crops, crop_frames)
        crops = np.ascontiguousarray(crop_starts, dtype=np.int32).reshape(len(self.starts), -1)
        num_crops = crops.shape[1]
        shape = (len(self.starts), 3 if derivative else 1, num_crops, crop_frames, self.num_filters)
        if out is None:
            out = np.empty(shape, dtype=np.float32)
        assert out.shape == shape and out.dtype == np.float32 and out.flags.c_contiguous
        _lib.check(self.ctx.lib.spb_feature_cubes_batch(
            self.ctx.handle, self.handle, _lib.ptr(arena), code, float(sample_scale), arena.shape[0],
            int(bool(derivative)), int(bool(normalize)), _lib.ptr(crops), num_crops, int(crop_frames),
            _lib.ptr(out)), self.ctx.handle)
        return out

    def feature_cubes_dev(self, d_samples: int, arena_len: int, d_out: int, crop_starts, crop_frames: int = 80,
                          derivative: bool = False, normalize: bool = False):
        """Device-pointer variant (float32 arena in, cubes out; ``crop_starts`` host int32 (U, crops))."""
        crops = np.ascontiguousarray(crop_starts, dtype=np.int32).reshape(len(self.starts), -1)
        _lib.check(self.ctx.lib.spb_feature_cubes_batch_dev(
            self.ctx.handle, self.handle, d_samples, arena_len, int(bool(derivative)), int(bool(normalize)),
            _lib.ptr(crops), crops.shape[1], int(crop_frames), d_out), self.ctx.handle)

    # ---- device-pointer entry points (asynchronous on ctx.stream()) -----------------
    def lmfe_cmvnw_dev(self, d_samples: int, arena_len: int, d_out: int):
        _lib.check(self.ctx.lib.spb_lmfe_cmvnw_batch_dev(self.ctx.handle, self.handle, d_samples,
                                                         arena_len, d_out), self.ctx.handle)

    def lmfe_dev(self, d_samples: int, arena_len: int, d_out: int):
        _lib.check(self.ctx.lib.spb_lmfe_batch_dev(self.ctx.handle, self.handle, d_samples, arena_len,
                                                   d_out), self.ctx.handle)

    def mfcc_delta_dev(self, d_samples: int, arena_len: int, d_out: int):
        _lib.check(self.ctx.lib.spb_mfcc_delta_batch_dev(self.ctx.handle, self.handle, d_samples,
                                                         arena_len, d_out), self.ctx.handle)

    def split(self, flat: np.ndarray):
        """Views of a batched result, one per utterance."""
        return [flat[self.frame_offsets[u]:self.frame_offsets[u + 1]] for u in range(len(self.starts))]


def lmfe_cmvnw(signals, **kw):
    """Convenience: list of 1-D signals -> list of (F_u, M) float32 arrays."""
    arena, starts, lengths = pack(signals)
    fb = FeatureBatch(starts, lengths, **kw)
    try:
        return [a.copy() for a in fb.split(fb.lmfe_cmvnw(arena))]
    finally:
        fb.close()


class CubeCollate:
    """``collate_fn`` for a DataLoader whose dataset yields raw ``(signal, label)`` pairs (or
    ``{'signal': ..., 'label': ...}`` dicts): ONE ragged batch call per mini-batch instead of the
    reference's per-utterance chain ``speech.lmfe`` (load_data.py:61-68) ->
    ``Compose([CMVN(), FeatureCube(cube_shape), ToTensor()])`` (train.py:194-210).

    Returns ``(features, labels)`` like the reference's loader does (train.py:76):
    ``features`` is (B, channels, num_crops, crop_frames, num_filters) float32 and ``labels``
    int64 (utils.ToTensor, utils.py:400-408).  With ``as_torch=True`` both are torch tensors
    and ``features`` is produced directly in device memory, so ``train_input.cuda()``
    (train.py:78) becomes a no-op; with ``as_torch=False`` they are numpy arrays.

    The defaults are the reference's training call: 25 ms frames, 10 ms stride, 40 filters,
    NFFT 1024 (load_data.py:61-68) and cube_shape (80, 40, 20) (train.py:199,204).  Crop starts
    come from the global numpy RNG in utterance order, as ``FeatureCube.__call__`` draws them
    (utils.py:373), unless ``rng`` is given.
    """

    def __init__(self, cube_shape=(80, 40, 20), derivative: bool = False, normalize: bool = False,
                 sampling_frequency=16000, frame_length=0.025, frame_stride=0.01, num_filters=None,
                 fft_length=1024, sample_scale: float = 1.0, as_torch: bool = True, rng=None, device=None):
        self.crop_frames, self.num_filters, self.num_crops = (int(v) for v in cube_shape[:3])
        if num_filters is not None:
            assert int(num_filters) == self.num_filters, "cube_shape[1] is the number of filters"
        self.derivative, self.normalize = bool(derivative), bool(normalize)
        self.kw = dict(sampling_frequency=sampling_frequency, frame_length=frame_length,
                       frame_stride=frame_stride, num_filters=self.num_filters, fft_length=fft_length,
                       device=device)
        self.sample_scale, self.as_torch, self.rng = float(sample_scale), bool(as_torch), rng

    @staticmethod
    def _unpack(sample):
        if isinstance(sample, dict):
            return sample["signal"], sample["label"]
        return sample[0], sample[1]

    def __call__(self, samples):
        signals, labels = zip(*(self._unpack(s) for s in samples))
        signals = [np.asarray(s).reshape(-1) for s in signals]
        pcm = all(s.dtype == np.int16 for s in signals)
        arena, starts, lengths = pack(signals, dtype=np.int16 if pcm else np.float32)
        fb = FeatureBatch(starts, lengths, **self.kw)
        try:
            if int(np.diff(fb.frame_offsets).min()) <= self.crop_frames:
                # np.random.randint(F - crop_frames) raises for F <= crop_frames (utils.py:373)
                raise ValueError("utterance shorter than crop_frames + 1 frames")
            crops = fb.draw_crop_starts(self.num_crops, self.crop_frames, self.rng)
            labels = np.asarray(labels, dtype=np.int64)
            if not self.as_torch:
                return fb.feature_cubes(arena, crops, crop_frames=self.crop_frames, derivative=self.derivative,
                                        normalize=self.normalize, sample_scale=self.sample_scale), labels
            import torch
            dev = torch.device("cuda", fb.ctx.device)
            d_in = torch.from_numpy(arena).to(dev)
            if d_in.dtype != torch.float32 or self.sample_scale != 1.0:
                d_in = d_in.to(torch.float32).mul_(self.sample_scale)  # exact for power-of-two scales
            out = torch.empty((len(signals), 3 if self.derivative else 1, self.num_crops, self.crop_frames,
                               self.num_filters), dtype=torch.float32, device=dev)
            torch.cuda.current_stream(dev).synchronize()
            fb.feature_cubes_dev(d_in.data_ptr(), d_in.numel(), out.data_ptr(), crops, self.crop_frames,
                                 self.derivative, self.normalize)
            fb.ctx.sync()
            return out, torch.from_numpy(labels)
        finally:
            fb.close()
