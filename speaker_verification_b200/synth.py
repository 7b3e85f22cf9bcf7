"""Deterministic synthetic corpora for parity tests and benchmarks.

No audio ships with the reference, so every
measurement uses synthetic 16 kHz mono audio: i.i.d. N(0, 0.1) samples as in the
reference's own test (speech_feature_extraction/tests/test_speechpy.py:13-14),
drawn from ``numpy.random.Generator(Philox(seed))`` so any host regenerates the
same stream.  Shapes follow BASELINE.json's configs.
"""
from __future__ import annotations

import numpy as np

SAMPLE_RATE = 16000


def audio(seed: int, num_samples: int, dtype=np.float32, sigma: float = 0.1) -> np.ndarray:
    """One utterance of N(0, sigma) noise."""
    rng = np.random.Generator(np.random.Philox(seed))
    return (sigma * rng.standard_normal(num_samples)).astype(dtype)


def pcm16(seed: int, num_samples: int, sigma: float = 0.1) -> np.ndarray:
    """One utterance of 16-bit PCM noise, round(N(0, sigma) * 32768) clipped to int16 (what a
    VoxCeleb wav holds; ``librosa.load`` in the reference's ``load_wav``, utils.py:170-173,
    decodes it to ``pcm / 32768`` float32)."""
    x = audio(seed, num_samples, np.float64, sigma) * 32768.0
    return np.clip(np.round(x), -32768, 32767).astype(np.int16)


def cfg3_lengths(num_utterances: int, seed: int = 3) -> np.ndarray:
    """VoxCeleb1-dev-shaped durations: clip(lognormal(ln 7.2 s, 0.5), 4 s, 20 s)
    in whole samples (int64)."""
    rng = np.random.Generator(np.random.Philox(seed))
    dur = np.clip(rng.lognormal(np.log(7.2), 0.5, size=num_utterances), 4.0, 20.0)
    return np.round(dur * SAMPLE_RATE).astype(np.int64)


def ragged_corpus(lengths: np.ndarray, seed: int, align: int = 8):
    """Concatenated fp32 arena + int64 offsets[U+1] for a ragged batch.

    Every utterance starts on a multiple of ``align`` samples (16 bytes for
    fp32) -- the layout the batch kernels' 128-bit loads prefer; the gap samples
    are zero and belong to no utterance.
    """
    lengths = np.asarray(lengths, dtype=np.int64)
    starts = np.zeros(len(lengths) + 1, dtype=np.int64)
    padded = (lengths + align - 1) // align * align
    np.cumsum(padded, out=starts[1:])
    rng = np.random.Generator(np.random.Philox(seed))
    arena = (0.1 * rng.standard_normal(int(starts[-1]), dtype=np.float32))
    for u in range(len(lengths)):
        arena[starts[u] + lengths[u]:starts[u + 1]] = 0.0
    return arena, starts[:-1].copy(), lengths
