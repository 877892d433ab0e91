"""Seeded synthetic utterances: harmonic tone complexes with a slow envelope, -30 dB noise,
near-silent edges, peak 0.5, quantised to int16.  Used by tests and bench.py (there are no datasets offline)."""
from __future__ import annotations

import wave
from pathlib import Path

import numpy as np


def utterance(rng: np.random.Generator, seconds: float, sr: int = 16000) -> np.ndarray:
    n = int(round(seconds * sr))
    t = np.arange(n, dtype=np.float64) / sr
    f0 = rng.uniform(90.0, 260.0)
    x = np.zeros(n)
    for h in range(1, int(rng.integers(3, 7)) + 1):
        x += rng.uniform(0.2, 1.0) / h * np.sin(2 * np.pi * f0 * h * t + rng.uniform(0, 2 * np.pi))
    env = 0.6 + 0.4 * np.sin(2 * np.pi * rng.uniform(0.3, 2.0) * t + rng.uniform(0, 2 * np.pi))
    x *= env
    x += rng.standard_normal(n) * (10 ** (-30 / 20)) * np.abs(x).max()
    edge = min(int(rng.uniform(0.1, 0.3) * sr), n // 4)
    if edge > 0:
        x[:edge] *= 1e-5
        x[n - edge:] *= 1e-5
    x *= 0.5 / max(np.abs(x).max(), 1e-9)
    return np.clip(np.round(x * 32768.0), -32768, 32767).astype(np.int16)


def make_set(n_utts: int, seed: int, lo: float = 3.0, hi: float = 10.0, sr: int = 16000):
    rng = np.random.default_rng(seed)
    return [utterance(rng, float(rng.uniform(lo, hi)), sr) for _ in range(n_utts)]


def write_wav(path, pcm: np.ndarray, sr: int = 16000):
    """pcm int16 [N] or [N, C]."""
    pcm = np.asarray(pcm, dtype=np.int16)
    ch = 1 if pcm.ndim == 1 else pcm.shape[1]
    with wave.open(str(path), "wb") as w:
        w.setnchannels(ch)
        w.setsampwidth(2)
        w.setframerate(sr)
        w.writeframes(pcm.tobytes())


def write_kaldi_dir(root, utts, sr: int = 16000, n_spk: int = 11, prefix: str = "utt"):
    """Write wav/ + wav.scp + utt2spk + text the way prepare_training_data.py lays a data dir out."""
    root = Path(root)
    (root / "wav").mkdir(parents=True, exist_ok=True)
    keys = []
    with open(root / "wav.scp", "w") as scp, open(root / "utt2spk", "w") as u2s, open(root / "text", "w") as txt:
        for i, pcm in enumerate(utts):
            key = f"{prefix}_{i:06d}"
            p = root / "wav" / f"{key}.wav"
            write_wav(p, pcm, sr)
            scp.write(f"{key} {p}\n")
            u2s.write(f"{key} spk{i % n_spk:02d}\n")
            txt.write(f"{key} synthetic {i}\n")
            keys.append(key)
    return keys
