"""Row N4 (24 kHz matcha mel), CPU side: the oracle restatement against golden vectors produced by the reference's
own mel_spectrogram (tests/golden/make_golden_matcha.py), its mel basis against transformers, frame-count rule."""
import importlib.util
from pathlib import Path

import numpy as np
import pytest
import torch

from fangyan_tts_b200.engine import matcha_num_frames
from oracle import frontend as F

GOLD = Path(__file__).resolve().parent / "golden"


def golden_inputs():
    spec = importlib.util.spec_from_file_location("make_golden_matcha", GOLD / "make_golden_matcha.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.inputs()


def rel_err(a, b):
    return float(np.abs(a - b).max() / np.abs(b).max())


def test_oracle_matches_reference_golden():
    gold = np.load(GOLD / "matcha.npz")
    for i, x in enumerate(golden_inputs()):
        assert len(x) == int(gold[f"n{i}"])
        got = F.matcha_mel(torch.from_numpy(x)[None])[0].numpy()
        assert got.shape == gold[f"mel{i}"].shape == (80, matcha_num_frames(len(x)))
        assert rel_err(got, gold[f"mel{i}"]) <= 1e-4          # north_star tolerance for mel features


def test_mel_basis_matches_transformers():
    au = pytest.importorskip("transformers.audio_utils")
    fb = au.mel_filter_bank(num_frequency_bins=961, num_mel_filters=80, min_frequency=0.0, max_frequency=12000.0,
                            sampling_rate=24000, norm="slaney", mel_scale="slaney").T
    mine = F.slaney_mel_filters(80, 1920, 24000)
    assert mine.shape == fb.shape == (80, 961)
    assert np.abs(mine - fb).max() <= 1e-6 * np.abs(fb).max()


def test_frame_count_rule():
    for n in (721, 959, 960, 12000, 48007, 151200):
        y = F.matcha_mel(torch.zeros(1, n))
        assert y.shape[-1] == matcha_num_frames(n)
    with pytest.raises(RuntimeError):
        F.matcha_mel(torch.zeros(1, 720))      # reflect padding needs more than 720 samples, as in the reference
