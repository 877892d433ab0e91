"""CPU tests that pin the oracle (oracle/) before anything is compared against it.

The reference has no tests or golden vectors for this path, so the pins are (a) the third-party callees
the reference uses, where they import here, and (b) tests/golden/stage12.npz, which was generated from those callees
(tests/golden/make_golden.py).
"""
import json
from pathlib import Path

import numpy as np
import pytest
import torch

from fangyan_tts_b200 import synth, synth_weights
from oracle import campplus as CP
from oracle import frontend as F
from oracle import s3tokenizer as S3

GOLD = Path(__file__).resolve().parent / "golden" / "stage12.npz"


@pytest.fixture(scope="module")
def gold():
    with np.load(GOLD) as z:
        return {k: z[k] for k in z.files}


def _wave(pcm):
    return torch.from_numpy(pcm.astype(np.float32) / 32768.0)[None]


def _rel(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def test_golden_meta(gold):
    meta = json.loads(str(gold["meta"]))
    assert meta["seed"] == 1986 and "torchaudio" in meta


def test_fbank_cmn_matches_golden(gold):
    for i in range(int(gold["n_utts"])):
        got = F.cmn(F.kaldi_fbank(_wave(gold[f"pcm_{i}"]))).numpy()
        ref = gold[f"fbank_cmn_{i}"]
        assert got.shape == ref.shape == (F.kaldi_num_frames(len(gold[f"pcm_{i}"])), 80)
        assert np.abs(got - ref).max() <= 1e-5


def test_logmel_matches_golden(gold):
    for i in range(int(gold["n_utts"])):
        got = F.whisper_log_mel(_wave(gold[f"pcm_{i}"]))[0].numpy()
        ref = gold[f"logmel_{i}"]
        assert got.shape == ref.shape == (128, len(gold[f"pcm_{i}"]) // 160)
        assert np.abs(got - ref).max() <= 5e-5      # float32 torch.stft vs float64 numpy spectrogram


def test_resample_matches_golden(gold):
    w = torch.from_numpy(gold["stereo24"].astype(np.float32).T / 32768.0)
    got = F.resample(w, 24000).numpy()
    assert got.shape == gold["resampled"].shape
    assert np.abs(got - gold["resampled"]).max() <= 1e-6


def test_fbank_bit_exact_vs_torchaudio():
    kaldi = pytest.importorskip("torchaudio.compliance.kaldi")
    for u in synth.make_set(4, 11, 0.3, 3.0):
        w = _wave(u)
        ref = kaldi.fbank(w, num_mel_bins=80, dither=0, sample_frequency=16000)
        assert torch.equal(F.kaldi_fbank(w), ref)


@pytest.mark.parametrize("sr", [8000, 22050, 24000, 44100, 48000])
def test_resample_vs_torchaudio(sr):
    ta = pytest.importorskip("torchaudio")
    w = torch.from_numpy(np.random.default_rng(sr).standard_normal((2, sr // 3 + 7)).astype(np.float32) * 0.1)
    ref = ta.transforms.Resample(sr, 16000)(w)
    got = F.resample(w, sr)
    assert got.shape == ref.shape and got.shape[1] == int(np.ceil(16000 * w.shape[1] / sr))
    assert float((got - ref).abs().max()) <= 1e-6


def test_logmel_vs_transformers():
    tr = pytest.importorskip("transformers")
    fe = tr.WhisperFeatureExtractor(feature_size=128)
    for u in synth.make_set(3, 12, 0.2, 4.0):
        w = _wave(u)
        ref = np.asarray(fe._np_extract_fbank_features(w.numpy().astype(np.float64), "cpu"), dtype=np.float32)[0]
        got = F.whisper_log_mel(w)[0].numpy()
        assert got.shape == ref.shape
        assert np.abs(got - ref).max() <= 5e-5      # fp32 torch.stft vs float64 numpy, log of small powers


def test_slaney_filters_vs_transformers():
    tr = pytest.importorskip("transformers")
    ref = tr.audio_utils.mel_filter_bank(201, 128, 0.0, 8000.0, 16000, norm="slaney", mel_scale="slaney").T
    assert np.abs(F.slaney_mel_filters() - ref).max() <= 1e-7


def test_frame_and_token_counts():
    # 3 s -> M 298, T 300, L 75; 10 s -> 998, 1000, 250; 30 s -> 2998, 3000, 750
    for secs, m, t, l in ((3, 298, 300, 75), (10, 998, 1000, 250), (30, 2998, 3000, 750)):
        n = secs * 16000
        assert F.kaldi_num_frames(n) == m and F.whisper_num_frames(n) == t and S3.token_len(t) == l
    assert F.kaldi_num_frames(399) == 0 and F.kaldi_num_frames(400) == 1
    assert S3.token_len(1) == 1 and S3.token_len(2) == 1 and S3.token_len(3) == 1 and S3.token_len(5) == 2


def test_stage2_gate_is_strict():
    assert F.stage2_gate(30 * 16000)            # exactly 30.0 s is tokenised (extract_speech_token.py:34)
    assert not F.stage2_gate(30 * 16000 + 1)


def _cam():
    m = CP.CamPlus()
    m.load_state_dict({k: torch.from_numpy(v) for k, v in synth_weights.campplus_state(seed=1986).items()}, strict=False)
    return m.eval()


def test_campplus_matches_golden(gold):
    m = _cam()
    for i in range(int(gold["n_utts"])):
        if f"emb_{i}" not in gold:
            continue
        got = CP.embed_one(m, torch.from_numpy(gold[f"fbank_cmn_{i}"])).numpy()
        assert got.shape == (192,)
        assert _rel(got, gold[f"emb_{i}"]) <= 1e-4


def test_campplus_batch1_semantics():
    """Zero-padding the input must change the result: the network has no internal masking, which is why the batched
    kernels re-mask after every layer."""
    m = _cam()
    feat = torch.randn(120, 80)
    a = CP.embed_one(m, feat)
    b = CP.embed_one(m, torch.cat([feat, torch.zeros(40, 80)]))
    assert float((a - b).abs().max()) > 1e-3


def test_s3tok_matches_golden(gold):
    m = S3.S3Tokenizer(S3.S3Config())
    m.load_state_dict({k: torch.from_numpy(v) for k, v in synth_weights.s3tok_state(seed=1986).items()})
    m.eval()
    for i in range(int(gold["n_utts"])):
        mel = torch.from_numpy(gold[f"logmel_{i}"])[None]
        ids = S3.tokenize_one(m, mel)
        ref, pre = gold[f"tok_{i}"], gold[f"pre_{i}"]
        assert len(ids) == len(ref) == S3.token_len(mel.shape[2])
        assert all(0 <= t < 6561 for t in ids)
        bad = [j for j, (a, b) in enumerate(zip(ids, ref.tolist())) if a != b]
        for j in bad:   # a different BLAS may flip an id only on a rounding boundary
            assert np.abs(np.abs(pre[j]) - 0.5).min() < 1e-3


def test_s3tok_padding_invariance():
    """The batched form with a length mask must equal the batch-1 run the reference does."""
    cfg = S3.S3Config(n_layer=2)
    m = S3.init_random(S3.S3Tokenizer(cfg))
    a, b = torch.randn(1, 128, 90), torch.randn(1, 128, 57)
    batch = torch.zeros(2, 128, 90)
    batch[0], batch[1, :, :57] = a[0], b[0]
    ids, lens = m.quantize(batch, torch.tensor([90, 57]))
    assert ids[0, :lens[0]].tolist() == S3.tokenize_one(m, a)
    assert ids[1, :lens[1]].tolist() == S3.tokenize_one(m, b)
