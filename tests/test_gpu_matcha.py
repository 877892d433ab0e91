"""Row N4 on the GPU: fy_matcha_mel against the reference-generated golden vectors and the oracle, packed ragged
batches, int16 input, batch invariance, and the CosyVoiceFrontEnd twin's _extract_speech_feat contract."""
import wave

import numpy as np
import pytest
import torch

from fangyan_tts_b200 import synth
from oracle import frontend as F
from test_matcha_cpu import GOLD, golden_inputs, rel_err

pytestmark = pytest.mark.gpu


def test_matches_reference_golden(engine):
    gold = np.load(GOLD / "matcha.npz")
    xs = golden_inputs()
    pcm = torch.from_numpy(np.concatenate(xs)).cuda()
    mel, frames = engine.matcha_mel(pcm, [len(x) for x in xs])
    mel = mel.cpu().numpy()
    o = 0
    for i, t in enumerate(frames):
        want = gold[f"mel{i}"].T                       # [T, 80], the layout both reference callers use
        assert want.shape == (t, 80)
        assert rel_err(mel[o:o + t], want) <= 1e-4
        o += t


def test_ragged_batch_int16_and_invariance(engine):
    rng = np.random.default_rng(77)
    utts = [synth.utterance(rng, d)[:n] for d, n in ((0.1, 721), (0.2, 2399), (1.0, 24000), (3.3, 79201), (0.5, 12001))]
    lens = [len(u) for u in utts]
    pcm = torch.from_numpy(np.concatenate(utts)).cuda()                     # int16 on the wire
    mel, frames = engine.matcha_mel(pcm, lens)
    mel = mel.cpu().numpy()
    o = 0
    for u, t in zip(utts, frames):
        want = F.matcha_mel(torch.from_numpy(u.astype(np.float32) / 32768.0)[None])[0].numpy().T
        assert want.shape == (t, 80)
        assert rel_err(mel[o:o + t], want) <= 1e-4
        alone, _ = engine.matcha_mel(torch.from_numpy(u).cuda(), [len(u)])
        assert np.array_equal(alone.cpu().numpy(), mel[o:o + t])           # bit-identical whatever the batch
        o += t
    with pytest.raises(ValueError):
        engine.matcha_mel(pcm[:720], [720])


def test_frontend_twin_speech_feat(engine, tmp_path):
    from fangyan_tts_b200.frontend import CosyVoiceFrontEndB200
    fe = CosyVoiceFrontEndB200.__new__(CosyVoiceFrontEndB200)
    fe.engine, fe.device = engine, engine.device
    rng = np.random.default_rng(5)
    x = synth.utterance(rng, 2.0)
    path = tmp_path / "p.wav"
    with wave.open(str(path), "wb") as w:
        w.setnchannels(1); w.setsampwidth(2); w.setframerate(16000); w.writeframes(x.tobytes())
    feat, n = fe._extract_speech_feat(str(path))
    wav24 = F.resample(torch.from_numpy(x.astype(np.float32) / 32768.0)[None], 16000, 24000)
    want = F.matcha_mel(wav24)[0].T
    assert feat.dtype == torch.float32 and feat.shape == (1, want.shape[0], 80) and feat.is_cuda
    assert n.dtype == torch.int32 and int(n[0]) == want.shape[0]
    assert rel_err(feat[0].cpu().numpy(), want.numpy()) <= 1e-4
