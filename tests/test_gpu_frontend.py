"""GPU parity: front-end kernels through the C ABI vs the CPU oracle (oracle/frontend.py)."""
import numpy as np
import pytest
import torch

from fangyan_tts_b200 import synth
from oracle import frontend as F

pytestmark = pytest.mark.gpu
REL_TOL = 1e-4  # north_star: fbank and log-mel within 1e-4 relative (max|a-b| / max|b|)


def _pack(utts):
    return torch.from_numpy(np.concatenate(utts)), [len(u) for u in utts]


def _rel(a, b):
    if a.numel() == 0:
        return 0.0
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("dtype", ["s16", "f32"])
def test_fbank_cmn_matches_oracle(engine, dtype):
    utts = synth.make_set(9, 1986, 0.5, 6.0) + [np.zeros(400, np.int16), synth.make_set(1, 3, 29.9, 30.0)[0]]
    pcm, lens = _pack(utts)
    x = pcm if dtype == "s16" else pcm.float() / 32768.0
    feat, frames = engine.fbank_cmn(x.cuda(), lens)
    feat = feat.cpu()
    off = 0
    for u, m in zip(utts, frames):
        ref = F.cmn(F.kaldi_fbank(torch.from_numpy(u.astype(np.float32) / 32768.0)[None]))
        assert ref.shape[0] == m == F.kaldi_num_frames(len(u))
        got = feat[off:off + m]
        assert _rel(got, ref) <= REL_TOL, (len(u), _rel(got, ref))
        off += m
    assert off == feat.shape[0]


def test_logmel_matches_oracle(engine):
    utts = synth.make_set(9, 1987, 0.5, 6.0) + [synth.make_set(1, 4, 29.9, 30.0)[0], synth.make_set(1, 5, 0.05, 0.06)[0]]
    pcm, lens = _pack(utts)
    mel, frames = engine.logmel(pcm.cuda(), lens)
    mel = mel.cpu()
    off = 0
    for u, t in zip(utts, frames):
        ref = F.whisper_log_mel(torch.from_numpy(u.astype(np.float32) / 32768.0)[None])[0]
        assert ref.shape == (128, t)
        got = mel[off * 128:(off + t) * 128].view(128, t)
        assert _rel(got, ref) <= REL_TOL, (len(u), _rel(got, ref))
        off += t


@pytest.mark.parametrize("sr", [24000, 44100, 48000, 8000])
def test_resample_matches_oracle(engine, sr):
    rng = np.random.default_rng(sr)
    lens = [int(sr * s) + k for s, k in ((0.7, 3), (1.3, 0), (0.05, 1))]
    waves = [torch.from_numpy(rng.standard_normal(n).astype(np.float32) * 0.1) for n in lens]
    out, out_len = engine.resample(torch.cat(waves).cuda(), lens, sr)
    out = out.cpu()
    off = 0
    for w, n in zip(waves, out_len):
        ref = F.resample(w[None], sr)[0]
        assert ref.numel() == n
        assert _rel(out[off:off + n], ref) <= 1e-5
        off += n


def test_gemm_bf16_matches_torch(engine):
    torch.manual_seed(0)
    for (m, n, k) in [(128, 256, 64), (300, 1280, 1280), (1000, 128, 160), (257, 32, 384), (4096, 5120, 1280)]:
        a = torch.randn(m, k, device="cuda").bfloat16()
        b = torch.randn(n, k, device="cuda").bfloat16()
        got = engine.gemm_bf16(a, b)
        ref = a.float() @ b.float().T
        err = float((got - ref).abs().max() / ref.abs().max())
        assert err < 2e-5, ((m, n, k), err)


def test_speaker_mean(engine):
    torch.manual_seed(1)
    emb = torch.randn(37, 192, device="cuda")
    spk = torch.arange(37, device="cuda", dtype=torch.int32) % 5
    ssum = torch.zeros(5, 192, device="cuda")
    cnt = torch.zeros(5, dtype=torch.int32, device="cuda")
    engine.speaker_sum(emb[:20].contiguous(), spk[:20].contiguous(), 5, ssum, cnt)
    engine.speaker_sum(emb[20:].contiguous(), spk[20:].contiguous(), 5, ssum, cnt)
    mean = engine.speaker_mean(ssum, cnt)
    for s in range(5):
        ref = emb[spk == s].mean(0)
        assert torch.allclose(mean[s], ref, atol=1e-6)
