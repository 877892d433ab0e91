"""GPU parity against the committed golden vectors (tests/golden/stage12.npz; generated from torchaudio / transformers
and the CPU oracle by tests/golden/make_golden.py).  Nothing here reads /root/reference or needs torchaudio."""
from pathlib import Path

import numpy as np
import pytest
import torch

from fangyan_tts_b200 import synth_weights
from fangyan_tts_b200.configs import CamPlusConfig, S3Config

pytestmark = pytest.mark.gpu
GOLD = Path(__file__).resolve().parent / "golden" / "stage12.npz"


@pytest.fixture(scope="module")
def gold():
    with np.load(GOLD) as z:
        return {k: z[k] for k in z.files}


def _rel(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def test_frontends_match_golden(engine, gold):
    n = int(gold["n_utts"])
    pcm = [gold[f"pcm_{i}"] for i in range(n)]
    lens = [len(p) for p in pcm]
    dev = torch.from_numpy(np.concatenate(pcm)).cuda()
    feat, frames = engine.fbank_cmn(dev, lens)
    mel, mframes = engine.logmel(dev, lens)
    feat, mel = feat.cpu().numpy(), mel.cpu().numpy()
    fo = mo = 0
    for i in range(n):
        ref = gold[f"fbank_cmn_{i}"]
        assert frames[i] == ref.shape[0]
        if ref.shape[0] > 1:     # a single all-silent frame is CMN'ed to exact zeros on both sides
            assert _rel(feat[fo:fo + frames[i]], ref) <= 1e-4
        else:
            assert np.abs(feat[fo:fo + frames[i]] - ref).max() <= 1e-5
        fo += frames[i]
        ref = gold[f"logmel_{i}"]
        assert mframes[i] == ref.shape[1]
        got = mel[mo * 128:(mo + mframes[i]) * 128].reshape(128, mframes[i])
        assert _rel(got, ref) <= 1e-4
        mo += mframes[i]


def test_resample_matches_golden(engine, gold):
    st = gold["stereo24"].astype(np.float32).T / 32768.0          # [2, N]
    out, lens = engine.resample(torch.from_numpy(np.ascontiguousarray(st).reshape(-1)).cuda(), [st.shape[1]] * 2, 24000)
    ref = gold["resampled"]
    assert lens == [ref.shape[1]] * 2
    assert np.abs(out.cpu().numpy().reshape(2, -1) - ref).max() <= 1e-6


def test_campplus_matches_golden(engine, gold):
    engine.load_campplus(synth_weights.campplus_state(seed=1986), CamPlusConfig())
    ids = [i for i in range(int(gold["n_utts"])) if f"emb_{i}" in gold]
    feats = [gold[f"fbank_cmn_{i}"] for i in ids]
    emb = engine.campplus(torch.from_numpy(np.concatenate(feats)).cuda(), [f.shape[0] for f in feats]).cpu().numpy()
    for j, i in enumerate(ids):
        ref = gold[f"emb_{i}"]
        cos = float(np.dot(emb[j], ref) / (np.linalg.norm(emb[j]) * np.linalg.norm(ref)))
        assert cos >= 0.9999, (i, cos)


def test_s3tok_matches_golden(engine, gold):
    cfg = S3Config()
    engine.load_s3tok(synth_weights.s3tok_state(cfg, seed=1986), cfg)
    n = int(gold["n_utts"])
    mels = [gold[f"logmel_{i}"] for i in range(n)]
    packed = torch.from_numpy(np.concatenate([m.reshape(-1) for m in mels])).cuda()
    tok, counts, pre = engine.s3tok(packed, [m.shape[1] for m in mels], want_pre=True)
    tok, pre = tok.cpu().numpy(), pre.cpu().numpy()
    o = 0
    for i in range(n):
        ref, rpre = gold[f"tok_{i}"], gold[f"pre_{i}"]
        assert counts[i] == len(ref)                              # token sequence lengths bit-exact
        got = tok[o:o + counts[i]]
        assert np.abs(pre[o:o + counts[i]] - rpre).max() <= 1e-3
        for j in np.nonzero(got != ref)[0]:                        # every mismatch sits on an FSQ rounding boundary
            assert np.abs(np.abs(rpre[j]) - 0.5).min() < 1e-3
        o += counts[i]
