"""GPU parity: CAMPPlus through the C ABI vs the CPU oracle (oracle/campplus.py), one utterance at a time."""
import numpy as np
import pytest
import torch

from fangyan_tts_b200 import synth
from oracle import campplus as CP
from oracle import frontend as F

pytestmark = pytest.mark.gpu
COS_TOL = 0.9999  # north_star: embeddings at cosine >= 0.9999


def test_campplus_matches_oracle(engine):
    cfg = CP.CamPlusConfig()
    model = CP.init_random(CP.CamPlus(cfg), seed=1986)
    engine.load_campplus(dict(model.state_dict()), cfg)
    utts = synth.make_set(7, 21, 0.6, 7.0) + [synth.make_set(1, 5, 0.065, 0.07)[0], synth.make_set(1, 6, 20.0, 21.0)[0]]
    feats = [F.cmn(F.kaldi_fbank(torch.from_numpy(u.astype(np.float32) / 32768.0)[None])) for u in utts]
    frames = [f.shape[0] for f in feats]
    emb = engine.campplus(torch.cat(feats).cuda(), frames).cpu()
    worst = 1.0
    for i, f in enumerate(feats):
        ref = CP.embed_one(model, f)
        cos = float(torch.nn.functional.cosine_similarity(emb[i], ref, dim=0))
        rel = float((emb[i] - ref).abs().max() / ref.abs().max())
        print(f"utt {i}: frames={frames[i]} cos={cos:.7f} rel={rel:.3e}")
        worst = min(worst, cos)
    assert worst >= COS_TOL, worst


def test_stage1_end_to_end(engine):
    """PCM -> fbank+CMN -> CAMPPlus on the device vs the oracle chain, plus the speaker mean."""
    cfg = CP.CamPlusConfig()
    model = CP.init_random(CP.CamPlus(cfg), seed=7)
    engine.load_campplus(dict(model.state_dict()), cfg)
    utts = synth.make_set(6, 33, 1.0, 4.0)
    pcm = torch.from_numpy(np.concatenate(utts)).cuda()
    feat, frames = engine.fbank_cmn(pcm, [len(u) for u in utts])
    emb = engine.campplus(feat, frames).cpu()
    for i, u in enumerate(utts):
        ref = CP.embed_one(model, F.cmn(F.kaldi_fbank(torch.from_numpy(u.astype(np.float32) / 32768.0)[None])))
        assert float(torch.nn.functional.cosine_similarity(emb[i], ref, dim=0)) >= COS_TOL
