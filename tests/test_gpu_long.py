"""BASELINE configs[3]: long-utterance stress -- clips up to the 30 s cap, a 256-utterance packed batch (2 encoder
layers here: the full-depth every-utterance comparisons live in test_gpu_parity_config1.py and bench.py --long-stress),
parity against the oracle on a sample with the boundary-distance check, and the size-independent checks on all of it:
token counts bit-exact, ids inside the FSQ vocabulary, embeddings finite."""
import numpy as np
import pytest
import torch

from fangyan_tts_b200 import synth, synth_weights
from fangyan_tts_b200.configs import CamPlusConfig, S3Config
from fangyan_tts_b200.engine import token_len
from oracle import campplus as CP
from oracle import frontend as F
from oracle import s3tokenizer as S3

pytestmark = pytest.mark.gpu


def test_long_utterance_batch_256(engine):
    cfg = S3Config(n_layer=2)
    engine.load_campplus(synth_weights.campplus_state(seed=3), CamPlusConfig())
    engine.load_s3tok(synth_weights.s3tok_state(cfg, seed=4), cfg)
    rng = np.random.default_rng(1986 + 3)
    pool = [synth.utterance(rng, 30.0) for _ in range(6)]
    utts = []
    for i in range(256):
        n = 480000 if i in (0, 255) else int(rng.uniform(1.0, 30.0) * 16000)     # two clips at exactly 30.0 s
        utts.append(np.ascontiguousarray(pool[i % 6][:n]))
    pcm = torch.from_numpy(np.concatenate(utts)).cuda()
    lens = [len(u) for u in utts]
    feat, frames = engine.fbank_cmn(pcm, lens)
    emb = engine.campplus(feat, frames).cpu()
    mel, mframes = engine.logmel(pcm, lens)
    tok, counts = engine.s3tok(mel, mframes)
    tok = tok.cpu()
    assert counts == [token_len(n // 160) for n in lens] and counts[0] == 750
    assert int(tok.min()) >= 0 and int(tok.max()) < 6561
    assert torch.isfinite(emb).all()
    cam = CP.CamPlus()
    cam.load_state_dict({k: torch.from_numpy(v) for k, v in synth_weights.campplus_state(seed=3).items()}, strict=False)
    cam.eval()
    s3 = S3.S3Tokenizer(S3.S3Config(n_layer=2))
    s3.load_state_dict({k: torch.from_numpy(v) for k, v in synth_weights.s3tok_state(cfg, seed=4).items()})
    s3.eval()
    off = np.concatenate([[0], np.cumsum(counts)])
    total = bad = 0
    for i in (0, 17, 64, 128, 200, 255):
        w = torch.from_numpy(utts[i].astype(np.float32) / 32768.0)[None]
        ref = CP.embed_one(cam, F.cmn(F.kaldi_fbank(w)))
        assert float(torch.nn.functional.cosine_similarity(emb[i], ref, dim=0)) >= 0.9999
        mel = F.whisper_log_mel(w)
        with torch.no_grad():
            hid, _ = s3.encode(mel, torch.tensor([mel.shape[2]]))
            pre = s3.fsq_pre(hid)[0]
        ids = S3.tokenize_one(s3, mel)
        got = tok[off[i]:off[i + 1]].tolist()
        total += len(ids)
        for j, (a, b) in enumerate(zip(got, ids)):
            if a != b:
                bad += 1
                assert float((pre[j].abs() - 0.5).abs().min()) < 1e-3, (i, j)    # only on an FSQ rounding boundary
    assert total >= 2000 and bad / total <= 1e-3, (bad, total)


def test_full_size_workload_properties(engine):
    """BASELINE configs[1] at full size (the bench workload: 1,000 utterances of 3-10 s, all 12 encoder layers, CTA-pair
    GEMMs with fused split segments), checked through size-independent properties: token counts bit-exact against the
    length rule, ids inside the vocabulary, finite embeddings, bit-identical repeat, and -- for a sample -- the same
    bits when the utterance is processed alone (the reference's batch-1 semantics) plus oracle parity."""
    import sys
    from pathlib import Path
    sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
    import bench
    from fangyan_tts_b200.pipeline import Extractor, Utt
    cfg = S3Config()
    cam_w, s3_w = synth_weights.campplus_state(), synth_weights.s3tok_state(cfg)
    engine.load_campplus(cam_w, CamPlusConfig())
    engine.load_s3tok(s3_w, cfg)
    utts = bench.workload(1000, 1986)
    ex = Extractor(engine, max_batch_seconds=2400.0)
    emb, toks = ex.extract([Utt(u[:, None]) for u in utts])
    assert emb.shape == (1000, 192) and np.isfinite(emb).all()
    assert [len(t) for t in toks] == [token_len(len(u) // 160) for u in utts]
    assert all(0 <= min(t) and max(t) < 6561 for t in toks)
    emb2, toks2 = ex.extract([Utt(u[:, None]) for u in utts])
    assert np.array_equal(emb, emb2) and toks == toks2
    cam = CP.CamPlus()
    cam.load_state_dict({k: torch.from_numpy(v) for k, v in cam_w.items()}, strict=False)
    cam.eval()
    s3 = S3.S3Tokenizer(S3.S3Config())
    s3.load_state_dict({k: torch.from_numpy(v) for k, v in s3_w.items()})
    s3.eval()
    total = bad = 0
    for i in (0, 333, 999):
        e1, t1 = ex.extract([Utt(utts[i][:, None])])
        assert np.array_equal(e1[0], emb[i]) and t1[0] == toks[i]
        w = torch.from_numpy(utts[i].astype(np.float32) / 32768.0)[None]
        ref = CP.embed_one(cam, F.cmn(F.kaldi_fbank(w)))
        assert float(torch.nn.functional.cosine_similarity(torch.from_numpy(emb[i]), ref, dim=0)) >= 0.9999
        ids = S3.tokenize_one(s3, F.whisper_log_mel(w))
        total += len(ids)
        bad += sum(a != b for a, b in zip(toks[i], ids))
    assert bad / total <= 1e-3, (bad, total)
