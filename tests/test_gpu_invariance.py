"""GPU property tests that hold at any size: every per-utterance result is bit-identical whatever else is packed
into the batch (the reference runs each utterance alone, so batching must not be observable), and repeatable."""
import numpy as np
import pytest
import torch

from fangyan_tts_b200 import synth, synth_weights
from fangyan_tts_b200.configs import CamPlusConfig, S3Config

pytestmark = pytest.mark.gpu


def _run(eng, utts, idx):
    pcm = torch.from_numpy(np.concatenate([utts[i] for i in idx])).cuda()
    lens = [len(utts[i]) for i in idx]
    feat, frames = eng.fbank_cmn(pcm, lens)
    emb = eng.campplus(feat, frames)
    mel, mframes = eng.logmel(pcm, lens)
    tok, counts, pre = eng.s3tok(mel, mframes, want_pre=True)
    torch.cuda.synchronize()
    return feat.cpu(), frames, mel.cpu(), mframes, emb.cpu(), tok.cpu(), pre.cpu(), counts


def test_results_do_not_depend_on_batch_composition(engine):
    engine.load_campplus(synth_weights.campplus_state(seed=11), CamPlusConfig())
    cfg = S3Config(n_layer=2)
    engine.load_s3tok(synth_weights.s3tok_state(cfg, seed=12), cfg)
    utts = synth.make_set(12, 1986, 0.3, 6.0) + [synth.make_set(1, 2, 29.0, 29.5)[0]]
    full = _run(engine, utts, list(range(len(utts))))
    again = _run(engine, utts, list(range(len(utts))))
    for a, b in zip(full, again):
        if torch.is_tensor(a):
            assert torch.equal(a, b)                                  # repeatable
    fo = np.concatenate([[0], np.cumsum(full[1])])
    mo = np.concatenate([[0], np.cumsum(full[3])])
    to = np.concatenate([[0], np.cumsum(full[7])])
    for sub in ([3], [12, 0, 7], [5, 4, 3, 2, 1], [11, 12]):
        part = _run(engine, utts, sub)
        f = m = t = 0
        for j, i in enumerate(sub):
            nf, nm, nt = part[1][j], part[3][j], part[7][j]
            assert torch.equal(part[0][f:f + nf], full[0][fo[i]:fo[i + 1]])                    # fbank + CMN
            assert torch.equal(part[2][m * 128:(m + nm) * 128], full[2][mo[i] * 128:mo[i + 1] * 128])   # log-mel
            assert torch.equal(part[4][j], full[4][i])                                           # embedding
            assert torch.equal(part[5][t:t + nt], full[5][to[i]:to[i + 1]])                      # token ids
            assert torch.equal(part[6][t:t + nt], full[6][to[i]:to[i + 1]])                      # pre-FSQ values
            f, m, t = f + nf, m + nm, t + nt


def test_extract_equals_single_stage_paths(engine):
    """Extractor.extract (one upload per batch, ramp-up batch, asynchronous result read-back over three streams) must
    return exactly what embed() + tokenize() return, in the caller's order, across repeated calls."""
    from fangyan_tts_b200 import synth_weights
    from fangyan_tts_b200.configs import CamPlusConfig, S3Config
    from fangyan_tts_b200.pipeline import Extractor, Utt
    cfg = S3Config(n_layer=2)
    engine.load_campplus(synth_weights.campplus_state(seed=21), CamPlusConfig())
    engine.load_s3tok(synth_weights.s3tok_state(cfg, seed=22), cfg)
    rng = np.random.default_rng(99)
    utts = [Utt(synth.utterance(rng, float(rng.uniform(0.5, 6.0)))[:, None]) for _ in range(150)]
    ex = Extractor(engine, max_batch_seconds=120.0, max_batch_utts=128)        # several batches + the ramp-up split
    emb_ref, tok_ref = ex.embed(utts), ex.tokenize(utts)
    for _ in range(2):
        emb, toks = ex.extract(utts)
        assert np.array_equal(emb, emb_ref)
        assert toks == tok_ref


def test_results_do_not_depend_on_workspace_contents(engine):
    """The caller-owned workspace is never cleared wholesale: every byte a kernel reads must have been written in the
    same call.  Poisoning it with 0xFF (NaN in fp16 / bf16 / fp32) between calls must not change a single bit."""
    cfg = S3Config(n_layer=2)
    engine.load_campplus(synth_weights.campplus_state(seed=11), CamPlusConfig())
    engine.load_s3tok(synth_weights.s3tok_state(cfg, seed=12), cfg)
    rng = np.random.default_rng(123)
    utts = [synth.utterance(rng, float(rng.uniform(0.4, 5.0))) for _ in range(9)]
    ref = _run(engine, utts, list(range(9)))
    for fill in (0xFF, 0x7F):
        engine._ws.fill_(fill)
        got = _run(engine, utts, list(range(9)))
        for a, b in zip(ref, got):
            if torch.is_tensor(a):
                assert torch.equal(a, b)
            else:
                assert a == b
    # a different batch shape over the poisoned buffer (offsets of every sub-buffer move)
    engine._ws.fill_(0xFF)
    sub = _run(engine, utts, [1, 4, 7])
    torch.cuda.synchronize()
    one = _run(engine, utts, [4])
    assert torch.equal(sub[4][1], one[4][0])                       # embedding of utterance 4
