"""north_star parity at BASELINE configs[0]/[1] shape, EVERY utterance against the oracle (not a sample):
100 utterances of 3-10 s, all 12 encoder layers (~16 k tokens), and a batch of 20-30 s clips (~20 k tokens).
Gates as written in BASELINE.json: embeddings cosine >= 0.9999; token ids >= 99.9 % equal with every mismatch within
1e-3 of an FSQ rounding boundary; token-sequence lengths bit-exact.  The oracle runs one utterance at a time at true
length on the host cores (bench.cpu_reference_run, a thread pool like the reference's), a few seconds for each case."""
import os
import sys
from pathlib import Path

import numpy as np
import pytest

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench
from fangyan_tts_b200 import synth, synth_weights
from fangyan_tts_b200.configs import CamPlusConfig, S3Config
from fangyan_tts_b200.pipeline import Extractor, Utt

pytestmark = pytest.mark.gpu


def _check(engine, utts, min_tokens):
    engine.load_campplus(synth_weights.campplus_state(), CamPlusConfig())
    engine.load_s3tok(synth_weights.s3tok_state(S3Config()), S3Config())
    emb, toks = Extractor(engine, max_batch_seconds=2400.0).extract([Utt(u[:, None]) for u in utts])
    ref = bench.cpu_reference_run(utts, len(os.sched_getaffinity(0)), want_outputs=True)
    rep = bench.parity_report(ref["out"], emb, toks)
    print(rep)
    assert rep["utts"] == len(utts) and rep["tokens"] >= min_tokens
    assert rep["len_exact"]
    assert rep["cos_min"] >= 0.9999, rep
    assert rep["token_match"] >= 0.999, rep
    assert rep["off_boundary_mismatches"] == 0, rep
    return rep


def test_config1_every_utterance_12_layers(engine):
    _check(engine, bench.workload(100, 1986 + 1), 15000)


def test_long_clips_every_utterance_12_layers(engine):
    rng = np.random.default_rng(1986 + 3)
    pool = [synth.utterance(rng, 30.0) for _ in range(4)]
    utts = [np.ascontiguousarray(pool[i % 4][: (480000 if i < 2 else int(rng.uniform(20.0, 30.0) * 16000))]) for i in range(32)]
    rep = _check(engine, utts, 18000)
    assert rep["tokens"] >= 2 * 750
