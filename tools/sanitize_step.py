"""A small full-path batch (both front-ends, CAMPPlus, 2-layer S3 encoder + FSQ, speaker mean) for compute-sanitizer:
   compute-sanitizer --tool memcheck|racecheck|synccheck python tools/sanitize_step.py"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

from fangyan_tts_b200 import synth, synth_weights
from fangyan_tts_b200.configs import CamPlusConfig, S3Config
from fangyan_tts_b200.engine import Engine

prec = sys.argv[1] if len(sys.argv) > 1 else "bf16x3"
eng = Engine(0, prec)
cfg = S3Config(n_layer=2)
eng.load_campplus(synth_weights.campplus_state(), CamPlusConfig())
eng.load_s3tok(synth_weights.s3tok_state(cfg), cfg)
utts = synth.make_set(5, 3, 0.6, 4.0) + [synth.make_set(1, 5, 11.0, 12.0)[0]]     # 1 .. 5 key tiles per head
lens = [len(u) for u in utts]
pcm = torch.from_numpy(np.concatenate(utts)).cuda()
feat, frames = eng.fbank_cmn(pcm, lens)
emb = eng.campplus(feat, frames)
ssum = torch.zeros(2, 192, device="cuda")
scnt = torch.zeros(2, dtype=torch.int32, device="cuda")
eng.speaker_sum(emb, torch.tensor([0, 1, 0, 1, 0, 1], dtype=torch.int32, device="cuda"), 2, ssum, scnt)
mean = eng.speaker_mean(ssum, scnt)
mel, mframes = eng.logmel(pcm, lens)
tok, counts = eng.s3tok(mel, mframes)
torch.cuda.synchronize()
assert torch.isfinite(emb).all() and torch.isfinite(mean).all() and int(tok.min()) >= 0 and int(tok.max()) < 6561
print("ok", prec, counts)
