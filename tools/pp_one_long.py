"""Stage 2 over 64 thirty-second clips (750 tokens each, 12 key tiles per head), one step: a short command line for ncu."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

from fangyan_tts_b200 import synth, synth_weights
from fangyan_tts_b200.configs import S3Config
from fangyan_tts_b200.engine import Engine

eng = Engine(0, sys.argv[1] if len(sys.argv) > 1 else "bf16x3")
cfg = S3Config()
eng.load_s3tok(synth_weights.s3tok_state(cfg), cfg)
base = synth.utterance(np.random.default_rng(3), 30.0)
lens = [len(base)] * 64
pcm = torch.from_numpy(np.tile(base, 64)).cuda()
mel, mframes = eng.logmel(pcm, lens)
eng.s3tok(mel, mframes)
torch.cuda.synchronize()
print("ok")
