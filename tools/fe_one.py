"""One launch of each front-end at batch N x 10 s (a short command line for ncu)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

from fangyan_tts_b200 import synth
from fangyan_tts_b200.engine import Engine

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
eng = Engine(0, "bf16x3")
base = synth.utterance(np.random.default_rng(1986 + 4), 10.0)
pcm = torch.from_numpy(np.tile(base, n)).cuda()
lens = [len(base)] * n
for _ in range(int(sys.argv[2]) if len(sys.argv) > 2 else 2):
    eng.fbank_cmn(pcm, lens)
    eng.logmel(pcm, lens)
torch.cuda.synchronize()
print("ok")
