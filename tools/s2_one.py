"""One stage-2 step (log-mel -> S3 encoder -> FSQ) over N bench-workload utterances: a short command line for ncu
(gemm launches in order: conv1, conv2, then QKV / O / fc1 / fc2 per layer)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

import bench
from fangyan_tts_b200 import synth_weights
from fangyan_tts_b200.configs import S3Config
from fangyan_tts_b200.engine import Engine

n = int(sys.argv[1]) if len(sys.argv) > 1 else 370
eng = Engine(0, sys.argv[2] if len(sys.argv) > 2 else "bf16x3")
cfg = S3Config()
eng.load_s3tok(synth_weights.s3tok_state(cfg), cfg)
utts = bench.workload(n, 1988)
lens = [len(u) for u in utts]
pcm = torch.from_numpy(np.concatenate(utts)).cuda()
mel, mframes = eng.logmel(pcm, lens)
eng.s3tok(mel, mframes)
torch.cuda.synchronize()
print("ok", sum(lens) / 16000.0, "audio-s")
