"""Run N device-resident steps of stage 1+2 over a small synthetic set (a short command line for ncu)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

import bench
from fangyan_tts_b200.engine import Engine
from fangyan_tts_b200 import synth_weights
from fangyan_tts_b200.configs import CamPlusConfig, S3Config


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 250
    prec = sys.argv[2] if len(sys.argv) > 2 else "bf16x3"
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    eng = Engine(0, prec)
    cam_cfg, s3_cfg = CamPlusConfig(), S3Config()
    eng.load_campplus(synth_weights.campplus_state(), cam_cfg)          # product-side random-init weights
    eng.load_s3tok(synth_weights.s3tok_state(s3_cfg), s3_cfg)
    utts = bench.workload(n, 1988)
    lens = [len(u) for u in utts]
    pcm = torch.from_numpy(np.concatenate(utts)).cuda()
    for _ in range(steps):
        feat, frames = eng.fbank_cmn(pcm, lens)
        eng.campplus(feat, frames)
        mel, mframes = eng.logmel(pcm, lens)
        eng.s3tok(mel, mframes)
    torch.cuda.synchronize()
    print("ok")


if __name__ == "__main__":
    main()
