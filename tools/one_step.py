"""Run N device-resident steps of stage 1+2 over a small synthetic set (a short command line for ncu)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

import bench
from fangyan_tts_b200.engine import Engine
from oracle import campplus as CP, s3tokenizer as S3


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 250
    prec = sys.argv[2] if len(sys.argv) > 2 else "bf16x3"
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    eng = Engine(0, prec)
    cam_cfg, s3_cfg = CP.CamPlusConfig(), S3.S3Config()
    eng.load_campplus(dict(CP.init_random(CP.CamPlus(cam_cfg)).state_dict()), cam_cfg)
    s3 = S3.init_random(S3.S3Tokenizer(s3_cfg))
    st = dict(s3.state_dict())
    st["rope.cos"], st["rope.sin"] = S3.rope_tables(2048, s3_cfg.d_head, s3_cfg.rope_theta)
    eng.load_s3tok(st, s3_cfg)
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
