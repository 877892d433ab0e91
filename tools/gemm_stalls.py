#!/usr/bin/env python3
"""Where do the GEMM roles wait?  Needs the debug build:
    FY_BUILD_TAG=prof FY_NVCC_EXTRA=-DFY_STALLPROF python -m fangyan_tts_b200.build
Runs one device-resident stage-2 step and prints, per GEMM kind, the share of each role's lifetime spent blocked:
producer on `empty`, MMA issuer on `full` (operands late) and on `tempty` (epilogue late), epilogue on `tfull`."""
import ctypes as C
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
os.environ.setdefault("FY_LIB", str(ROOT / "fangyan_tts_b200" / "libfangyan_b200_prof.so"))
sys.path.insert(0, str(ROOT))
import numpy as np
import torch

import bench
from fangyan_tts_b200 import _lib, synth_weights
from fangyan_tts_b200.configs import S3Config
from fangyan_tts_b200.engine import Engine


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 250
    prec = sys.argv[2] if len(sys.argv) > 2 else "bf16x3"
    eng = Engine(0, prec)
    cfg = S3Config()
    eng.load_s3tok(synth_weights.s3tok_state(cfg, seed=2), cfg)
    utts = bench.workload(n, 1988)
    lens = [len(u) for u in utts]
    pcm = torch.from_numpy(np.concatenate(utts)).cuda()
    lib = C.CDLL(os.environ["FY_LIB"])
    buf = (C.c_ulonglong * 40)()
    mel, mframes = eng.logmel(pcm, lens)
    eng.s3tok(mel, mframes)
    assert lib.fy_debug_gemm_stalls(buf, 1) == 0
    eng.s3tok(mel, mframes)
    assert lib.fy_debug_gemm_stalls(buf, 1) == 0
    names = ["QKV", "fc1", "O", "fc2", "other"]
    for s, nm in enumerate(names):
        v = [int(buf[8 * s + i]) for i in range(8)]
        if v[7] == 0:
            continue
        print("%-5s workers*launches=%d  producer: wait_empty %.1f%%   mma: wait_full %.1f%% wait_tempty %.1f%% "
              "(life %.2f Mclk/worker)   epilogue(warp4): wait_tfull %.1f%%" % (
                  nm, v[7], 100.0 * v[0] / max(v[1], 1), 100.0 * v[2] / max(v[4], 1), 100.0 * v[3] / max(v[4], 1),
                  v[4] / v[7] / 1e6, 100.0 * v[5] / max(v[6], 1)))


if __name__ == "__main__":
    main()
