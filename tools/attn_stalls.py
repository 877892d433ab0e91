#!/usr/bin/env python3
"""Where do the tcgen05 attention roles wait?  Needs the debug build:
    FY_BUILD_TAG=prof FY_NVCC_EXTRA=-DFY_STALLPROF python -m fangyan_tts_b200.build
Runs one device-resident stage-2 step and prints the share of each role's lifetime spent blocked on each barrier."""
import ctypes as C
import os

os.environ["FY_ATTN_TC"] = "1"
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
    buf = (C.c_ulonglong * 16)()
    mel, mframes = eng.logmel(pcm, lens)
    eng.s3tok(mel, mframes)
    assert lib.fy_debug_attn_stalls(buf, 1) == 0
    eng.s3tok(mel, mframes)
    assert lib.fy_debug_attn_stalls(buf, 1) == 0
    v = [int(b) for b in buf]
    n = max(v[15], 1)
    pl, il, sl = max(v[3], 1), max(v[10], 1), max(v[14], 1)
    print("CTAs*launches=%d  life/CTA: producer %.1f issuer %.1f softmax %.1f kclk" % (n, pl / n / 1e3, il / n / 1e3, sl / n / 1e3))
    print("producer: k_empty %.1f%% v_empty %.1f%%" % tuple(100.0 * v[i] / pl for i in (0, 1)))
    print("issuer  : k_full %.1f%% s_empty %.1f%% v_full %.1f%% p_full %.1f%% o_empty %.1f%% q_full %.1f%%" % tuple(
        100.0 * v[i] / il for i in (4, 5, 6, 7, 8, 9)))
    print("softmax (thread 0): s_full %.1f%% o_full %.1f%% bar.sync %.1f%%" % tuple(100.0 * v[i] / sl for i in (11, 12, 13)))


if __name__ == "__main__":
    main()
