#!/usr/bin/env python3
"""Where does the ping-pong tcgen05 attention kernel spend its time?  Needs the debug build:
    FY_BUILD_TAG=prof FY_NVCC_EXTRA=-DFY_STALLPROF python -m fangyan_tts_b200.build
Runs one device-resident stage-2 step and prints, per role, the share of its lifetime in each phase."""
import ctypes as C
import os
import sys
from pathlib import Path

os.environ["FY_ATTN_TC"] = "2"
ROOT = Path(__file__).resolve().parent.parent
os.environ.setdefault("FY_LIB", str(ROOT / "fangyan_tts_b200" / "libfangyan_b200_prof.so"))
sys.path.insert(0, str(ROOT))
import numpy as np
import torch

import bench
from fangyan_tts_b200 import synth, synth_weights
from fangyan_tts_b200.configs import S3Config
from fangyan_tts_b200.engine import Engine


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 250
    prec = sys.argv[2] if len(sys.argv) > 2 else "bf16x3"
    secs = float(sys.argv[3]) if len(sys.argv) > 3 else 0.0
    eng = Engine(0, prec)
    cfg = S3Config()
    eng.load_s3tok(synth_weights.s3tok_state(cfg, seed=2), cfg)
    if secs > 0:
        base = synth.utterance(np.random.default_rng(3), secs)
        utts = [base] * n
    else:
        utts = bench.workload(n, 1988)
    lens = [len(u) for u in utts]
    pcm = torch.from_numpy(np.concatenate(utts)).cuda()
    lib = C.CDLL(os.environ["FY_LIB"])
    buf = (C.c_ulonglong * 32)()
    mel, mframes = eng.logmel(pcm, lens)
    eng.s3tok(mel, mframes)
    assert lib.fy_debug_pp_stalls(buf, 1) == 0
    eng.s3tok(mel, mframes)
    assert lib.fy_debug_pp_stalls(buf, 1) == 0
    v = [int(b) for b in buf]
    n_cta = max(v[6], 1)
    il, sl, pl = max(v[5], 1), max(v[14], 1), max(v[18], 1)   # issuer: summed over the two issuing warps
    print(f"{prec} {len(utts)} utts ({'bench workload' if secs <= 0 else str(secs) + ' s clips'}): CTAs*launches={n_cta}  "
          f"life/CTA: issuer {il / n_cta / 2e3:.1f} softmax {sl / n_cta / 1e3:.1f} producer {pl / n_cta / 1e3:.1f} kclk, "
          f"prologue {v[19] / n_cta / 1e3:.2f} kclk")
    print("issuer  : q_full %.1f%% k_full %.1f%% v_full %.1f%% p_full %.1f%% o_empty %.1f%%" % tuple(100.0 * v[i] / il for i in range(5)))
    print("softmax (thread 0): s_full %.1f%% S load %.1f%% max+ex2 %.1f%% pack+store %.1f%% o_full %.1f%% head epilogue %.1f%% Q staging %.1f%%"
          % tuple(100.0 * v[i] / sl for i in (8, 9, 10, 11, 12, 13, 15)))
    print("  head epilogue: O load %.1f%% normalise+transpose %.1f%% stores %.1f%%;  of max+ex2: mask+max %.1f%%"
          % tuple(100.0 * v[i] / sl for i in (20, 21, 22, 23)))
    print("producer: k_empty %.1f%% v_empty %.1f%%" % tuple(100.0 * v[i] / pl for i in (16, 17)))


if __name__ == "__main__":
    main()
