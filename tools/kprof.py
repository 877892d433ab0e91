"""Per-kernel time breakdown of one bench step using torch.profiler (CUPTI sees the library's kernels too)."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch
from torch.profiler import ProfilerActivity, profile

import bench
from fangyan_tts_b200.engine import Engine
from fangyan_tts_b200.pipeline import plan_batches
from fangyan_tts_b200 import synth_weights
from fangyan_tts_b200.configs import CamPlusConfig, S3Config


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 250
    prec = sys.argv[2] if len(sys.argv) > 2 else "bf16x3"
    eng = Engine(0, prec)
    cam_cfg, s3_cfg = CamPlusConfig(), S3Config()
    eng.load_campplus(synth_weights.campplus_state(), cam_cfg)          # product-side random-init weights
    eng.load_s3tok(synth_weights.s3tok_state(s3_cfg), s3_cfg)
    utts = bench.workload(n, 1988)
    lens = [len(u) for u in utts]
    pcm = torch.from_numpy(np.concatenate(utts)).cuda()

    def step():
        feat, frames = eng.fbank_cmn(pcm, lens)
        eng.campplus(feat, frames)
        mel, mframes = eng.logmel(pcm, lens)
        eng.s3tok(mel, mframes)

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        step()
        torch.cuda.synchronize()
    rows = {}
    evs = sorted([e for e in prof.events() if e.device_type.name == "CUDA"], key=lambda e: e.time_range.start)
    if os.environ.get("KPROF_SEQ"):
        for e in evs[:int(os.environ["KPROF_SEQ"])]:
            t = e.device_time if hasattr(e, "device_time") else e.cuda_time
            print(f"{t:9.1f} us {e.name.replace('(anonymous namespace)::', '').split('(')[0][:60]}")
    for e in prof.events():
        if e.device_type.name != "CUDA":
            continue
        name = e.name.replace("(anonymous namespace)::", "").split("(")[0].replace("void ", "")
        r = rows.setdefault(name, [0, 0.0])
        r[0] += 1
        r[1] += e.device_time if hasattr(e, "device_time") else e.cuda_time
    tot = sum(v[1] for v in rows.values())
    print(f"{n} utts, {sum(lens)/16000:.0f} audio-s, kernel time total {tot/1e3:.2f} ms")
    for k, v in sorted(rows.items(), key=lambda kv: -kv[1][1])[:24]:
        print(f"{v[1]/1e3:9.2f} ms {100*v[1]/tot:5.1f}% n={v[0]:4d} avg={v[1]/v[0]:8.1f}us {k[:80]}")


if __name__ == "__main__":
    main()
