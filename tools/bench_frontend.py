#!/usr/bin/env python3
"""BASELINE configs[4]: front-end-only sweep (fbank+CMN and log-mel kernels), batch 1-1024 fixed 10 s utterances,
reported against the HBM roofline (MEASURED_PEAKS.json).  Algorithmic bytes per audio-second:
fbank+CMN 64,000 B (int16 in + fp32 feat out), log-mel 83,200 B (int16 in + fp32 mel out).  Prints one JSON line per
batch size.  CUDA events on the launch stream, 3 warm-ups, inputs rotated through > 126 MB so L2 does not hold them."""
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

from fangyan_tts_b200 import synth
from fangyan_tts_b200.engine import Engine


def main():
    peaks = json.loads((Path(__file__).resolve().parent.parent / "MEASURED_PEAKS.json").read_text()) \
        if (Path(__file__).resolve().parent.parent / "MEASURED_PEAKS.json").exists() else {"hbm_gbs": 6650.0}
    hbm = peaks["hbm_gbs"]
    eng = Engine(0, "bf16x3")
    rng = np.random.default_rng(1986 + 4)
    base = synth.utterance(rng, 10.0)
    for batch in (1, 4, 16, 64, 256, 1024):
        n_sets = max(2, int(np.ceil(160e6 / (batch * len(base) * 2))))     # rotate > L2 worth of PCM
        n_sets = min(n_sets, 64)
        sets = [torch.from_numpy(np.tile(np.roll(base, 37 * k), batch)).cuda() for k in range(n_sets)]
        lens = [len(base)] * batch
        secs = batch * 10.0
        res = {"config": "front-end sweep", "batch": batch, "audio_s": secs}
        for name, fn, bytes_per_s in (("fbank_cmn", eng.fbank_cmn, 64000.0), ("logmel", eng.logmel, 83200.0)):
            for k in range(3):
                fn(sets[k % n_sets], lens)
            iters = max(5, min(50, int(2000 / batch) + 5))
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            a.record()
            for k in range(iters):
                fn(sets[k % n_sets], lens)
            b.record()
            torch.cuda.synchronize()
            ms = a.elapsed_time(b) / iters
            gbs = bytes_per_s * secs / (ms * 1e-3) / 1e9
            res[name] = {"ms": ms, "achieved_GBps": gbs, "frac_of_measured_hbm": gbs / hbm,
                         "audio_hours_per_s": secs / 3600.0 / (ms * 1e-3)}
        # row N4: 24 kHz matcha mel on the same number of 10 s clips (48,000 B int16 in + 16,000 B mel out per audio-s)
        base24 = np.resize(base, 240000)
        sets24 = [torch.from_numpy(np.tile(np.roll(base24, 37 * k), batch)).cuda() for k in range(min(n_sets, 8))]
        lens24 = [len(base24)] * batch
        for k in range(3):
            eng.matcha_mel(sets24[k % len(sets24)], lens24)
        iters = max(5, min(50, int(2000 / batch) + 5))
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        for k in range(iters):
            eng.matcha_mel(sets24[k % len(sets24)], lens24)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / iters
        gbs = 64000.0 * secs / (ms * 1e-3) / 1e9
        res["matcha_mel_24k"] = {"ms": ms, "achieved_GBps": gbs, "frac_of_measured_hbm": gbs / hbm,
                                 "audio_hours_per_s": secs / 3600.0 / (ms * 1e-3)}
        del sets24
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
