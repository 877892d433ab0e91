#!/usr/bin/env python3
"""Batch-1 latency of the CosyVoiceFrontEnd twin (contract B2, cosyvoice/cli/frontend.py:95-118) on one prompt:
_extract_speech_token and _extract_spk_embedding, eager launches against the CUDA-graph replay, 12 encoder layers.
Wall time per call with a device synchronise on both sides (file read + decode + upload + kernels + nothing read back:
the reference returns device tensors too) and the kernel part alone (CUDA events).  One JSON line."""
import json
import sys
import tempfile
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

from fangyan_tts_b200 import synth, synth_weights
from fangyan_tts_b200.configs import CamPlusConfig, S3Config
from fangyan_tts_b200.frontend import CosyVoiceFrontEndB200


def main():
    secs = float(sys.argv[1]) if len(sys.argv) > 1 else 5.0
    reps = 30
    tmp = Path(tempfile.mkdtemp(prefix="fy_twin_"))
    np.savez(tmp / "campplus.npz", **synth_weights.campplus_state(CamPlusConfig()))
    np.savez(tmp / "s3tok.npz", **synth_weights.s3tok_state(S3Config()))
    wav = tmp / "prompt.wav"
    synth.write_wav(wav, synth.utterance(np.random.default_rng(1986), secs))
    out = {"metric": "twin_batch1_latency_ms", "prompt_s": secs, "reps": reps, "layers": 12, "precision": "bf16x3"}
    for mode, use in (("eager", False), ("graph", True)):
        fe = CosyVoiceFrontEndB200(str(tmp / "campplus.npz"), str(tmp / "s3tok.npz"), use_graphs=use)
        for name, fn in (("_extract_speech_token", fe._extract_speech_token), ("_extract_spk_embedding", fe._extract_spk_embedding)):
            for _ in range(3):
                fn(str(wav))
            torch.cuda.synchronize()
            wall, dev = [], []
            for _ in range(reps):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                a.record()
                fn(str(wav))
                b.record()
                torch.cuda.synchronize()
                wall.append(1e3 * (time.perf_counter() - t0))
                dev.append(a.elapsed_time(b))
            out.setdefault(name, {})[mode] = {"wall_ms_median": float(np.median(wall)), "wall_ms_p90": float(np.percentile(wall, 90)),
                                              "stream_ms_median": float(np.median(dev))}
        fe.engine.close()
    for name in ("_extract_speech_token", "_extract_spk_embedding"):
        out[name]["speedup_wall"] = out[name]["eager"]["wall_ms_median"] / out[name]["graph"]["wall_ms_median"]
    print(json.dumps(out))


if __name__ == "__main__":
    main()
