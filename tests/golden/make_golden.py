#!/usr/bin/env python3
"""Generate tests/golden/stage12.npz: small input/output vectors for every stage of the hot path.

Run in the build container (needs torchaudio + transformers; nothing here runs on the GPU box):
    python tests/golden/make_golden.py

Sources of truth, per array:
  fbank_cmn_*  torchaudio.compliance.kaldi.fbank(num_mel_bins=80, dither=0, sample_frequency=16000) then feat - mean(0):
               the reference's literal callee and arguments (CosyVoice/tools/extract_embedding.py:29-33).
  resampled    torchaudio.transforms.Resample(24000, 16000): the literal callee (extract_embedding.py:27-28).
  logmel_*     transformers.WhisperFeatureExtractor(feature_size=128)._np_extract_fbank_features on the unpadded
               waveform: an independent implementation of whisper.log_mel_spectrogram(n_mels=128)
               (extract_speech_token.py:38; openai-whisper itself is not installable offline).
  emb_*, tok_*, pre_*  the CPU oracle networks (oracle/campplus.py, oracle/s3tokenizer.py; PARITY UNPINNED restatements)
               with fangyan_tts_b200.synth_weights seeds 1986, fed the arrays above, one utterance at a time.
"""
import json
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent.parent
sys.path.insert(0, str(ROOT))

import torchaudio  # noqa: E402
import torchaudio.compliance.kaldi as kaldi  # noqa: E402
import transformers  # noqa: E402

from fangyan_tts_b200 import synth, synth_weights  # noqa: E402
from oracle import campplus as CP, s3tokenizer as S3  # noqa: E402

SEED = 1986


def main():
    utts = synth.make_set(3, SEED, 0.4, 1.4)
    utts.append(np.zeros(400, np.int16))                       # exactly one fbank frame, all-silent: log floors
    rng = np.random.default_rng(SEED)
    stereo24 = np.stack([synth.utterance(rng, 0.5, 24000), synth.utterance(rng, 0.5, 24000)], axis=1)
    out = {"n_utts": np.array(len(utts)), "stereo24": stereo24}
    fe = transformers.WhisperFeatureExtractor(feature_size=128)
    cam = CP.CamPlus()
    cam.load_state_dict({k: torch.from_numpy(v) for k, v in synth_weights.campplus_state(seed=SEED).items()}, strict=False)
    cam.eval()
    s3 = S3.S3Tokenizer(S3.S3Config())
    s3.load_state_dict({k: torch.from_numpy(v) for k, v in synth_weights.s3tok_state(seed=SEED).items()})
    s3.eval()
    for i, u in enumerate(utts):
        w = torch.from_numpy(u.astype(np.float32) / 32768.0)[None]
        feat = kaldi.fbank(w, num_mel_bins=80, dither=0, sample_frequency=16000)
        feat = feat - feat.mean(dim=0, keepdim=True)
        mel = torch.from_numpy(np.asarray(fe._np_extract_fbank_features(w.numpy().astype(np.float64), "cpu"),
                                          dtype=np.float32))      # [1, 128, T]
        out[f"pcm_{i}"] = u
        out[f"fbank_cmn_{i}"] = feat.numpy()
        out[f"logmel_{i}"] = mel[0].numpy()
        with torch.no_grad():
            if feat.shape[0] >= 4:
                out[f"emb_{i}"] = CP.embed_one(cam, feat).numpy()
            hid, _ = s3.encode(mel, torch.tensor([mel.shape[2]]))
            out[f"pre_{i}"] = s3.fsq_pre(hid)[0].numpy()
            out[f"tok_{i}"] = np.array(S3.tokenize_one(s3, mel), dtype=np.int32)
    w24 = torch.from_numpy(stereo24.astype(np.float32).T / 32768.0)                     # [2, N]
    out["resampled"] = torchaudio.transforms.Resample(24000, 16000)(w24).numpy()
    meta = {"torch": torch.__version__, "torchaudio": torchaudio.__version__, "transformers": transformers.__version__,
            "seed": SEED, "reference_pins": {"torchaudio": "2.3.1", "openai-whisper": "20231117", "onnxruntime": "1.18.0"}}
    out["meta"] = np.array(json.dumps(meta))
    path = Path(__file__).resolve().parent / "stage12.npz"
    np.savez_compressed(path, **out)
    print(path, path.stat().st_size, "bytes", meta)


if __name__ == "__main__":
    main()
