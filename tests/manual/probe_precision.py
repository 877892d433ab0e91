"""GPU probe: CAMPPlus cosine and S3 token match of each precision mode against the CPU oracle (decides defaults)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent.parent))
import numpy as np
import torch

from fangyan_tts_b200 import synth, synth_weights
from fangyan_tts_b200.configs import CamPlusConfig, S3Config
from fangyan_tts_b200.engine import Engine
from oracle import campplus as CP, frontend as F, s3tokenizer as S3


def main():
    utts = synth.make_set(10, 77, 2.0, 10.0)
    cam = CP.CamPlus()
    cam.load_state_dict({k: torch.from_numpy(v) for k, v in synth_weights.campplus_state().items()}, strict=False)
    cam.eval()
    s3 = S3.S3Tokenizer()
    s3.load_state_dict({k: torch.from_numpy(v) for k, v in synth_weights.s3tok_state().items()})
    s3.eval()
    torch.set_num_threads(16)
    waves = [torch.from_numpy(u.astype(np.float32) / 32768.0)[None] for u in utts]
    with torch.no_grad():
        refs = [CP.embed_one(cam, F.cmn(F.kaldi_fbank(w))) for w in waves]
        mels = [F.whisper_log_mel(w) for w in waves]
        ids, pres = [], []
        for m in mels:
            hid, _ = s3.encode(m, torch.tensor([m.shape[2]]))
            pres.append(s3.fsq_pre(hid)[0])
            ids.append(torch.tensor(S3.tokenize_one(s3, m), dtype=torch.int32))
    ref_ids, ref_pre = torch.cat(ids), torch.cat(pres)
    pcm = torch.from_numpy(np.concatenate(utts)).cuda()
    lens = [len(u) for u in utts]
    for prec in ("bf16x3", "bf16"):
        eng = Engine(0, prec)
        eng.load_campplus(synth_weights.campplus_state(), CamPlusConfig())
        eng.load_s3tok(synth_weights.s3tok_state(), S3Config())
        feat, frames = eng.fbank_cmn(pcm, lens)
        emb = eng.campplus(feat, frames).cpu()
        cos = [float(torch.nn.functional.cosine_similarity(emb[i], refs[i], dim=0)) for i in range(len(utts))]
        mel, mframes = eng.logmel(pcm, lens)
        tok, counts, pre = eng.s3tok(mel, mframes, want_pre=True)
        tok, pre = tok.cpu(), pre.cpu()
        mism = (tok != ref_ids)
        d = (ref_pre.abs() - 0.5).abs().min(dim=1).values
        print(f"{prec}: campplus cos min={min(cos):.7f} mean={np.mean(cos):.7f} | tokens n={ref_ids.numel()} "
              f"match={1 - mism.float().mean():.5f} max|pre err|={float((pre - ref_pre).abs().max()):.3e} "
              f"mismatches off-boundary(>1e-3)={int((mism & (d > 1e-3)).sum())}")
        eng.close()


if __name__ == "__main__":
    main()
