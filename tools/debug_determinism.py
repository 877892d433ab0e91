"""GPU debug aid: are per-utterance results independent of batch composition and repeatable?"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch
from fangyan_tts_b200 import synth, synth_weights
from fangyan_tts_b200.configs import CamPlusConfig, S3Config
from fangyan_tts_b200.engine import Engine

eng = Engine(0, "bf16x3")
eng.load_campplus(synth_weights.campplus_state(seed=11), CamPlusConfig())
cfg = S3Config(n_layer=2)
eng.load_s3tok(synth_weights.s3tok_state(cfg, seed=12), cfg)
utts = synth.make_set(9, 1986, 0.5, 4.0)


def run(idx):
    pcm = torch.from_numpy(np.concatenate([utts[i] for i in idx])).cuda()
    lens = [len(utts[i]) for i in idx]
    feat, frames = eng.fbank_cmn(pcm, lens)
    emb = eng.campplus(feat, frames)
    mel, mframes = eng.logmel(pcm, lens)
    tok, counts, pre = eng.s3tok(mel, mframes, want_pre=True)
    torch.cuda.synchronize()
    return feat.cpu(), frames, emb.cpu(), pre.cpu(), counts


fa, fra, ea, pa, ca = run(list(range(9)))
fb, frb, eb, pb, cb = run(list(range(9)))
print("repeat: feat", float((fa - fb).abs().max()), "emb", float((ea - eb).abs().max()), "pre", float((pa - pb).abs().max()))
sub = [0, 2, 5]
fs, frs, es, ps, cs = run(sub)
off = np.concatenate([[0], np.cumsum(fra)])
offp = np.concatenate([[0], np.cumsum(ca)])
o = 0
op = 0
for j, i in enumerate(sub):
    df = float((fa[off[i]:off[i + 1]] - fs[o:o + frs[j]]).abs().max())
    de = float((ea[i] - es[j]).abs().max() / ea[i].abs().max())
    dp = float((pa[offp[i]:offp[i + 1]] - ps[op:op + cs[j]]).abs().max())
    print(f"utt {i}: feat diff {df:.3e}  emb rel diff {de:.3e}  pre diff {dp:.3e}")
    o += frs[j]
    op += cs[j]
