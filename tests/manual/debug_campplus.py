"""GPU debugging aid: compare CAMPPlus intermediate activations with the oracle for one utterance."""
import ctypes as C
import sys
from pathlib import Path

import numpy as np
import torch
import torch.nn.functional as Fn

sys.path.insert(0, str(Path(__file__).resolve().parent.parent.parent))
from fangyan_tts_b200 import synth
from fangyan_tts_b200.engine import Engine
from oracle import campplus as CP
from oracle import frontend as F


def rel(a, b):
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def main():
    secs = float(sys.argv[1]) if len(sys.argv) > 1 else 1.5
    cfg = CP.CamPlusConfig()
    m = CP.init_random(CP.CamPlus(cfg), seed=1986)
    eng = Engine(0, "bf16x3")
    eng.load_campplus(dict(m.state_dict()), cfg)
    u = synth.make_set(1, 21, secs, secs + 0.01)[0]
    feat = F.cmn(F.kaldi_fbank(torch.from_numpy(u.astype(np.float32) / 32768.0)[None]))
    T = feat.shape[0]
    emb = eng.campplus(feat.cuda(), [T]).cpu()
    torch.cuda.synchronize()
    fo = np.array([0, T], dtype=np.int32)
    out = (C.c_int64 * 20)()
    eng.lib.fy_campplus_layout(eng.h, C.c_void_p(fo.ctypes.data), 1, out, 20)
    L = list(out)
    R50, R100 = L[0], L[1]
    ws = eng._ws

    def bf(off, n):
        return ws[off:off + n * 2].view(torch.bfloat16).float().cpu()

    def f32(off, n):
        return ws[off:off + n * 4].view(torch.float32).cpu()

    with torch.no_grad():
        x = feat.unsqueeze(0).permute(0, 2, 1).unsqueeze(1)
        h = m.head
        a = Fn.relu(h.bn1(h.conv1(x)))            # [1,32,80,T]
        got = f32(L[2], R100 * 82 * 32).view(R100, 82, 32)[2:2 + T, 1:81]   # [T,80,32]
        print("stem", rel(got, a[0].permute(2, 1, 0)))
        names = [(5, 42, 40), (6, 42, 40), (9, 22, 20), (10, 22, 20)]
        for i, blk in enumerate(h.blocks):
            a = blk(a)
            off, fp, f = names[i]
            got = f32(L[off], R100 * fp * 32).view(R100, fp, 32)
            print(f"res{i}", rel(got[2:2 + T, 1:1 + f], a[0].permute(2, 1, 0)),
                  "pads", float(got[:, 0].abs().max()), float(got[:, fp - 1].abs().max()), float(got[:2].abs().max()), float(got[2 + T:].abs().max()))
        a = Fn.relu(h.bn2(h.conv2(a)))             # [1,32,10,T]
        got = f32(L[11], R100 * 320).view(R100, 10, 32)
        print("fcm", rel(got[2:2 + T], a[0].permute(2, 1, 0)), "guards", float(got[:2].abs().max()), float(got[2 + T:].abs().max()))
        a = a.reshape(1, 320, T)
        a = m.tdnn_nl(m.tdnn(a))                   # [1,128,T2]
        T2 = a.shape[2]
        widths = [512, 1024, 1024, 512]
        cat = f32(L[12], R50 * 512).view(R50, 512)
        print("tdnn", rel(cat[:T2, :128], a[0].T), "T2", T2, "guards", float(cat[T2:, :128].abs().max()))
        for b, (layers, tr) in enumerate(zip(m.blocks, m.transits)):
            for l, lay in enumerate(layers):
                y = lay(a)
                if l < 2 or l == len(layers) - 1:
                    cat = f32(L[12 + b], R50 * widths[b]).view(R50, widths[b])
                    c0 = a.shape[1]
                    print(f"block{b} layer{l}", rel(cat[:T2, c0:c0 + 32], y[0].T))
                a = torch.cat([a, y], dim=1)
            a = tr(a)
            cat = f32(L[13 + b], R50 * widths[b + 1]).view(R50, widths[b + 1])
            print(f"transit{b}", rel(cat[:T2, :a.shape[1]], a[0].T))
        ref = CP.embed_one(m, feat)
        print("emb", rel(emb[0], ref), float(Fn.cosine_similarity(emb[0], ref, dim=0)))


if __name__ == "__main__":
    main()
