"""can a TWO-product operand split meet the token gate?  Measured, not estimated.

The S3 encoder's linear layers / strided convolutions are bilinear, so what a tensor-core split computes can be
reproduced exactly (up to fp32 accumulation order) on any machine: round the operands to the pieces the scheme would
feed the MMA (bf16 / fp16 "hi", residual "lo"), run the listed piece products in fp32 and add them.  Attention's own
products and everything elementwise stay fp32, so a scheme's figure here is a LOWER bound of its error in a kernel.
Reference = the plain fp32 oracle.  Each row: pre-FSQ max |error|, token match, mismatches further than 1e-3 from a
rounding boundary (north_star gate: match >= 99.9 % and 0 such mismatches), on >= 10 k tokens of the bench workload.

    python tests/manual/probe_precision_emul.py [n_utts=64] > profiles/r02_precision_schemes.txt
"""
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent.parent))
import numpy as np
import torch
import torch.nn.functional as F

import bench
from fangyan_tts_b200 import synth_weights
from oracle import frontend as FE, s3tokenizer as S3

bf = lambda x: x.bfloat16().float()
hf = lambda x: x.half().float()


def pieces(x, fmt):
    hi = fmt(x)
    return hi, fmt(x - hi)


# scheme = (activation format, weight format, list of (activation piece, weight piece)); piece 0 = hi, 1 = lo
SCHEMES = {
    "fp32 (reference arithmetic)": None,
    "bf16 single pass (1 product)": (bf, bf, [(0, 0)]),
    "bf16x3: hi.hi + hi.lo + lo.hi (3 products, shipped)": (bf, bf, [(0, 0), (0, 1), (1, 0)]),
    "fp16 single pass (1 product)": (hf, hf, [(0, 0)]),
    "(a) fp16 hi+lo activations x fp16 weights (2 products)": (hf, hf, [(0, 0), (1, 0)]),
    "(b) fp16 activations x bf16 hi+lo weights (2 products)": (hf, bf, [(0, 0), (0, 1)]),
    "(b') fp16 activations x fp16 hi+lo weights (2 products)": (hf, hf, [(0, 0), (0, 1)]),
    "fp16x3: hi.hi + hi.lo + lo.hi (3 products)": (hf, hf, [(0, 0), (0, 1), (1, 0)]),
}
MIXED = "(c) bf16x3 on fc1/fc2, scheme (a) on conv/QKV/O"


def make_op(scheme_for):
    def op(kind, name, x, w, b, **kw):
        sch = scheme_for(name)
        f = F.linear if kind == "linear" else F.conv1d
        if sch is None:
            return f(x, w, b, **kw)
        fa, fw, prods = sch
        xp, wp = pieces(x, fa), pieces(w, fw)
        out = None
        for i, j in prods:
            y = f(xp[i], wp[j], None, **kw)
            out = y if out is None else out + y
        return out if b is None else out + (b.view(1, -1, 1) if kind == "conv" else b)
    return op


@torch.no_grad()
def encode(m, mel, op):
    """S3Tokenizer.encode + fsq_pre for one utterance at true length (oracle/s3tokenizer.py), GEMMs through `op`."""
    cfg = m.cfg
    x = F.gelu(op("conv", "conv1", mel, m.conv1.weight, m.conv1.bias, stride=2, padding=1))
    x = F.gelu(op("conv", "conv2", x, m.conv2.weight, m.conv2.bias, stride=2, padding=1))
    x = x.permute(0, 2, 1)
    t = x.shape[1]
    cos, sin = S3.rope_tables(t, cfg.d_head, cfg.rope_theta)
    h, dh = cfg.n_head, cfg.d_head
    for i, blk in enumerate(m.blocks):
        a = blk.attn
        y = blk.attn_ln(x)
        q = op("linear", "qkv", y, a.query.weight, a.query.bias).view(1, t, h, dh)
        k = op("linear", "qkv", y, a.key.weight, None).view(1, t, h, dh)
        v = op("linear", "qkv", y, a.value.weight, a.value.bias).view(1, t, h, dh)
        q, k = S3.apply_rope(q, cos, sin), S3.apply_rope(k, cos, sin)
        vm = v.reshape(1, t, -1)
        mem = a.fsmn(F.pad(vm.transpose(1, 2), (a.pad, a.pad))).transpose(1, 2) + vm
        sc = dh ** -0.25
        w = torch.softmax((q.permute(0, 2, 1, 3) * sc) @ (k.permute(0, 2, 3, 1) * sc), dim=-1)
        o = (w @ v.permute(0, 2, 1, 3)).permute(0, 2, 1, 3).reshape(1, t, -1)
        x = x + op("linear", "out", o, a.out.weight, a.out.bias) + mem
        y = blk.mlp_ln(x)
        x = x + op("linear", "fc2", F.gelu(op("linear", "fc1", y, blk.fc1.weight, blk.fc1.bias)), blk.fc2.weight, blk.fc2.bias)
    return torch.tanh(F.linear(x, m.project_down.weight, m.project_down.bias)) * cfg.fsq_scale


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    utts = bench.workload(n, 1986 + 2)
    m = S3.S3Tokenizer()
    m.load_state_dict({k: torch.from_numpy(v) for k, v in synth_weights.s3tok_state().items()})
    m.eval()
    mels = [FE.whisper_log_mel(torch.from_numpy(u.astype(np.float32) / 32768.0)[None]) for u in utts]
    pw = 3.0 ** torch.arange(8, dtype=torch.float32)
    ids_of = lambda pre: ((pre.round() + 1.0) * pw).sum(-1).int()
    runs = [(name, (lambda s: (lambda _n: s))(s)) for name, s in SCHEMES.items()]
    a, x3 = SCHEMES["(a) fp16 hi+lo activations x fp16 weights (2 products)"], SCHEMES["bf16x3: hi.hi + hi.lo + lo.hi (3 products, shipped)"]
    runs.append((MIXED, lambda name: x3 if name in ("fc1", "fc2") else a))
    ref = None
    print(f"# {n} utterances of the bench workload (seed 1988), 12 layers, random-init weights; torch {torch.__version__} CPU fp32")
    print(f"# {'scheme':58s} {'tokens':>7s} {'max|pre err|':>12s} {'rms err':>10s} {'match':>9s} {'mismatch':>8s} {'off-boundary':>12s}  gate")
    for name, sf in runs:
        t0 = time.time()
        pre = torch.cat([encode(m, mel, make_op(sf))[0] for mel in mels])
        if ref is None:
            ref = pre
            with torch.no_grad():
                check = torch.cat([m.fsq_pre(m.encode(mel, torch.tensor([mel.shape[2]]))[0])[0] for mel in mels[:4]])
            assert torch.equal(check, pre[: check.shape[0]]) or float((check - pre[: check.shape[0]]).abs().max()) < 3e-5, \
                "this file's encode() must restate the oracle's"
        err = (pre - ref).abs()
        bad = ids_of(pre) != ids_of(ref)
        dist = (ref.abs() - 0.5).abs().min(dim=1).values
        off = int((bad & (dist >= 1e-3)).sum())
        match = 1.0 - float(bad.float().mean())
        ok = "PASS" if match >= 0.999 and off == 0 else "FAIL"
        print(f"  {name:58s} {pre.shape[0]:7d} {float(err.max()):12.3e} {float(err.pow(2).mean().sqrt()):10.3e} {match:9.5f} "
              f"{int(bad.sum()):8d} {off:12d}  {ok}   ({time.time() - t0:.0f} s)", flush=True)


if __name__ == "__main__":
    main()
