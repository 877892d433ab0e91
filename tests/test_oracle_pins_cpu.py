"""Independent pins for sub-blocks of the S3-tokenizer oracle.  The whole network cannot be pinned (no
speech_tokenizer_v3.onnx, no s3tokenizer package offline: DESIGN.md section 2), but two of its building blocks exist in
this image as third-party code of the same lineage, written by other people: the rotate-half RoPE of
transformers' LLaMA, and the FSMN memory of FunASR's SAN-M attention as shipped in vllm.  CPU only."""
import subprocess
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent


def test_rope_matches_transformers_rotate_half():
    llama = pytest.importorskip("transformers.models.llama.modeling_llama")
    from oracle.s3tokenizer import apply_rope, rope_tables
    torch.manual_seed(3)
    b, t, h, d = 2, 37, 4, 64
    x = torch.randn(b, t, h, d)
    cos, sin = rope_tables(t, d, 10000.0)
    ours = apply_rope(x, cos, sin)
    # transformers: q [B, H, T, D], cos / sin [B, T, D], unsqueeze_dim=1
    q = x.permute(0, 2, 1, 3)
    theirs, _ = llama.apply_rotary_pos_emb(q, q, cos[None].expand(b, -1, -1), sin[None].expand(b, -1, -1))
    assert torch.equal(ours, theirs.permute(0, 2, 1, 3))
    # and the angle table is the standard one: position * theta^(-2i/d), both halves equal
    inv = 1.0 / (10000.0 ** (torch.arange(0, d, 2, dtype=torch.float32) / d))
    ang = torch.arange(t, dtype=torch.float32)[:, None] * inv[None]
    assert torch.equal(cos, torch.cat([ang, ang], -1).cos()) and torch.equal(sin, torch.cat([ang, ang], -1).sin())


_SANM_PROBE = r"""
import sys, types, torch
sys.path.insert(0, %r)
from vllm.model_executor.models.funasr import MultiHeadedAttentionSANM as SANM
from oracle.s3tokenizer import FsmnAttention, S3Config
torch.manual_seed(5)
cfg = S3Config(d_model=64, n_head=4, n_layer=1, fsmn_kernel=31)
att = FsmnAttention(cfg)
torch.nn.init.normal_(att.fsmn.weight, std=0.2)
b, t = 3, 83
v = torch.randn(b, t, cfg.d_model)
lens = torch.tensor([83, 40, 7])
mask = (torch.arange(t)[None] < lens[:, None]).float()[..., None]
ours = att.fsmn_memory(v, mask)
stub = types.SimpleNamespace(fsmn_block=att.fsmn, pad_fn=torch.nn.ConstantPad1d((15, 15), 0.0))
theirs = SANM.forward_fsmn(stub, v.clone(), mask)
print("MAXDIFF", float((ours - theirs).abs().max()), "EQUAL", bool(torch.equal(ours, theirs)))
"""


def test_fsmn_memory_matches_funasr_sanm_in_vllm():
    """Runs in a subprocess: importing vllm takes ~35 s on CPU and must not leak its state into this test session."""
    try:
        r = subprocess.run([sys.executable, "-c", _SANM_PROBE % str(ROOT)], capture_output=True, text=True, timeout=240)
    except subprocess.TimeoutExpired:
        pytest.skip("importing vllm took more than 240 s")
    line = [l for l in r.stdout.splitlines() if l.startswith("MAXDIFF")]
    if r.returncode != 0 or not line:
        if "No module named" in r.stderr or "ImportError" in r.stderr:
            pytest.skip("vllm's FunASR model is not importable here")
        pytest.fail(r.stderr[-1500:])
    assert line[0].endswith("EQUAL True"), line[0]
