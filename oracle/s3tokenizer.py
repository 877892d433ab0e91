"""CPU oracle for the S3 speech tokenizer v3 encoder + FSQ (rows A9, A10) -- TEST INFRASTRUCTURE ONLY.

The reference runs an opaque `speech_tokenizer_v3.onnx` through onnxruntime
(CosyVoice/tools/extract_speech_token.py:40-48, cosyvoice/cli/frontend.py:99-103); its PyTorch
twin is the un-vendored `s3tokenizer` pip package (runtime/triton_trtllm/requirements.txt:2,
model_repo/audio_tokenizer/1/model.py:84-85).  Neither is under /root/reference.  What the
reference pins: inputs fp32 [1, 128, T] + int32 [1]; ids in [0, 6561) at 25 Hz
(examples/dialect/cosyvoice3/conf/cosyvoice3.yaml:13,26).

PARITY UNPINNED: restated from the published S3Tokenizer v2/v3 design (xingchensong/S3Tokenizer):
two stride-2 Conv1d + GELU, pre-LN transformer blocks with rotary position embedding on q/k, an
FSMN memory branch (depthwise Conv1d k=31 over the masked values, plus identity), GELU MLP, then
FSQ: Linear(d, 8) -> tanh -> *0.999 -> round -> +1 -> base-3 digits.  Hyper-parameters live in
S3Config.  Always run one utterance at a time at true length, as the reference does.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.nn as nn
import torch.nn.functional as F


@dataclass
class S3Config:
    n_mels: int = 128
    d_model: int = 1280
    n_head: int = 20
    n_layer: int = 12
    d_ff: int = 5120
    fsmn_kernel: int = 31
    fsq_dim: int = 8
    fsq_levels: int = 3
    rope_theta: float = 10000.0
    ln_eps: float = 1e-5            # attention LayerNorm (and the MLP one unless mlp_ln_eps is set)
    mlp_ln_eps: float | None = None  # the two norms of a block may carry different epsilons in the real graph (UNVERIFIED)
    fsq_scale: float = 0.9990000128746033

    @property
    def d_head(self):
        return self.d_model // self.n_head

    @property
    def vocab(self):
        return self.fsq_levels ** self.fsq_dim


def conv_out_len(t: int) -> int:
    """Conv1d(k=3, s=2, p=1) output length."""
    return (t + 2 - 3) // 2 + 1


def token_len(t_mel: int) -> int:
    return conv_out_len(conv_out_len(t_mel))


def rope_tables(n: int, d_head: int, theta: float):
    """cos/sin [n, d_head] with the two halves equal (rotate-half convention)."""
    inv = 1.0 / (theta ** (torch.arange(0, d_head, 2, dtype=torch.float32)[: d_head // 2] / d_head))
    ang = torch.outer(torch.arange(n, dtype=torch.float32), inv)
    ang = torch.cat([ang, ang], dim=-1)
    return ang.cos(), ang.sin()


def apply_rope(x, cos, sin):
    """x [B, T, H, D]; rotate-half: x*cos + cat(-x[D/2:], x[:D/2])*sin."""
    d = x.shape[-1]
    rot = torch.cat([-x[..., d // 2:], x[..., : d // 2]], dim=-1)
    return x * cos[None, :, None, :] + rot * sin[None, :, None, :]


class FsmnAttention(nn.Module):
    def __init__(self, cfg: S3Config):
        super().__init__()
        d = cfg.d_model
        self.cfg = cfg
        self.query = nn.Linear(d, d)
        self.key = nn.Linear(d, d, bias=False)
        self.value = nn.Linear(d, d)
        self.out = nn.Linear(d, d)
        self.fsmn = nn.Conv1d(d, d, cfg.fsmn_kernel, groups=d, bias=False)
        self.pad = (cfg.fsmn_kernel - 1) // 2

    def fsmn_memory(self, v, mask_pad):
        """FSMN memory of the value projection: depthwise conv over time (zero padding, no bias) + identity, masked
        before and after.  The same function as FunASR's SAN-M `forward_fsmn`, which S3Tokenizer's attention block
        derives from; tests/test_oracle_pins_cpu.py checks it against the copy of that code shipped in vllm."""
        vm = v * mask_pad
        mem = self.fsmn(F.pad(vm.transpose(1, 2), (self.pad, self.pad))).transpose(1, 2)
        return (mem + vm) * mask_pad

    def forward(self, x, mask_pad, bias, cos, sin):
        # x [B, T, d]; mask_pad [B, T, 1] (1 = valid); bias [B, 1, 1, T] additive
        b, t, d = x.shape
        h, dh = self.cfg.n_head, self.cfg.d_head
        q = self.query(x).view(b, t, h, dh)
        k = self.key(x).view(b, t, h, dh)
        v = self.value(x).view(b, t, h, dh)
        q, k = apply_rope(q, cos, sin), apply_rope(k, cos, sin)
        mem = self.fsmn_memory(v.reshape(b, t, d), mask_pad)
        scale = dh ** -0.25
        qk = (q.permute(0, 2, 1, 3) * scale) @ (k.permute(0, 2, 3, 1) * scale) + bias
        w = torch.softmax(qk.float(), dim=-1).to(q.dtype)
        o = (w @ v.permute(0, 2, 1, 3)).permute(0, 2, 1, 3).reshape(b, t, d)
        return self.out(o) + mem


class Block(nn.Module):
    def __init__(self, cfg: S3Config):
        super().__init__()
        d = cfg.d_model
        self.attn_ln = nn.LayerNorm(d, eps=cfg.ln_eps)
        self.attn = FsmnAttention(cfg)
        self.mlp_ln = nn.LayerNorm(d, eps=cfg.ln_eps if cfg.mlp_ln_eps is None else cfg.mlp_ln_eps)
        self.fc1 = nn.Linear(d, cfg.d_ff)
        self.fc2 = nn.Linear(cfg.d_ff, d)

    def forward(self, x, mask_pad, bias, cos, sin):
        x = x + self.attn(self.attn_ln(x), mask_pad, bias, cos, sin)
        return x + self.fc2(F.gelu(self.fc1(self.mlp_ln(x))))


class S3Tokenizer(nn.Module):
    def __init__(self, cfg: S3Config | None = None):
        super().__init__()
        self.cfg = cfg = cfg or S3Config()
        self.conv1 = nn.Conv1d(cfg.n_mels, cfg.d_model, 3, stride=2, padding=1)
        self.conv2 = nn.Conv1d(cfg.d_model, cfg.d_model, 3, stride=2, padding=1)
        self.blocks = nn.ModuleList(Block(cfg) for _ in range(cfg.n_layer))
        self.project_down = nn.Linear(cfg.d_model, cfg.fsq_dim)

    @staticmethod
    def _mask(lens, t):
        return (torch.arange(t)[None, :] < lens[:, None]).float()

    def encode(self, mel, mel_len):
        """mel [B, 128, T], mel_len [B] -> hidden [B, L, d], lens [B]."""
        m = self._mask(mel_len, mel.shape[2]).unsqueeze(1)
        x = F.gelu(self.conv1(mel * m))
        l1 = (mel_len + 2 - 3) // 2 + 1
        m = self._mask(l1, x.shape[2]).unsqueeze(1)
        x = F.gelu(self.conv2(x * m))
        l2 = (l1 + 2 - 3) // 2 + 1
        m = self._mask(l2, x.shape[2])
        x = x.permute(0, 2, 1)
        cos, sin = rope_tables(x.shape[1], self.cfg.d_head, self.cfg.rope_theta)
        mask_pad = m.unsqueeze(-1)
        bias = (1.0 - m)[:, None, None, :] * -1.0e10
        for blk in self.blocks:
            x = blk(x, mask_pad, bias, cos, sin)
        return x, l2

    def fsq_pre(self, hidden):
        """Pre-rounding FSQ value tanh(Wx+b)*0.999 in [-1, 1]; [.., 8]."""
        return torch.tanh(self.project_down(hidden).float()) * self.cfg.fsq_scale

    def quantize(self, mel, mel_len):
        hidden, lens = self.encode(mel, mel_len)
        h = self.fsq_pre(hidden).round() + 1.0
        powers = float(self.cfg.fsq_levels) ** torch.arange(self.cfg.fsq_dim, dtype=torch.float32)
        ids = (h * powers).sum(-1).int()
        return ids, lens


def init_random(model: S3Tokenizer, seed: int = 1986) -> S3Tokenizer:
    """Seeded random weights: fan-in scaled so activations neither vanish nor blow up through
    the stack; the FSQ projection is scaled so pre-tanh values spread over about (-2, 2) and
    all three levels occur."""
    g = torch.Generator().manual_seed(seed)
    cfg = model.cfg
    with torch.no_grad():
        for name, p in model.named_parameters():
            if p.dim() >= 2:
                fan_in = p[0].numel()
                std = (1.0 / fan_in) ** 0.5
                if name.endswith("out.weight") or name.endswith("fc2.weight"):
                    std *= 0.5
                if "fsmn" in name:
                    std = 0.1
                p.copy_(torch.randn(p.shape, generator=g) * std)
            elif "ln" in name and name.endswith("weight"):
                p.copy_(0.8 + 0.4 * torch.rand(p.shape, generator=g))
            else:
                p.copy_(torch.randn(p.shape, generator=g) * 0.05)
        # hidden rows have RMS ~ O(1..few); aim pre-tanh std ~ 1
        model.project_down.weight.mul_(0.25)
    return model.eval()


@torch.no_grad()
def tokenize_one(model: S3Tokenizer, mel: torch.Tensor):
    """mel [1, 128, T] -> list[int] of length token_len(T); batch 1 with feat_len = T,
    as extract_speech_token.py:40-48."""
    ids, lens = model.quantize(mel, torch.tensor([mel.shape[2]]))
    return ids[0, : int(lens[0])].tolist()
