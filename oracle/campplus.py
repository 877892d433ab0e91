"""CPU oracle for the CAMPPlus speaker network (row A6) -- TEST INFRASTRUCTURE ONLY.

The reference runs an opaque `campplus.onnx` through onnxruntime
(CosyVoice/tools/extract_embedding.py:36-41, cosyvoice/cli/frontend.py:115-116); neither the
file nor a definition of the network is under /root/reference.  The only I/O the reference pins
is input [1, M, 80] fp32 -> output [1, 192] fp32, M in [4, 3000]
(runtime/triton_trtllm/model_repo/speaker_embedding/1/model.py:84-89,112-113).

PARITY UNPINNED: this is a restatement, from the published CAM++ design (3D-Speaker
`speakerlab/models/campplus`, Wang et al. 2023), of the architecture that file is believed to hold.
All layer hyper-parameters live in CamPlusConfig so they can be corrected if a real ONNX graph
ever becomes available.  The network is always run one utterance at a time at its true length, as
the reference does; that unbatched run defines the masking semantics the batched kernels must match.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import torch
import torch.nn as nn
import torch.nn.functional as F


@dataclass
class CamPlusConfig:
    feat_dim: int = 80
    embed_dim: int = 192
    fcm_channels: int = 32
    fcm_blocks: tuple = (2, 2)
    init_channels: int = 128
    growth: int = 32
    bn_size: int = 4
    block_layers: tuple = (12, 24, 16)
    block_kernel: tuple = (3, 3, 3)
    block_dilation: tuple = (1, 2, 2)
    seg_len: int = 100
    bn_eps: float = 1e-5


def _bn_relu(c):
    return nn.Sequential(nn.BatchNorm1d(c), nn.ReLU())


class ResBlock2d(nn.Module):
    def __init__(self, cin, cout, stride):
        super().__init__()
        self.conv1 = nn.Conv2d(cin, cout, 3, stride=(stride, 1), padding=1, bias=False)
        self.bn1 = nn.BatchNorm2d(cout)
        self.conv2 = nn.Conv2d(cout, cout, 3, stride=1, padding=1, bias=False)
        self.bn2 = nn.BatchNorm2d(cout)
        self.has_sc = stride != 1 or cin != cout
        if self.has_sc:
            self.sc_conv = nn.Conv2d(cin, cout, 1, stride=(stride, 1), bias=False)
            self.sc_bn = nn.BatchNorm2d(cout)

    def forward(self, x):
        y = F.relu(self.bn1(self.conv1(x)))
        y = self.bn2(self.conv2(y))
        s = self.sc_bn(self.sc_conv(x)) if self.has_sc else x
        return F.relu(y + s)


class FCM(nn.Module):
    """2-D convolutional head over (freq, time); frequency is strided 80 -> 40 -> 20 -> 10."""

    def __init__(self, cfg: CamPlusConfig):
        super().__init__()
        c = cfg.fcm_channels
        self.conv1 = nn.Conv2d(1, c, 3, stride=1, padding=1, bias=False)
        self.bn1 = nn.BatchNorm2d(c)
        blocks = []
        for nb in cfg.fcm_blocks:
            for i in range(nb):
                blocks.append(ResBlock2d(c, c, 2 if i == 0 else 1))
        self.blocks = nn.ModuleList(blocks)
        self.conv2 = nn.Conv2d(c, c, 3, stride=(2, 1), padding=1, bias=False)
        self.bn2 = nn.BatchNorm2d(c)
        self.out_channels = c * (cfg.feat_dim // 8)

    def forward(self, x):            # x [B, F, T]
        x = x.unsqueeze(1)
        x = F.relu(self.bn1(self.conv1(x)))
        for b in self.blocks:
            x = b(x)
        x = F.relu(self.bn2(self.conv2(x)))
        return x.reshape(x.shape[0], x.shape[1] * x.shape[2], x.shape[3])


class CamLayer(nn.Module):
    """Context-aware-masking TDNN layer: local dilated conv gated by a sigmoid mask computed
    from the global mean plus the 100-frame segment mean of the layer input."""

    def __init__(self, cin, cout, k, dilation, seg_len):
        super().__init__()
        self.local = nn.Conv1d(cin, cout, k, padding=(k - 1) // 2 * dilation, dilation=dilation, bias=False)
        self.lin1 = nn.Conv1d(cin, cin // 2, 1)
        self.lin2 = nn.Conv1d(cin // 2, cout, 1)
        self.seg_len = seg_len

    def seg_pool(self, x):
        seg = F.avg_pool1d(x, self.seg_len, self.seg_len, ceil_mode=True)
        seg = seg.unsqueeze(-1).expand(*seg.shape, self.seg_len).reshape(*seg.shape[:-1], -1)
        return seg[..., : x.shape[-1]]

    def forward(self, x):
        y = self.local(x)
        ctx = x.mean(-1, keepdim=True) + self.seg_pool(x)
        m = torch.sigmoid(self.lin2(F.relu(self.lin1(ctx))))
        return y * m


class DenseLayer(nn.Module):
    def __init__(self, cin, cfg: CamPlusConfig, k, dilation):
        super().__init__()
        bnc = cfg.bn_size * cfg.growth
        self.nl1 = _bn_relu(cin)
        self.lin1 = nn.Conv1d(cin, bnc, 1, bias=False)
        self.nl2 = _bn_relu(bnc)
        self.cam = CamLayer(bnc, cfg.growth, k, dilation, cfg.seg_len)

    def forward(self, x):
        return self.cam(self.nl2(self.lin1(self.nl1(x))))


class Transit(nn.Module):
    def __init__(self, cin, cout):
        super().__init__()
        self.nl = _bn_relu(cin)
        self.lin = nn.Conv1d(cin, cout, 1, bias=False)

    def forward(self, x):
        return self.lin(self.nl(x))


class CamPlus(nn.Module):
    def __init__(self, cfg: CamPlusConfig | None = None):
        super().__init__()
        self.cfg = cfg = cfg or CamPlusConfig()
        self.head = FCM(cfg)
        c = self.head.out_channels
        self.tdnn = nn.Conv1d(c, cfg.init_channels, 5, stride=2, padding=2, bias=False)
        self.tdnn_nl = _bn_relu(cfg.init_channels)
        c = cfg.init_channels
        self.blocks = nn.ModuleList()
        self.transits = nn.ModuleList()
        for nl, k, d in zip(cfg.block_layers, cfg.block_kernel, cfg.block_dilation):
            layers = nn.ModuleList()
            for i in range(nl):
                layers.append(DenseLayer(c + i * cfg.growth, cfg, k, d))
            self.blocks.append(layers)
            c += nl * cfg.growth
            self.transits.append(Transit(c, c // 2))
            c //= 2
        self.out_nl = _bn_relu(c)
        self.dense = nn.Conv1d(2 * c, cfg.embed_dim, 1, bias=False)
        self.dense_bn = nn.BatchNorm1d(cfg.embed_dim, affine=False)

    def forward(self, feat):          # feat [B, T, 80]
        x = self.head(feat.permute(0, 2, 1))
        x = self.tdnn_nl(self.tdnn(x))
        for layers, tr in zip(self.blocks, self.transits):
            for l in layers:
                x = torch.cat([x, l(x)], dim=1)
            x = tr(x)
        x = self.out_nl(x)
        stats = torch.cat([x.mean(-1), x.std(-1, unbiased=True)], dim=-1)
        return self.dense_bn(self.dense(stats.unsqueeze(-1))).squeeze(-1)


def init_random(model: CamPlus, seed: int = 1986) -> CamPlus:
    """Seeded random weights: fan-in scaled normals, BatchNorm running
    statistics and affine terms moved away from (0, 1) so folding errors show."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for m in model.modules():
            if isinstance(m, (nn.Conv1d, nn.Conv2d)):
                fan_in = m.weight[0].numel()
                m.weight.copy_(torch.randn(m.weight.shape, generator=g) * (1.6 / fan_in) ** 0.5)
                if m.bias is not None:
                    m.bias.copy_(torch.randn(m.bias.shape, generator=g) * 0.1)
            elif isinstance(m, (nn.BatchNorm1d, nn.BatchNorm2d)):
                n = m.num_features
                m.running_mean.copy_(torch.randn(n, generator=g) * 0.1)
                m.running_var.copy_(0.5 + torch.rand(n, generator=g))
                if m.affine:
                    m.weight.copy_(0.8 + 0.4 * torch.rand(n, generator=g))
                    m.bias.copy_(torch.randn(n, generator=g) * 0.1)
    return model.eval()


@torch.no_grad()
def embed_one(model: CamPlus, feat: torch.Tensor) -> torch.Tensor:
    """feat [M, 80] (already CMN'ed) -> [192]; batch 1, as extract_embedding.py:36-41."""
    return model(feat.unsqueeze(0))[0]
