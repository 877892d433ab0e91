"""Hyper-parameter tables of the two networks (product side; the oracle keeps its own copy under oracle/).

Neither architecture is defined under /root/reference: the reference only holds paths to `campplus.onnx` and
`speech_tokenizer_v3.onnx` (cosyvoice/cli/cosyvoice.py:204-209, tools/extract_embedding.py:94,
tools/extract_speech_token.py:90).  The defaults below restate the published CAM++ (3D-Speaker) and S3Tokenizer
v2/v3 designs and are UNVERIFIED against the real ONNX graphs, which is why they are data and not
constants baked into kernels.  Override with a JSON file via FY_CAMPPLUS_CONFIG / FY_S3TOK_CONFIG.
"""
from __future__ import annotations

import json
import os
from dataclasses import dataclass, fields


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
    def d_head(self) -> int:
        return self.d_model // self.n_head

    @property
    def vocab(self) -> int:
        return self.fsq_levels ** self.fsq_dim


def _from_env(cls, var: str):
    cfg = cls()
    path = os.environ.get(var)
    if path:
        with open(path) as f:
            d = json.load(f)
        known = {f.name for f in fields(cls)}
        unknown = set(d) - known
        if unknown:
            raise KeyError(f"{path}: unknown {cls.__name__} keys {sorted(unknown)}")
        for k, v in d.items():
            setattr(cfg, k, tuple(v) if isinstance(v, list) else v)
    return cfg


def campplus_config() -> CamPlusConfig:
    return _from_env(CamPlusConfig, "FY_CAMPPLUS_CONFIG")


def s3_config() -> S3Config:
    return _from_env(S3Config, "FY_S3TOK_CONFIG")
