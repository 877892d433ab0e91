"""Seeded random-init weights of the two named architectures (checkpoints are unavailable offline).

Returns {tensor name: float32 ndarray} with the names fy_load_campplus / fy_load_s3tok look up (the state-dict names
of oracle/campplus.py and oracle/s3tokenizer.py, so tests can load the very same tensors into the CPU oracle).
Statistics: fan-in scaled normals so activations neither vanish nor saturate through the stack; BatchNorm running
statistics and affine terms moved away from (0, 1) so a folding bug shows; the FSQ projection scaled so pre-tanh
values spread over about (-2, 2) and all three levels occur.
"""
from __future__ import annotations

from typing import Dict

import numpy as np

from .configs import CamPlusConfig, S3Config


class _Shapes:
    """Stand-in for the random generator that records shapes instead of drawing values (campplus_shapes / s3tok_shapes)."""

    @staticmethod
    def standard_normal(shape, dtype=None):
        return np.broadcast_to(np.float32(0), shape if isinstance(shape, tuple) else (shape,))

    @staticmethod
    def random(n):
        return np.broadcast_to(np.float32(0), (n,))


class _ShapeDict(dict):
    def __setitem__(self, k, v):
        super().__setitem__(k, tuple(np.shape(v)))


def _bn(out: Dict[str, np.ndarray], rng, name: str, n: int, affine: bool = True):
    out[name + ".running_mean"] = (rng.standard_normal(n) * 0.1).astype(np.float32)
    out[name + ".running_var"] = (0.5 + rng.random(n)).astype(np.float32)
    if affine:
        out[name + ".weight"] = (0.8 + 0.4 * rng.random(n)).astype(np.float32)
        out[name + ".bias"] = (rng.standard_normal(n) * 0.1).astype(np.float32)


def _conv(out, rng, name: str, shape, bias: bool = False, gain: float = 1.6):
    fan_in = int(np.prod(shape[1:]))
    out[name + ".weight"] = (rng.standard_normal(shape, dtype=np.float32) * np.float32((gain / fan_in) ** 0.5))
    if bias:
        out[name + ".bias"] = (rng.standard_normal(shape[0]) * 0.1).astype(np.float32)


def campplus_shapes(cfg: CamPlusConfig | None = None) -> Dict[str, tuple]:
    """{tensor name: full shape} of everything fy_load_campplus looks up."""
    return dict(campplus_state(cfg, _shapes=True))


def s3tok_shapes(cfg: S3Config | None = None) -> Dict[str, tuple]:
    return dict(s3tok_state(cfg, _shapes=True))


def campplus_state(cfg: CamPlusConfig | None = None, seed: int = 1986, _shapes: bool = False) -> Dict[str, np.ndarray]:
    cfg = cfg or CamPlusConfig()
    rng = _Shapes() if _shapes else np.random.default_rng(seed)
    o: Dict[str, np.ndarray] = _ShapeDict() if _shapes else {}
    c = cfg.fcm_channels
    _conv(o, rng, "head.conv1", (c, 1, 3, 3))
    _bn(o, rng, "head.bn1", c)
    i = 0
    for nb in cfg.fcm_blocks:
        for j in range(nb):
            pre = f"head.blocks.{i}"
            _conv(o, rng, pre + ".conv1", (c, c, 3, 3))
            _bn(o, rng, pre + ".bn1", c)
            _conv(o, rng, pre + ".conv2", (c, c, 3, 3))
            _bn(o, rng, pre + ".bn2", c)
            if j == 0:
                _conv(o, rng, pre + ".sc_conv", (c, c, 1, 1))
                _bn(o, rng, pre + ".sc_bn", c)
            i += 1
    _conv(o, rng, "head.conv2", (c, c, 3, 3))
    _bn(o, rng, "head.bn2", c)
    ch = c * (cfg.feat_dim // 8)
    _conv(o, rng, "tdnn", (cfg.init_channels, ch, 5))
    _bn(o, rng, "tdnn_nl.0", cfg.init_channels)
    ch = cfg.init_channels
    bnc = cfg.bn_size * cfg.growth
    for b, (nl, k, _d) in enumerate(zip(cfg.block_layers, cfg.block_kernel, cfg.block_dilation)):
        for l in range(nl):
            pre = f"blocks.{b}.{l}"
            cin = ch + l * cfg.growth
            _bn(o, rng, pre + ".nl1.0", cin)
            _conv(o, rng, pre + ".lin1", (bnc, cin, 1))
            _bn(o, rng, pre + ".nl2.0", bnc)
            _conv(o, rng, pre + ".cam.local", (cfg.growth, bnc, k))
            _conv(o, rng, pre + ".cam.lin1", (bnc // 2, bnc, 1), bias=True)
            _conv(o, rng, pre + ".cam.lin2", (cfg.growth, bnc // 2, 1), bias=True)
        ch += nl * cfg.growth
        _bn(o, rng, f"transits.{b}.nl.0", ch)
        _conv(o, rng, f"transits.{b}.lin", (ch // 2, ch, 1))
        ch //= 2
    _bn(o, rng, "out_nl.0", ch)
    _conv(o, rng, "dense", (cfg.embed_dim, 2 * ch, 1))
    _bn(o, rng, "dense_bn", cfg.embed_dim, affine=False)
    return o


def s3tok_state(cfg: S3Config | None = None, seed: int = 1986, _shapes: bool = False) -> Dict[str, np.ndarray]:
    cfg = cfg or S3Config()
    rng = _Shapes() if _shapes else np.random.default_rng(seed)
    o: Dict[str, np.ndarray] = _ShapeDict() if _shapes else {}
    d, f = cfg.d_model, cfg.d_ff

    def lin(name, n_out, n_in, bias=True, gain=1.0, extra=()):
        shape = (n_out, n_in) + tuple(extra)
        fan_in = int(np.prod(shape[1:]))
        o[name + ".weight"] = rng.standard_normal(shape, dtype=np.float32) * np.float32(gain / fan_in ** 0.5)
        if bias:
            o[name + ".bias"] = (rng.standard_normal(n_out) * 0.05).astype(np.float32)

    def ln(name):
        o[name + ".weight"] = (0.8 + 0.4 * rng.random(d)).astype(np.float32)
        o[name + ".bias"] = (rng.standard_normal(d) * 0.05).astype(np.float32)

    lin("conv1", d, cfg.n_mels, extra=(3,))
    lin("conv2", d, d, extra=(3,))
    for i in range(cfg.n_layer):
        p = f"blocks.{i}."
        ln(p + "attn_ln")
        lin(p + "attn.query", d, d)
        lin(p + "attn.key", d, d, bias=False)
        lin(p + "attn.value", d, d)
        lin(p + "attn.out", d, d, gain=0.5)
        o[p + "attn.fsmn.weight"] = (rng.standard_normal((d, 1, cfg.fsmn_kernel)) * 0.1).astype(np.float32)
        ln(p + "mlp_ln")
        lin(p + "fc1", f, d)
        lin(p + "fc2", d, f, gain=0.5)
    lin("project_down", cfg.fsq_dim, d, gain=0.25)
    return o
