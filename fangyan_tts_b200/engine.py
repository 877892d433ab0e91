"""Python host side of the C ABI: one Engine per GPU.  PyTorch is used only for device memory and streams."""
from __future__ import annotations

import ctypes as C
import math
from typing import Sequence

import numpy as np
import torch

from . import _lib
from ._lib import FyCamConfig, FyS3Config, FyTensor, check

PREC = {"bf16": 0, "bf16x3": 1}
EMBED_DIM = 192


def kaldi_num_frames(n: int) -> int:
    """kaldi.fbank snip_edges=True frame count (extract_embedding.py:29-32)."""
    return 0 if n < 400 else 1 + (n - 400) // 160


def whisper_num_frames(n: int) -> int:
    """whisper.log_mel_spectrogram frame count (extract_speech_token.py:38)."""
    return n // 160


MATCHA_PAD = 720   # (n_fft - hop) / 2 of the 24 kHz mel


def matcha_num_frames(n: int) -> int:
    """Frames of matcha mel_spectrogram(n_fft 1920, hop 480, center=False) after its 720 + 720 reflect padding."""
    return 1 + (n - 480) // 480 if n >= 480 else 0


def token_len(t: int) -> int:
    """Two Conv1d(k=3, s=2, p=1) in the S3 encoder."""
    l1 = (t - 1) // 2 + 1
    return (l1 - 1) // 2 + 1


def _offsets(lengths: Sequence[int]) -> np.ndarray:
    off = np.zeros(len(lengths) + 1, dtype=np.int64)
    np.cumsum(np.asarray(lengths, dtype=np.int64), out=off[1:])
    if off[-1] >= 2 ** 31:
        raise ValueError("batch too large for int32 offsets")
    return off.astype(np.int32)


def _ptr(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


class Engine:
    """Owns a fy_ctx on one CUDA device.  Fails loudly when the library or a GPU is missing."""

    def __init__(self, device: int = 0, precision: str = "bf16x3"):
        self.lib = _lib.load()
        if not torch.cuda.is_available():
            raise _lib.FyError("no CUDA device: fangyan_b200 has no CPU path")
        self.device = torch.device("cuda", device)
        self.precision = precision
        torch.cuda.set_device(self.device)
        torch.cuda.init()
        torch.zeros(1, device=self.device)  # make sure the primary context exists
        h = C.c_void_p()
        check(self.lib.fy_create(device, PREC[precision], C.byref(h)))
        self.h = h
        self._ws = None
        self.s3_cfg = None
        # offsets arrays already on the device, keyed by their bytes: filled by an eager run, so that the same call
        # sequence can then be captured in a CUDA graph without any host-to-device copy (frontend.py, batch-1 twin)
        self.cache_offsets = False
        self._off_cache = {}
        self.use_stream(torch.cuda.current_stream(self.device))

    def close(self):
        if getattr(self, "h", None):
            self.lib.fy_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def use_stream(self, stream: torch.cuda.Stream):
        check(self.lib.fy_set_stream(self.h, C.c_void_p(stream.cuda_stream)))

    # ------------------------------------------------------------------ helpers
    def _dev_i32(self, a: np.ndarray) -> torch.Tensor:
        a = np.ascontiguousarray(a, dtype=np.int32)
        if not self.cache_offsets:
            return torch.from_numpy(a).to(self.device, non_blocking=True)
        key = a.tobytes()
        t = self._off_cache.get(key)
        if t is None:
            if len(self._off_cache) >= 256:
                self._off_cache.pop(next(iter(self._off_cache)))
            t = self._off_cache[key] = torch.from_numpy(a).to(self.device)
        return t

    def workspace(self, nbytes: int) -> torch.Tensor:
        if self._ws is None or self._ws.numel() < nbytes:
            self._ws = None
            self._ws = torch.empty(int(nbytes * 1.25) + 4096, dtype=torch.uint8, device=self.device)
        return self._ws

    @staticmethod
    def _pcm_dtype(pcm: torch.Tensor) -> int:
        if pcm.dtype == torch.int16:
            return 0
        if pcm.dtype == torch.float32:
            return 1
        raise TypeError(f"pcm must be int16 or float32, got {pcm.dtype}")

    # ------------------------------------------------------------------ weights
    def _tensor_table(self, state: dict):
        keep = []
        arr = (FyTensor * len(state))()
        for i, (name, t) in enumerate(state.items()):
            if isinstance(t, np.ndarray):
                if t.dtype.kind != "f":
                    t = np.zeros(0, np.float32) if t.dtype.kind in "iu" and t.ndim == 0 else t.astype(np.float32)
                a = np.ascontiguousarray(t, dtype=np.float32)
            else:
                a = np.ascontiguousarray(t.detach().cpu().to(torch.float32).numpy())
            keep.append(a)
            if a.ndim > 4:
                raise ValueError(f"{name}: rank {a.ndim} > 4")
            arr[i].name = name.encode()
            arr[i].h_data = a.ctypes.data
            arr[i].ndim = a.ndim
            for j, s in enumerate(a.shape):
                arr[i].shape[j] = s
        return arr, keep

    def load_campplus(self, state: dict, cfg):
        from .weights import check_shapes
        check_shapes(state, "campplus", cfg)          # full shapes (the C loader checks element counts)
        arr, keep = self._tensor_table(state)
        c = FyCamConfig()
        c.feat_dim, c.embed_dim, c.fcm_channels = cfg.feat_dim, cfg.embed_dim, cfg.fcm_channels
        c.init_channels, c.growth, c.bn_size, c.n_blocks = cfg.init_channels, cfg.growth, cfg.bn_size, len(cfg.block_layers)
        for i, (nl, k, d) in enumerate(zip(cfg.block_layers, cfg.block_kernel, cfg.block_dilation)):
            c.layers[i], c.kernel[i], c.dilation[i] = nl, k, d
        c.seg_len, c.bn_eps = cfg.seg_len, cfg.bn_eps
        check(self.lib.fy_load_campplus(self.h, arr, len(state), C.byref(c)))

    def load_s3tok(self, state: dict, cfg):
        from .weights import check_shapes
        check_shapes(state, "s3tok", cfg)
        arr, keep = self._tensor_table(state)
        c = FyS3Config(cfg.n_mels, cfg.d_model, cfg.n_head, cfg.n_layer, cfg.d_ff, cfg.fsmn_kernel, cfg.fsq_dim,
                       cfg.rope_theta, cfg.ln_eps, cfg.fsq_scale, cfg.mlp_ln_eps if cfg.mlp_ln_eps is not None else 0.0)
        check(self.lib.fy_load_s3tok(self.h, arr, len(state), C.byref(c)))
        self.s3_cfg = cfg

    # ------------------------------------------------------------------ A2
    def resample(self, wave: torch.Tensor, lengths: Sequence[int], orig_freq: int, new_freq: int = 16000):
        """wave: packed float32 [sum N]; returns (packed float32, out_lengths)."""
        g = math.gcd(orig_freq, new_freq)
        o, n = orig_freq // g, new_freq // g
        out_len = [int(math.ceil(n * l / o)) for l in lengths]
        in_off, out_off = _offsets(lengths), _offsets(out_len)
        out = torch.empty(int(out_off[-1]), dtype=torch.float32, device=self.device)
        d_in, d_out = self._dev_i32(in_off), self._dev_i32(out_off)
        check(self.lib.fy_resample_prepare(self.h, orig_freq, new_freq))     # builds the filter bank on first use
        check(self.lib.fy_resample(self.h, _ptr(wave), _ptr(d_in), _ptr(out), _ptr(d_out), len(lengths),
                                   int(out_off[-1]), orig_freq, new_freq))
        return out, out_len

    # ------------------------------------------------------------------ A3 + A4
    def fbank_cmn(self, pcm: torch.Tensor, lengths: Sequence[int]):
        """pcm: packed int16/float32 mono 16 kHz; returns (feat [sum M, 80] fp32, frame counts)."""
        frames = [kaldi_num_frames(l) for l in lengths]
        pcm_off, fr_off = _offsets(lengths), _offsets(frames)
        total = int(fr_off[-1])
        feat = torch.empty((total, 80), dtype=torch.float32, device=self.device)
        d_po, d_fo = self._dev_i32(pcm_off), self._dev_i32(fr_off)
        nb = self.lib.fy_fbank_workspace_bytes(len(lengths), total)
        ws = self.workspace(nb)
        check(self.lib.fy_fbank_cmn(self.h, _ptr(pcm), self._pcm_dtype(pcm), _ptr(d_po), _ptr(d_fo), len(lengths), total,
                                    _ptr(feat), _ptr(ws), ws.numel()))
        return feat, frames

    # ------------------------------------------------------------------ A5
    def logmel(self, pcm: torch.Tensor, lengths: Sequence[int]):
        """Returns (mel packed fp32: utterance b is [128, T_b] at 128 * frame_off[b], frame counts)."""
        frames = [whisper_num_frames(l) for l in lengths]
        pcm_off, fr_off = _offsets(lengths), _offsets(frames)
        total = int(fr_off[-1])
        mel = torch.empty(total * 128, dtype=torch.float32, device=self.device)
        d_po, d_fo = self._dev_i32(pcm_off), self._dev_i32(fr_off)
        nb = self.lib.fy_logmel_workspace_bytes(len(lengths), total)
        ws = self.workspace(nb)
        check(self.lib.fy_logmel(self.h, _ptr(pcm), self._pcm_dtype(pcm), _ptr(d_po), _ptr(d_fo), len(lengths), total,
                                 max(frames) if frames else 0, _ptr(mel), _ptr(ws), ws.numel()))
        return mel, frames

    # ------------------------------------------------------------------ N4: 24 kHz matcha mel
    def matcha_mel(self, pcm: torch.Tensor, lengths: Sequence[int]):
        """24 kHz packed PCM -> (mel packed fp32 [sum T_b, 80] time-major, frame counts).  The reference's
        feat_extractor (matcha mel_spectrogram, n_fft 1920 / hop 480 / center=False) followed by its callers'
        transpose(0, 1): cosyvoice/cli/frontend.py:120-125, cosyvoice/dataset/processor.py:156-181."""
        for i, n in enumerate(lengths):
            if n <= MATCHA_PAD:   # torch's reflect padding refuses these too
                raise ValueError(f"utterance {i}: {n} samples at 24 kHz, need more than {MATCHA_PAD} for reflect padding")
        frames = [matcha_num_frames(n) for n in lengths]
        pcm_off, fr_off = _offsets(lengths), _offsets(frames)
        total = int(fr_off[-1])
        mel = torch.empty((total, 80), dtype=torch.float32, device=self.device)
        d_po, d_fo = self._dev_i32(pcm_off), self._dev_i32(fr_off)
        check(self.lib.fy_matcha_mel(self.h, _ptr(pcm), self._pcm_dtype(pcm), _ptr(d_po), _ptr(d_fo), len(lengths),
                                     total, _ptr(mel)))
        return mel, frames

    # ------------------------------------------------------------------ A6
    def campplus(self, feat: torch.Tensor, frames: Sequence[int]) -> torch.Tensor:
        fr_off = _offsets(frames)
        d_fo = self._dev_i32(fr_off)
        emb = torch.empty((len(frames), EMBED_DIM), dtype=torch.float32, device=self.device)
        nb = self.lib.fy_campplus_workspace_bytes(self.h, len(frames), int(fr_off[-1]))
        ws = self.workspace(nb)
        check(self.lib.fy_campplus_forward(self.h, _ptr(feat), _ptr(d_fo), C.c_void_p(fr_off.ctypes.data), len(frames),
                                           _ptr(emb), _ptr(ws), ws.numel()))
        return emb

    # ------------------------------------------------------------------ A9 + A10
    def s3tok(self, mel: torch.Tensor, frames: Sequence[int], want_pre: bool = False):
        """mel as returned by logmel().  Returns (tokens packed int32, token counts[, fsq_pre [sum L, 8]])."""
        toks = [token_len(t) if t > 0 else 0 for t in frames]
        fr_off, tk_off = _offsets(frames), _offsets(toks)
        d_fo, d_to = self._dev_i32(fr_off), self._dev_i32(tk_off)
        total = int(tk_off[-1])
        out = torch.empty(total, dtype=torch.int32, device=self.device)
        pre = torch.empty((total, 8), dtype=torch.float32, device=self.device) if want_pre else None
        nb = self.lib.fy_s3tok_workspace_bytes(self.h, len(frames), int(fr_off[-1]))
        ws = self.workspace(nb)
        check(self.lib.fy_s3tok_encode_fsq(self.h, _ptr(mel), _ptr(d_fo), C.c_void_p(fr_off.ctypes.data), len(frames),
                                           _ptr(out), _ptr(d_to), _ptr(pre), _ptr(ws), ws.numel()))
        return (out, toks, pre) if want_pre else (out, toks)

    # ------------------------------------------------------------------ A7
    def speaker_sum(self, emb: torch.Tensor, spk_id: torch.Tensor, n_spk: int, ssum: torch.Tensor, count: torch.Tensor):
        check(self.lib.fy_speaker_sum(self.h, _ptr(emb), _ptr(spk_id), emb.shape[0], n_spk, _ptr(ssum), _ptr(count)))

    def speaker_mean(self, ssum: torch.Tensor, count: torch.Tensor) -> torch.Tensor:
        mean = torch.empty_like(ssum)
        check(self.lib.fy_speaker_mean_finish(self.h, _ptr(ssum), _ptr(count), ssum.shape[0], _ptr(mean)))
        return mean

    # ------------------------------------------------------------------ generic GEMM
    def gemm_bf16(self, a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
        """a [M, K] bf16, b [N, K] bf16 -> fp32 [M, N] = a @ b.T"""
        assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.is_contiguous() and b.is_contiguous()
        m, k = a.shape
        n = b.shape[0]
        out = torch.empty((m, n), dtype=torch.float32, device=self.device)
        check(self.lib.fy_gemm_bf16(self.h, _ptr(a), _ptr(b), _ptr(out), m, n, k))
        return out
