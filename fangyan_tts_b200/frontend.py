"""Online twin of stages 1-2 (contract B2): the two CosyVoiceFrontEnd methods, backed by the sm_100a kernels.

Mirrors cosyvoice/cli/frontend.py:95-118 of the reference (method names, argument meaning, return types, asserts):
    _extract_speech_token(prompt_wav)  -> (int32 [1, L], int32 [1]) on the device; asserts the clip is <= 30 s
    _extract_spk_embedding(prompt_wav) -> float32 [1, 192] on the device
    _extract_speech_feat(prompt_wav)   -> (float32 [1, T, 80], int32 [1]) on the device (24 kHz matcha mel, row N4)
and its loader cosyvoice/utils/file_utils.py:44-50 (channel MEAN first, then resample; sample rate must be >= 16 kHz)
-- note this differs from the stage CLIs, which take channel 0 for stage 1.
"""
from __future__ import annotations

import numpy as np
import torch

from .audio_io import read_wav
from .configs import campplus_config, s3_config
from .engine import Engine
from .weights import load_state

SR = 16000


class CosyVoiceFrontEndB200:
    """Construct with the same two model paths CosyVoiceFrontEnd takes (frontend.py:32-48)."""

    def __init__(self, campplus_model: str, speech_tokenizer_model: str, device: int = 0, precision: str = "bf16x3",
                 use_graphs: bool = True, max_graphs: int = 16):
        """use_graphs: a prompt is one utterance, so each method is ~30 (embedding) to ~130 (tokens) kernel launches of
        a few microseconds each and the launch path is the latency.  The launch sequence of a given prompt length is
        captured once in a CUDA graph (the C ABI's launch calls neither synchronise nor allocate) and replayed for every
        later prompt of that length; up to max_graphs lengths are kept."""
        self.engine = Engine(device, precision)
        self.device = self.engine.device
        self.use_graphs, self.max_graphs = use_graphs, max_graphs
        self._graphs = {}
        self.engine.cache_offsets = use_graphs

        def tensors(path, kind, cfg):
            return {k: np.ascontiguousarray(v, dtype=np.float32) for k, v in load_state(path, kind, cfg).items()
                    if v.dtype.kind == "f"}

        cam_cfg, s3_cfg = campplus_config(), s3_config()
        self.engine.load_campplus(tensors(campplus_model, "campplus", cam_cfg), cam_cfg)
        self.engine.load_s3tok(tensors(speech_tokenizer_model, "s3tok", s3_cfg), s3_cfg)

    def load_wav(self, wav, target_sr: int = SR, min_sr: int = SR) -> torch.Tensor:
        """file_utils.load_wav: float32 [N] mono at 16 kHz on the device."""
        x, sr = read_wav(wav, str(wav))
        x = x.astype(np.float32) / 32768.0 if x.dtype == np.int16 else x.astype(np.float32)
        mono = torch.from_numpy(np.ascontiguousarray(x.mean(axis=1, dtype=np.float32))).to(self.device)
        if sr != target_sr:
            assert sr >= min_sr, 'wav sample rate {} must be greater than {}'.format(sr, target_sr)
            mono, _ = self.engine.resample(mono, [mono.numel()], sr, target_sr)
        return mono

    def _graphed(self, kind: str, speech: torch.Tensor, fn):
        """fn(static speech tensor) -> output tensor, eagerly the first time a (kind, length) is seen -- which also puts
        the offsets arrays on the device and sizes the workspace --, then captured and replayed."""
        if not self.use_graphs:
            return fn(speech)
        eng = self.engine
        key = (kind, speech.numel())
        ent = self._graphs.get(key)
        if ent is not None and ent[3] != (eng._ws.data_ptr() if eng._ws is not None else 0):
            ent = None                                   # the workspace was regrown since: captured pointers are stale
        if ent is None:
            static_in = speech.clone()
            fn(static_in)
            torch.cuda.synchronize(self.device)
            g = torch.cuda.CUDAGraph()
            try:
                with torch.cuda.graph(g):
                    eng.use_stream(torch.cuda.current_stream(self.device))
                    out = fn(static_in)
            finally:
                eng.use_stream(torch.cuda.current_stream(self.device))
            if len(self._graphs) >= self.max_graphs:
                self._graphs.pop(next(iter(self._graphs)))
            ent = self._graphs[key] = (g, static_in, out, eng._ws.data_ptr())
        g, static_in, out, _ = ent
        static_in.copy_(speech)
        g.replay()
        return out.clone()

    def _extract_speech_token(self, prompt_wav):
        speech = self.load_wav(prompt_wav, SR)
        assert speech.numel() / SR <= 30, 'do not support extract speech token for audio longer than 30s'

        def run(x):
            mel, frames = self.engine.logmel(x, [x.numel()])
            return self.engine.s3tok(mel, frames)[0]

        tok = self._graphed("tok", speech, run)
        speech_token = tok.view(1, -1).to(torch.int32)
        speech_token_len = torch.tensor([speech_token.shape[1]], dtype=torch.int32, device=self.device)
        return speech_token, speech_token_len

    def _extract_spk_embedding(self, prompt_wav):
        speech = self.load_wav(prompt_wav, SR)

        def run(x):
            feat, frames = self.engine.fbank_cmn(x, [x.numel()])
            return self.engine.campplus(feat, frames)

        return self._graphed("spk", speech, run).view(1, -1)

    def _extract_speech_feat(self, prompt_wav):
        """frontend.py:120-125: load at 24 kHz -> matcha mel -> (float32 [1, T, 80], int32 [1]) on the device."""
        speech = self.load_wav(prompt_wav, 24000)
        mel, frames = self.engine.matcha_mel(speech, [speech.numel()])
        speech_feat = mel.view(1, frames[0], 80)
        speech_feat_len = torch.tensor([speech_feat.shape[1]], dtype=torch.int32, device=self.device)
        return speech_feat, speech_feat_len
