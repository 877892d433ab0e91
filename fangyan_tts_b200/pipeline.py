"""Host pipeline for stages 1-2: length-sorted packed batches -> C-ABI kernels -> per-utterance results.

Mirrors tools/extract_embedding.py::single_job and tools/extract_speech_token.py::single_job of the reference, but
for a whole batch per launch sequence instead of one utterance per onnxruntime call.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import List, Sequence

import numpy as np
import torch

from .engine import Engine, kaldi_num_frames, token_len, whisper_num_frames

SR = 16000
MAX_TOKEN_SECONDS = 30  # tools/extract_speech_token.py:34: strictly longer than 30 s -> []
RAMP_UTTS = 48          # Extractor.extract: size of the first (ramp-up) batch


@dataclass
class Utt:
    """One decoded utterance: samples [N, C] (int16 or float32) and its sample rate."""
    samples: np.ndarray
    sr: int = SR


def plan_batches(durations: Sequence[float], max_seconds: float, max_utts: int) -> List[List[int]]:
    """Sort by duration and cut into batches bounded by total audio seconds and utterance count."""
    order = sorted(range(len(durations)), key=lambda i: durations[i])
    batches, cur, tot = [], [], 0.0
    for i in order:
        if cur and (tot + durations[i] > max_seconds or len(cur) >= max_utts):
            batches.append(cur)
            cur, tot = [], 0.0
        cur.append(i)
        tot += durations[i]
    if cur:
        batches.append(cur)
    return batches


def shard_indices(durations: Sequence[float], rank: int, world: int) -> List[int]:
    """Deal utterances (longest first) to the currently lightest rank: equal audio per rank, no duplication
    (unlike DistributedSampler, examples/grpo/cosyvoice2/infer_dataset.py:311, which pads by repeating samples)."""
    order = sorted(range(len(durations)), key=lambda i: -durations[i])
    load = [0.0] * world
    mine = []
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        load[r] += durations[i]
        if r == rank:
            mine.append(i)
    return sorted(mine)


class Extractor:
    def __init__(self, engine: Engine, max_batch_seconds: float = 2400.0, max_batch_utts: int = 1024):
        self.eng = engine
        self.max_batch_seconds = max_batch_seconds
        self.max_batch_utts = max_batch_utts
        self._pin = None
        self._copy_stream = None
        self._d2h_stream = None
        self._res = [(None, None), (None, None)]

    # ------------------------------------------------------------------ device-side preparation
    def _to_device_16k(self, utts: Sequence[Utt], mono: str):
        """Returns (packed pcm on device (int16 or float32), lengths).  mono = 'first' (stage 1: kaldi.fbank takes
        channel 0, extract_embedding.py:29) or 'mean' (stage 2, extract_speech_token.py:32-33).
        Fast path: everything 16 kHz mono int16 -> one pinned copy, int16 on the wire."""
        dev = self.eng.device
        simple = all(u.sr == SR and u.samples.dtype == np.int16 and (u.samples.shape[1] == 1 or mono == "first")
                     for u in utts)
        if simple:
            lens = [u.samples.shape[0] for u in utts]
            total = int(sum(lens))
            if self._pin is None or self._pin.numel() < total:
                self._pin = torch.empty(max(total, 1), dtype=torch.int16).pin_memory()
            host = self._pin[:total].numpy()
            o = 0
            for u, n in zip(utts, lens):
                host[o:o + n] = u.samples[:, 0]
                o += n
            return self._pin[:total].to(dev, non_blocking=True), lens
        # general path: float32, resample per sample-rate group on the device, then channel handling
        out, lens = [None] * len(utts), [0] * len(utts)
        groups = {}
        for i, u in enumerate(utts):
            x = u.samples
            x = x.astype(np.float32) / 32768.0 if x.dtype == np.int16 else x.astype(np.float32)
            chans = [x[:, 0]] if (mono == "first" or x.shape[1] == 1) else [x[:, c] for c in range(x.shape[1])]
            groups.setdefault(u.sr, []).append((i, chans))
        for sr, items in groups.items():
            flat = [c for _, chans in items for c in chans]
            ln = [len(c) for c in flat]
            packed = torch.from_numpy(np.concatenate(flat) if flat else np.zeros(0, np.float32)).to(dev)
            if sr != SR:
                packed, ln = self.eng.resample(packed, ln, sr, SR)
            o = 0
            k = 0
            for i, chans in items:
                nch = len(chans)
                n = ln[k]
                seg = packed[o:o + n * nch].view(nch, n) if nch > 1 else packed[o:o + n].view(1, n)
                out[i] = seg.mean(dim=0) if nch > 1 else seg[0]
                lens[i] = n
                o += n * nch
                k += nch
        return torch.cat(out) if out else torch.zeros(0, device=dev), lens

    # ------------------------------------------------------------------ stage 1
    def embed_batch(self, utts: Sequence[Utt]) -> torch.Tensor:
        """[n, 192] fp32 on the device for one batch."""
        pcm, lens = self._to_device_16k(utts, "first")
        for i, n in enumerate(lens):
            if kaldi_num_frames(n) < 3:
                raise ValueError(f"utterance {i} of the batch is too short for CAMPPlus ({n} samples at 16 kHz)")
        feat, frames = self.eng.fbank_cmn(pcm, lens)
        return self.eng.campplus(feat, frames)

    def embed(self, utts: Sequence[Utt]) -> np.ndarray:
        """Stage 1 for every utterance, original order: float32 [n, 192] on the host."""
        out = np.zeros((len(utts), 192), dtype=np.float32)
        dur = [u.samples.shape[0] / u.sr for u in utts]
        for idx in plan_batches(dur, self.max_batch_seconds, self.max_batch_utts):
            emb = self.embed_batch([utts[i] for i in idx]).cpu().numpy()
            out[idx] = emb
        return out

    # ------------------------------------------------------------------ both stages, one upload
    def extract(self, utts: Sequence[Utt]):
        """Stages 1 and 2 over the same utterances: (embeddings float32 [n, 192], token lists) on the host, original
        order.  16 kHz mono PCM16 input (what prepare_training_data.py writes) is uploaded once per batch as int16 and
        feeds both front-ends; results stay on the device until the last batch has been launched, so host staging of
        batch k+1 overlaps the kernels of batch k.  Anything else takes the two single-stage paths."""
        fused = all(u.sr == SR and u.samples.dtype == np.int16 and u.samples.shape[1] == 1 and
                    u.samples.shape[0] / SR <= MAX_TOKEN_SECONDS and u.samples.shape[0] >= 400 for u in utts)
        if not fused:
            return self.embed(utts), self.tokenize(utts)
        dur = [u.samples.shape[0] / SR for u in utts]
        batches = plan_batches(dur, self.max_batch_seconds, self.max_batch_utts)
        if len(batches[0]) > 2 * RAMP_UTTS:      # short ramp-up batch: the GPU starts after staging 48 clips, not 1,000
            batches[0:1] = [batches[0][:RAMP_UTTS], batches[0][RAMP_UTTS:]]
        total = sum(u.samples.shape[0] for u in utts)
        if self._pin is None or self._pin.numel() < total:
            self._pin = torch.empty(max(total, 1), dtype=torch.int16).pin_memory()
        host = self._pin.numpy()
        dev = self.eng.device
        main = torch.cuda.current_stream(dev)
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(dev)
            self._d2h_stream = torch.cuda.Stream(dev)
        out = np.zeros((len(utts), 192), dtype=np.float32)
        toks: List[List[int]] = [[] for _ in utts]

        def finish(item):
            """Host side of one batch: wait for its device-to-host copies, then unpack (runs under later batches)."""
            idx, counts, emb_h, tok_h, done = item
            done.synchronize()
            out[idx] = emb_h.numpy()
            th = tok_h.numpy()
            p = 0
            for i, n in zip(idx, counts):
                toks[i] = th[p:p + n].tolist()
                p += n

        pending, o = [], 0
        for k, idx in enumerate(batches):
            lens = [utts[i].samples.shape[0] for i in idx]
            b0 = o
            for i, n in zip(idx, lens):
                host[o:o + n] = utts[i].samples[:, 0]
                o += n
            # upload on a side stream: the copy of batch k+1 runs under the kernels of batch k
            with torch.cuda.stream(self._copy_stream):
                pcm = self._pin[b0:o].to(dev, non_blocking=True)
                ready = torch.cuda.Event()
                ready.record(self._copy_stream)
            main.wait_event(ready)
            pcm.record_stream(main)
            feat, frames = self.eng.fbank_cmn(pcm, lens)
            emb = self.eng.campplus(feat, frames)
            mel, mframes = self.eng.logmel(pcm, lens)
            tok, counts = self.eng.s3tok(mel, mframes)
            computed = torch.cuda.Event()
            computed.record(main)
            # results leave on a third stream into pinned buffers (two alternating sets), unpacked one batch later
            emb_h, tok_h = self._result_buffers(k & 1, emb.shape[0], tok.numel())
            with torch.cuda.stream(self._d2h_stream):
                self._d2h_stream.wait_event(computed)
                emb_h.copy_(emb, non_blocking=True)
                tok_h.copy_(tok, non_blocking=True)
                done = torch.cuda.Event()
                done.record(self._d2h_stream)
            emb.record_stream(self._d2h_stream)
            tok.record_stream(self._d2h_stream)
            pending.append((idx, counts, emb_h, tok_h, done))
            if len(pending) == 2:
                finish(pending.pop(0))
        for item in pending:
            finish(item)
        return out, toks

    def _result_buffers(self, slot: int, n_emb: int, n_tok: int):
        """Pinned host buffers for one batch's results (grown on demand, reused across calls)."""
        e, t = self._res[slot]
        if e is None or e.shape[0] < n_emb:
            e = torch.empty((max(n_emb, 256), 192), dtype=torch.float32).pin_memory()
        if t is None or t.numel() < n_tok:
            t = torch.empty(max(n_tok, 4096), dtype=torch.int32).pin_memory()
        self._res[slot] = (e, t)
        return e[:n_emb], t[:n_tok]

    # ------------------------------------------------------------------ stage 2
    def tokenize_batch(self, utts: Sequence[Utt]):
        """(packed int32 tokens on the device, token counts) for one batch; every utterance must pass the gate."""
        pcm, lens = self._to_device_16k(utts, "mean")
        mel, frames = self.eng.logmel(pcm, lens)
        return self.eng.s3tok(mel, frames)

    def tokenize(self, utts: Sequence[Utt]) -> List[List[int]]:
        """Stage 2 for every utterance, original order.  Longer than 30 s (after resampling) -> [] as in the
        reference; the gate is evaluated on the 16 kHz length like extract_speech_token.py:34."""
        res: List[List[int]] = [[] for _ in utts]
        keep, dur = [], []
        for i, u in enumerate(utts):
            n16 = u.samples.shape[0] if u.sr == SR else int(math.ceil(SR * u.samples.shape[0] / u.sr))
            if n16 / SR > MAX_TOKEN_SECONDS or whisper_num_frames(n16) < 1:
                continue
            keep.append(i)
            dur.append(n16 / SR)
        for b in plan_batches(dur, self.max_batch_seconds, self.max_batch_utts):
            idx = [keep[j] for j in b]
            tok, counts = self.tokenize_batch([utts[i] for i in idx])
            host = tok.cpu().numpy()
            o = 0
            for i, n in zip(idx, counts):
                res[i] = host[o:o + n].tolist()
                o += n
        return res
