"""Host pipeline for stages 1-2: length-sorted packed batches -> C-ABI kernels -> per-utterance results.

Mirrors tools/extract_embedding.py::single_job and tools/extract_speech_token.py::single_job of the reference, but
for a whole batch per launch sequence instead of one utterance per onnxruntime call.
"""
from __future__ import annotations

import math
import time
from dataclasses import dataclass
from typing import List, Optional, Sequence

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


def shard_assignment(durations: Sequence[float], world: int) -> np.ndarray:
    """rank_of[i] for every utterance: deal utterances (longest first) to the currently lightest rank -- equal audio per
    rank, no duplication (unlike DistributedSampler, examples/grpo/cosyvoice2/infer_dataset.py:311, which pads by
    repeating samples).  Deterministic, so every rank derives the whole job's layout without communicating."""
    import heapq
    order = sorted(range(len(durations)), key=lambda i: -durations[i])
    heap = [(0.0, k) for k in range(world)]
    rank_of = np.zeros(len(durations), dtype=np.int32)
    for i in order:
        load, k = heapq.heappop(heap)
        rank_of[i] = k
        heapq.heappush(heap, (load + durations[i], k))
    return rank_of


def shard_indices(durations: Sequence[float], rank: int, world: int) -> List[int]:
    """This rank's utterances under shard_assignment, in wav.scp order."""
    return np.nonzero(shard_assignment(durations, world) == rank)[0].tolist()


def split_tokens(flat: np.ndarray, counts: np.ndarray) -> List[List[int]]:
    """Packed ids + per-utterance counts -> list of Python int lists (the reference's value type,
    extract_speech_token.py:48).  One tolist() of the whole array, then list slices: no per-utterance numpy call."""
    ids = flat.tolist()
    out, o = [], 0
    for n in counts.tolist():
        out.append(ids[o:o + n])
        o += n
    return out


class PinnedRing:
    """Host->device staging of packed int16 PCM: `slots` pinned buffers used in turn, each guarded by the event recorded
    after its last copy, so a slot is never rewritten while the DMA that reads it may still be queued (the copy runs on
    a side stream under the previous batch's kernels; the caller's stream waits for it)."""

    def __init__(self, device: torch.device, slots: int = 4):
        self.device = device
        self.slots: List[Optional[torch.Tensor]] = [None] * slots
        self.events: List[Optional[torch.cuda.Event]] = [None] * slots
        self.k = 0
        self.stream: Optional[torch.cuda.Stream] = None
        self.pack_seconds = 0.0

    def acquire(self, total: int):
        """Next slot (after its previous copy has finished), grown to `total` samples: (slot index, int16 numpy view)."""
        i = self.k % len(self.slots)
        self.k += 1
        if self.events[i] is not None:
            self.events[i].synchronize()             # the previous copy out of this slot has finished
        if self.slots[i] is None or self.slots[i].numel() < total:
            self.slots[i] = None
            self.slots[i] = torch.empty(max(int(total * 1.25), 1 << 16), dtype=torch.int16).pin_memory()
        return i, self.slots[i][:total].numpy()

    def submit(self, i: int, total: int) -> torch.Tensor:
        """Start the host->device copy of slot i's first `total` samples on the side stream; the current stream waits
        for it.  Returns the device tensor."""
        if self.stream is None:
            self.stream = torch.cuda.Stream(self.device)
        main = torch.cuda.current_stream(self.device)
        with torch.cuda.stream(self.stream):
            dev = self.slots[i][:total].to(self.device, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.stream)
        self.events[i] = ev
        main.wait_event(ev)
        dev.record_stream(main)
        return dev

    def upload(self, arrays: Sequence[np.ndarray], total: int) -> torch.Tensor:
        """Concatenate 1-D int16 arrays into the next slot and start its copy; returns the device tensor, which the
        current stream may use (it waits for the copy)."""
        t0 = time.perf_counter()
        i, view = self.acquire(total)
        if arrays:
            np.concatenate(arrays, out=view)
        dev = self.submit(i, total)
        self.pack_seconds += time.perf_counter() - t0
        return dev


class Extractor:
    def __init__(self, engine: Engine, max_batch_seconds: float = 2400.0, max_batch_utts: int = 1024):
        self.eng = engine
        self.max_batch_seconds = max_batch_seconds
        self.max_batch_utts = max_batch_utts
        self.ring = PinnedRing(engine.device)
        self._d2h_stream = None
        self._res = [(None, None), (None, None)]

    # ------------------------------------------------------------------ device-side preparation
    def _to_device_16k(self, utts: Sequence[Utt], mono: str):
        """Returns (packed pcm on device (int16 or float32), lengths).  mono = 'first' (stage 1: kaldi.fbank takes
        channel 0, extract_embedding.py:29) or 'mean' (stage 2, extract_speech_token.py:32-33).
        Fast path: everything 16 kHz mono int16 -> one pinned copy on the upload ring, int16 on the wire."""
        dev = self.eng.device
        simple = all(u.sr == SR and u.samples.dtype == np.int16 and (u.samples.shape[1] == 1 or mono == "first")
                     for u in utts)
        if simple:
            lens = [u.samples.shape[0] for u in utts]
            return self.ring.upload([u.samples[:, 0] for u in utts], int(sum(lens))), lens
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
        return self.embed_pcm(*self._to_device_16k(utts, "first"))

    def embed_pcm(self, pcm: torch.Tensor, lens: Sequence[int]) -> torch.Tensor:
        """Stage 1 on packed mono 16 kHz PCM already on the device (int16 or float32)."""
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
    def extract(self, utts: Sequence[Utt], shard=None):
        """Stages 1 and 2 over the same utterances: (embeddings float32 [n, 192], token lists) on the host, original
        order.  16 kHz mono PCM16 input (what prepare_training_data.py writes) is uploaded once per batch as int16 and
        feeds both front-ends; results stay on the device until the last batch has been launched, so host staging of
        batch k+1 overlaps the kernels of batch k.  Anything else takes the two single-stage paths.

        shard = (index, n_total): `utts` are this rank's share of an n_total-utterance job (index[j] = position of
        utts[j] in the job, see shard_indices).  Results then stay on the device, are gathered to rank 0 over the
        process group (distributed.gather_stage1/2) and only rank 0 returns (embeddings [n_total, 192], n_total token
        lists); the other ranks return (None, None)."""
        if shard is not None:
            return self._extract_sharded(utts, shard[0], shard[1])
        fused = all(u.sr == SR and u.samples.dtype == np.int16 and u.samples.shape[1] == 1 and
                    u.samples.shape[0] / SR <= MAX_TOKEN_SECONDS and u.samples.shape[0] >= 400 for u in utts)
        if not fused:
            return self.embed(utts), self.tokenize(utts)
        dur = [u.samples.shape[0] / SR for u in utts]
        batches = plan_batches(dur, self.max_batch_seconds, self.max_batch_utts)
        if len(batches[0]) > 2 * RAMP_UTTS:      # short ramp-up batch: the GPU starts after staging 48 clips, not 1,000
            batches[0:1] = [batches[0][:RAMP_UTTS], batches[0][RAMP_UTTS:]]
        dev = self.eng.device
        main = torch.cuda.current_stream(dev)
        if self._d2h_stream is None:
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

        pending = []
        for k, idx in enumerate(batches):
            lens = [utts[i].samples.shape[0] for i in idx]
            # upload on the ring's side stream: the copy of batch k+1 runs under the kernels of batch k
            pcm = self.ring.upload([utts[i].samples[:, 0] for i in idx], int(sum(lens)))
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

    def _extract_sharded(self, utts: Sequence[Utt], index: Sequence[int], n_total: int):
        from . import distributed as D
        for u in utts:
            if not (u.sr == SR and u.samples.dtype == np.int16 and u.samples.shape[1] == 1 and
                    400 <= u.samples.shape[0] <= MAX_TOKEN_SECONDS * SR):
                raise ValueError("sharded extract takes 16 kHz mono PCM16 clips of 25 ms .. 30 s (the CLIs handle the rest)")
        dev = self.eng.device
        dur = [u.samples.shape[0] / SR for u in utts]
        order, emb_p, tok_p, cnt_p = [], [], [], []
        batches = plan_batches(dur, self.max_batch_seconds, self.max_batch_utts)
        if batches and len(batches[0]) > 2 * RAMP_UTTS:      # short ramp-up batch, as in extract()
            batches[0:1] = [batches[0][:RAMP_UTTS], batches[0][RAMP_UTTS:]]
        for idx in batches:
            lens = [utts[i].samples.shape[0] for i in idx]
            pcm = self.ring.upload([utts[i].samples[:, 0] for i in idx], int(sum(lens)))
            feat, frames = self.eng.fbank_cmn(pcm, lens)
            emb_p.append(self.eng.campplus(feat, frames))
            mel, mframes = self.eng.logmel(pcm, lens)
            tok, counts = self.eng.s3tok(mel, mframes)
            tok_p.append(tok)
            cnt_p.extend(counts)
            order.extend(index[i] for i in idx)
        gidx = torch.from_numpy(np.asarray(order, dtype=np.int64)).to(dev)
        cnt = torch.from_numpy(np.asarray(cnt_p, dtype=np.int64)).to(dev)
        emb = torch.cat(emb_p) if emb_p else torch.zeros(0, 192, device=dev)
        tok = torch.cat(tok_p) if tok_p else torch.zeros(0, dtype=torch.int32, device=dev)
        full = D.gather_stage1(gidx, emb, n_total)
        counts_full, flat = D.gather_stage2(gidx, cnt, tok, n_total)
        if full is None:
            return None, None
        emb_h = full.cpu().numpy()
        return emb_h, split_tokens(flat.cpu().numpy(), counts_full.cpu().numpy())

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
        return self.tokenize_pcm(*self._to_device_16k(utts, "mean"))

    def tokenize_pcm(self, pcm: torch.Tensor, lens: Sequence[int]):
        """Stage 2 on packed mono 16 kHz PCM already on the device: (packed int32 tokens, token counts)."""
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
