"""Multi-GPU plumbing for stages 1-2: one process per GPU, utterances sharded by duration, no data-path collective.

The only communication is the result exchange at the end: an all_gather of per-utterance results
(embeddings [n_r, 192] fp32, token ids + lengths, utterance indices) and an all_reduce(SUM) of the per-speaker
embedding sums and counts.  Backend is NCCL over NVLink on GPUs; the same code runs over gloo on CPU tensors, which is
how tests/test_distributed.py exercises it with world_size 2.
"""
from __future__ import annotations

import os
from typing import List, Sequence

import torch
import torch.distributed as dist


def init(rank: int, world: int, backend: str = "nccl", port: int = 29531, device: torch.device | None = None):
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", str(port))
    kw = {}
    if backend == "nccl" and device is not None:
        kw["device_id"] = device
    dist.init_process_group(backend, rank=rank, world_size=world, **kw)


def world() -> int:
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def rank() -> int:
    return dist.get_rank() if dist.is_available() and dist.is_initialized() else 0


def all_gather_rows(t: torch.Tensor) -> List[torch.Tensor]:
    """Gather a [n_r, ...] tensor whose leading size differs per rank; returns the per-rank pieces (all ranks)."""
    if world() == 1:
        return [t]
    n = torch.tensor([t.shape[0]], dtype=torch.int64, device=t.device)
    counts = [torch.zeros_like(n) for _ in range(world())]
    dist.all_gather(counts, n)
    counts = [int(c.item()) for c in counts]
    mx = max(max(counts), 1)
    pad = torch.zeros((mx,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    pad[: t.shape[0]] = t
    out = [torch.empty_like(pad) for _ in range(world())]
    dist.all_gather(out, pad)
    return [o[:c] for o, c in zip(out, counts)]


def gather_stage1(index: torch.Tensor, emb: torch.Tensor, n_total: int) -> torch.Tensor:
    """index [n_r] int64 (positions in wav.scp order), emb [n_r, 192] -> [n_total, 192] in wav.scp order."""
    idx = torch.cat(all_gather_rows(index))
    e = torch.cat(all_gather_rows(emb))
    out = torch.zeros((n_total, emb.shape[1]), dtype=emb.dtype, device=emb.device)
    out[idx] = e
    return out


def gather_stage2(index: torch.Tensor, counts: torch.Tensor, tokens: torch.Tensor, n_total: int):
    """index [n_r], counts [n_r] (tokens per utterance, 0 for gated ones), tokens packed [sum counts] (int32; ids are
    < 6561 so they travel as int16).  Returns (counts [n_total], list of per-utterance int tensors)."""
    idx = torch.cat(all_gather_rows(index))
    cnt = torch.cat(all_gather_rows(counts))
    wire = tokens.to(torch.int16).contiguous().view(torch.uint8).view(-1, 2)   # bytes: every backend moves uint8
    tok = torch.cat(all_gather_rows(wire)).contiguous().view(torch.int16).view(-1).to(torch.int32)
    pieces = torch.split(tok, cnt.tolist())
    out: List[torch.Tensor] = [None] * n_total  # type: ignore
    full = torch.zeros(n_total, dtype=counts.dtype, device=counts.device)
    for i, c, p in zip(idx.tolist(), cnt.tolist(), pieces):
        out[i] = p
        full[i] = c
    return full, out


def reduce_speakers(ssum: torch.Tensor, scnt: torch.Tensor):
    """all_reduce(SUM) of per-speaker embedding sums [S, 192] and counts [S] (tools/extract_embedding.py:50-56)."""
    if world() > 1:
        dist.all_reduce(ssum, op=dist.ReduceOp.SUM)
        dist.all_reduce(scnt, op=dist.ReduceOp.SUM)
    return ssum, scnt
