"""Multi-GPU plumbing for stages 1-2: one process per GPU, utterances sharded by duration, no data-path collective.

The only communication is the result exchange at the end: a gather-v TO RANK 0 of per-utterance
results (utterance indices, embeddings [n_r, 192] fp32, token counts, token ids as int16) and an all_reduce(SUM) of the
per-speaker embedding sums and counts (tools/extract_embedding.py:45-58 builds the same two dicts in one process).
Only rank 0 writes files, so only rank 0 receives: every other rank sends its rows exactly once, unpadded
(point-to-point sends posted as one batch; NCCL runs them as one grouped kernel over NVLink).  Backend is NCCL on
GPUs; the same code runs over gloo on CPU tensors, which is how tests/test_distributed_cpu.py exercises it with
world_size 2.
"""
from __future__ import annotations

import os
from typing import List, Optional, Tuple

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


def _as_bytes(t: torch.Tensor) -> torch.Tensor:
    """[n, ...] of any dtype -> [n, row_bytes] uint8 view (NCCL has no int16; every backend moves uint8)."""
    t = t.contiguous()
    row = t.element_size()
    for s in t.shape[1:]:
        row *= s
    return t.view(torch.uint8).view(t.shape[0], row)


def row_counts(n_rows: int, device: torch.device) -> List[int]:
    """Rows every rank holds (one tiny all_gather; the only step all ranks take part in symmetrically)."""
    if world() == 1:
        return [n_rows]
    n = torch.tensor([n_rows], dtype=torch.int64, device=device)
    out = torch.empty(world(), dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(out, n)
    return [int(c) for c in out.tolist()]


def gather_rows(tensors: List[torch.Tensor], counts: Optional[List[List[int]]] = None, dst: int = 0):
    """Gather-v of several [n_r, ...] tensors (leading sizes differ per rank and per tensor) to rank `dst`.

    Returns, on `dst`, the list of concatenated tensors in rank order; on every other rank None.  `counts[j]` are the
    per-rank leading sizes of tensors[j] (computed with row_counts when omitted).  Nothing is padded and nothing is
    sent to a rank that does not need it: one send per (rank, tensor), all posted together."""
    w, r = world(), rank()
    if w == 1:
        return list(tensors)
    if counts is None:
        counts = [row_counts(t.shape[0], t.device) for t in tensors]
    ops, outs, views = [], [], []
    for t, cnt in zip(tensors, counts):
        b = _as_bytes(t)
        if r == dst:
            out = torch.empty((sum(cnt),) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
            ob = _as_bytes(out)
            o = 0
            for src, c in enumerate(cnt):
                if src == dst:
                    ob[o:o + c].copy_(b)
                elif c:
                    ops.append(dist.P2POp(dist.irecv, ob[o:o + c], src))
                o += c
            outs.append(out)
        elif b.shape[0]:
            views.append(b)
            ops.append(dist.P2POp(dist.isend, b, dst))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    return outs if r == dst else None


def warm_up(device: torch.device) -> None:
    """Build everything the result exchange will need, ahead of time: the communicator (first collective) and the
    point-to-point connections of every rank to rank 0 (first send / receive per peer; ~1.4 s at 8 ranks when it happens
    in front of the real gather).  The CLIs call this from a side thread under the extraction loop."""
    if world() == 1:
        return
    t = torch.zeros(1, device=device)
    dist.all_reduce(t)
    gather_rows([torch.full((1, 8), rank(), dtype=torch.uint8, device=device)], [[1] * world()])
    if device.type == "cuda":
        torch.cuda.synchronize(device)


def gather_stage1(index: torch.Tensor, emb: torch.Tensor, n_total: int,
                  rows: Optional[List[int]] = None) -> Optional[torch.Tensor]:
    """index [n_r] int64 (positions in wav.scp order), emb [n_r, 192] -> on rank 0 [n_total, 192] in wav.scp order;
    None on the other ranks.  `rows` = utterances per rank when the caller already knows them (the sharding is
    deterministic), which saves the count exchange and its host synchronisation."""
    cnt = rows if rows is not None else row_counts(index.shape[0], index.device)
    got = gather_rows([index, emb], [cnt, cnt])
    if got is None:
        return None
    idx, e = got
    out = torch.zeros((n_total, emb.shape[1]), dtype=emb.dtype, device=emb.device)
    out[idx] = e
    return out


def gather_stage2(index: torch.Tensor, counts: torch.Tensor, tokens: torch.Tensor, n_total: int,
                  rows: Optional[List[int]] = None,
                  tok_rows: Optional[List[int]] = None) -> Tuple[Optional[torch.Tensor], Optional[torch.Tensor]]:
    """index [n_r], counts [n_r] (tokens per utterance), tokens packed [sum counts] int32 in the order of `index`
    (ids are < 6561 so they travel as int16).  On rank 0 returns (counts [n_total] int64 -- 0 for utterances nobody
    reported, e.g. the > 30 s ones --, tokens packed int32 in wav.scp order); (None, None) elsewhere.  The ragged
    re-ordering is one scatter on the device: no per-utterance host loop.  rows / tok_rows: per-rank utterance and token
    totals when known in advance (token counts follow from the lengths), see gather_stage1."""
    rows = rows if rows is not None else row_counts(index.shape[0], index.device)
    toks = tok_rows if tok_rows is not None else row_counts(tokens.shape[0], tokens.device)
    got = gather_rows([index, counts.to(torch.int64), tokens.to(torch.int16)], [rows, rows, toks])
    if got is None:
        return None, None
    idx, cnt, tok = got
    full = torch.zeros(n_total, dtype=torch.int64, device=index.device)
    full[idx] = cnt
    dst_off = torch.cumsum(full, 0) - full                     # where each utterance starts in wav.scp order
    src_off = torch.cumsum(cnt, 0) - cnt                       # ... and in arrival order
    shift = torch.repeat_interleave(dst_off[idx] - src_off, cnt)
    pos = torch.arange(tok.shape[0], device=tok.device) + shift
    out = torch.empty(tok.shape[0], dtype=torch.int32, device=tok.device)
    out[pos] = tok.to(torch.int32)
    return full, out


def reduce_speakers(ssum: torch.Tensor, scnt: torch.Tensor):
    """all_reduce(SUM) of per-speaker embedding sums [S, 192] and counts [S] (tools/extract_embedding.py:50-56)."""
    if world() > 1:
        dist.all_reduce(ssum, op=dist.ReduceOp.SUM)
        dist.all_reduce(scnt, op=dist.ReduceOp.SUM)
    return ssum, scnt


def checksum(index: torch.Tensor, rows: torch.Tensor) -> torch.Tensor:
    """Order-independent int64 fingerprint of (index, row bits): sum over rows of (index + 1) * (sum of the row's
    32-bit patterns).  Ranks all_reduce(SUM) their local value; it must equal the fingerprint of the gathered array, so
    a dropped, duplicated or misplaced row shows (used by bench.py and the N-GPU tests)."""
    bits = rows.contiguous().view(torch.int32).to(torch.int64).view(rows.shape[0], -1).sum(1)
    return ((index.to(torch.int64) + 1) * bits).sum()


def checksum_ragged(index: torch.Tensor, counts: torch.Tensor, tokens: torch.Tensor) -> torch.Tensor:
    """The same fingerprint for ragged token rows: sum over tokens of (utterance index + 1) * (id + 1)."""
    w = torch.repeat_interleave(index.to(torch.int64) + 1, counts.to(torch.int64))
    return (w * (tokens.to(torch.int64) + 1)).sum()
