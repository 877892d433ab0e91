"""world_size-2 gloo test of the multi-GPU plumbing: duration sharding, the result gather and the
per-speaker reduction, with CPU tensors standing in for per-rank GPU results."""
import os

import numpy as np
import torch
import torch.multiprocessing as mp

from fangyan_tts_b200 import distributed as D
from fangyan_tts_b200.pipeline import shard_indices

N, S = 53, 4


def _fake_results(n):
    rng = np.random.default_rng(7)
    dur = rng.uniform(1.0, 12.0, n)
    emb = torch.from_numpy(rng.standard_normal((n, 192)).astype(np.float32))
    counts = torch.tensor([0 if i % 9 == 4 else int(d * 25) for i, d in enumerate(dur)], dtype=torch.int64)
    toks = [torch.from_numpy(rng.integers(0, 6561, int(c)).astype(np.int32)) for c in counts]
    spk = torch.arange(n) % S
    return dur.tolist(), emb, counts, toks, spk


def _worker(rank, world, port, out_path):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    D.init(rank, world, backend="gloo", port=port)
    # communicator + peer connections from a side thread while this thread keeps working, as the CLIs do
    import threading
    warm = threading.Thread(target=D.warm_up, args=(torch.device("cpu"),), daemon=True)
    warm.start()
    busy = sum(float((torch.randn(64, 64) @ torch.randn(64, 64)).sum()) for _ in range(50))
    warm.join(timeout=120)
    assert not warm.is_alive() and busy == busy
    dur, emb, counts, toks, spk = _fake_results(N)
    mine = shard_indices(dur, rank, world)
    idx = torch.tensor(mine, dtype=torch.int64)
    full = D.gather_stage1(idx, emb[idx], N)
    cnt, flat = D.gather_stage2(idx, counts[idx], torch.cat([toks[i] for i in mine]), N)
    ssum = torch.zeros(S, 192).index_add_(0, spk[idx], emb[idx])
    scnt = torch.zeros(S, dtype=torch.int32).index_add_(0, spk[idx], torch.ones(len(mine), dtype=torch.int32))
    D.reduce_speakers(ssum, scnt)
    local = torch.stack([D.checksum(idx, emb[idx]), D.checksum_ragged(idx, counts[idx], torch.cat([toks[i] for i in mine]))])
    torch.distributed.all_reduce(local)
    if rank == 0:
        assert full.shape == (N, 192) and cnt.shape == (N,)
        everyone = torch.arange(N)
        assert int(local[0]) == int(D.checksum(everyone, full))          # nothing dropped, duplicated or misplaced
        assert int(local[1]) == int(D.checksum_ragged(everyone, cnt, flat))
        torch.save({"emb": full, "cnt": cnt, "tok": list(torch.split(flat, cnt.tolist())), "ssum": ssum, "scnt": scnt},
                   out_path)
    else:
        assert full is None and cnt is None and flat is None             # only rank 0 receives
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()


def test_gather_and_speaker_reduce_world2(tmp_path):
    out = tmp_path / "r0.pt"
    mp.spawn(_worker, args=(2, 29731 + os.getpid() % 200, str(out)), nprocs=2, join=True)
    got = torch.load(out)
    dur, emb, counts, toks, spk = _fake_results(N)
    assert torch.equal(got["emb"], emb)                      # wav.scp order restored, values untouched
    assert torch.equal(got["cnt"], counts)
    for i in range(N):
        assert torch.equal(got["tok"][i], toks[i])           # ragged token lists bit-exact (int16 on the wire)
    ref_sum = torch.zeros(S, 192).index_add_(0, spk, emb)
    assert torch.allclose(got["ssum"], ref_sum, atol=1e-5)
    assert got["scnt"].tolist() == [int((spk == s).sum()) for s in range(S)]


def test_single_process_helpers_are_identity():
    t = torch.arange(6).view(3, 2)
    assert D.world() == 1 and D.rank() == 0
    assert torch.equal(D.gather_rows([t])[0], t) and D.row_counts(3, t.device) == [3]
    idx = torch.tensor([2, 0, 1])
    full, flat = D.gather_stage2(idx, torch.tensor([1, 2, 3]), torch.tensor([9, 5, 6, 1, 2, 3], dtype=torch.int32), 4)
    assert full.tolist() == [2, 3, 1, 0] and flat.tolist() == [5, 6, 1, 2, 3, 9]   # wav.scp order, ragged
