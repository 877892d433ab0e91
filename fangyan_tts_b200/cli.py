"""Drop-in command lines for stages 1-2 (contract B1).

Same flags, inputs and outputs as the reference's tools/extract_embedding.py:61-102 and
tools/extract_speech_token.py:61-96:

    extract_embedding.py    --dir D --onnx_path campplus.onnx            [--num_thread 8]
        reads D/wav.scp, D/utt2spk         writes D/utt2embedding.pt, D/spk2embedding.pt
    extract_speech_token.py --dir D --onnx_path speech_tokenizer_v3.onnx [--num_thread 8]
        reads D/wav.scp                    writes D/utt2speech_token.pt

What differs is inside: utterances are sorted by length, packed into batches and run through the sm_100a kernels;
with more than one visible GPU the script spawns one worker per GPU (or joins an existing torchrun launch), shards
the utterances by duration and gathers the results over NCCL so run.sh needs no edit.  `--num_thread` sizes the
WAV-reader pool (the reference uses it for its per-utterance worker threads).  There is no CPU fallback: without a
GPU or the built library the command exits non-zero and writes nothing, as the reference does on any worker error
(extract_embedding.py:49).

Environment knobs (all optional): FY_GPUS (number of GPUs, default all visible), FY_PRECISION (bf16x3 | bf16),
FY_BATCH_SECONDS, FY_BATCH_UTTS, FY_CAMPPLUS_CONFIG / FY_S3TOK_CONFIG (JSON hyper-parameter overrides).
"""
from __future__ import annotations

import argparse
import collections
import itertools
import json
import logging
import math
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor
from typing import List, Sequence

import numpy as np

from . import kaldi_io, parquet_pack

SR = 16000
log = logging.getLogger("fangyan_b200")


def parse_args(argv, prog):
    ap = argparse.ArgumentParser(prog=prog)
    ap.add_argument("--dir", type=str, required=True)
    ap.add_argument("--onnx_path", type=str, required=True)
    ap.add_argument("--num_thread", type=int, default=8)
    return ap.parse_args(argv)


def scan_durations(keys: Sequence[str], paths: Sequence[str], pool: ThreadPoolExecutor):
    """Header-only pass: (frames, channels, sample_rate) per utterance, so batches can be planned before decoding."""
    return list(pool.map(lambda kp: kaldi_io.wav_header(kp[1], kp[0]), zip(keys, paths)))


def len16k(frames: int, sr: int) -> int:
    """Length after torchaudio Resample(sr -> 16000): ceil(16000 * n / sr) (functional.py:1432)."""
    if sr == SR:
        return frames
    g = math.gcd(sr, SR)
    return int(math.ceil((SR // g) * frames / (sr // g)))


def prefetch(batches, load_one, pool: ThreadPoolExecutor, depth: int = 2):
    """Yield (batch indices, decoded utterances) while the reader pool already works on the next `depth` batches."""
    it = iter(batches)
    pending = collections.deque()
    for b in itertools.islice(it, depth):
        pending.append((b, [pool.submit(load_one, i) for i in b]))
    while pending:
        b, futs = pending.popleft()
        nxt = next(it, None)
        if nxt is not None:
            pending.append((nxt, [pool.submit(load_one, i) for i in nxt]))
        yield b, [f.result() for f in futs]   # a decode error propagates here with the utterance key in its message


def _dist_env():
    """(rank, world, local_rank) when launched under torchrun, else None."""
    if "RANK" in os.environ and "WORLD_SIZE" in os.environ:
        return int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
    return None


def _visible_gpus_fast():
    """Number of GPUs this process will see, WITHOUT importing torch or creating a CUDA context (both cost seconds):
    the driver's /proc listing filtered by CUDA_VISIBLE_DEVICES.  None when that cannot be told cheaply."""
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis is not None:
        ids = [x for x in vis.split(",") if x.strip() != ""]
        return len(ids) if all(x.strip().lstrip("-").isdigit() or x.strip().startswith(("GPU-", "MIG-")) for x in ids) else None
    try:
        return len(os.listdir("/proc/driver/nvidia/gpus")) or None
    except OSError:
        return None


def _n_gpus() -> int:
    import torch
    if not torch.cuda.is_available():
        from ._lib import FyError
        raise FyError("no CUDA device visible: fangyan_b200 has no CPU path (the reference's onnxruntime CPU fallback "
                      "is deliberately not reproduced)")
    n = torch.cuda.device_count()
    want = int(os.environ.get("FY_GPUS", n))
    return max(1, min(n, want))


def _seconds_since_exec() -> float:
    """Seconds since this process was created (interpreter start, imports ...), from /proc; 0.0 when unknown."""
    try:
        with open("/proc/self/stat") as f:
            start_ticks = int(f.read().rsplit(")", 1)[1].split()[19])
        with open("/proc/uptime") as f:
            up = float(f.read().split()[0])
        return max(0.0, up - start_ticks / os.sysconf("SC_CLK_TCK"))
    except (OSError, ValueError, IndexError):
        return 0.0


def _worker(local_rank: int, world: int, stage: int, args, port: int, spawned: bool):
    import torch
    from . import distributed as D
    from .audio_io import read_s16_into, read_wav, scan_wavs
    from .configs import campplus_config, s3_config
    from .engine import Engine, kaldi_num_frames
    from .pipeline import Extractor, Utt, plan_batches, shard_indices, split_tokens
    from .weights import load_state

    env = _dist_env()
    rank = env[0] if env and not spawned else local_rank
    t_entry = time.perf_counter()
    since_exec = _seconds_since_exec()                     # interpreter start + imports, for the time split
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", str(port))
        # the job's only traffic is a few point-to-point sends and two tiny all_reduces: no NVLink-SHARP set-up
        os.environ.setdefault("NCCL_NVLS_ENABLE", "0")
        torch.cuda.set_device(local_rank)
        # no device_id: the communicator is NOT built here (8 ranks: ~10 s) but by the first collective, below
        D.init(rank, world, "nccl", port, None)
    t_start = time.perf_counter()
    t_init = t_start - t_entry
    warm = None
    if world > 1:
        # NCCL builds its communicator at the first collective (1.5 s at 2 ranks, ~10 s at 8): a side thread does that
        # under the weight upload, the header scan AND the extraction loop; it is joined in front of the result gather
        import threading

        def _warm():
            try:
                torch.cuda.set_device(local_rank)
                D.warm_up(torch.device("cuda", local_rank))
            except BaseException:       # a rank whose warm-up failed must not go on to the real exchange: its peers
                import traceback        # would pair its sends with their warm-up receives.  End it; the launcher's
                traceback.print_exc()   # watcher then ends every rank and nothing is written.
                sys.stderr.flush()
                os._exit(1)

        warm = threading.Thread(target=_warm, daemon=True)
        warm.start()
    utt2wav = kaldi_io.read_table(os.path.join(args.dir, "wav.scp"))
    keys, paths = list(utt2wav.keys()), list(utt2wav.values())
    pool = ThreadPoolExecutor(max_workers=max(1, args.num_thread))
    n_io = max(1, args.num_thread)
    # header pass over every file (native threads, GIL released) runs under the engine set-up and the weight upload
    scan_job = pool.submit(lambda: (time.perf_counter(), scan_wavs(keys, paths, n_io), time.perf_counter()))
    eng = Engine(local_rank, os.environ.get("FY_PRECISION", "bf16x3"))
    cfg = campplus_config() if stage == 1 else s3_config()
    # .onnx files are mapped by graph topology (weights.py); .npz / .pt carry our tensor names
    state = {k: np.ascontiguousarray(v, dtype=np.float32)
             for k, v in load_state(args.onnx_path, "campplus" if stage == 1 else "s3tok", cfg).items() if v.dtype.kind == "f"}
    if stage == 1:
        eng.load_campplus(state, cfg)
    else:
        eng.load_s3tok(state, cfg)
    del state

    t_weights = time.perf_counter() - t_start
    t0 = time.perf_counter()
    s0, scan, s1 = scan_job.result()
    t_scan_wait, t_scan = time.perf_counter() - t0, s1 - s0
    n16 = [len16k(int(fr), int(sr)) for fr, sr in zip(scan["frames"], scan["sample_rate"])]
    dur = [n / SR for n in n16]
    mine = shard_indices(dur, rank, world)
    # plain 16 kHz mono PCM16 (what prepare_training_data.py writes): read by the library straight into the pinned ring
    direct = scan["simple"] & (scan["sample_rate"] == SR) & (scan["channels"] == 1)
    if os.environ.get("FY_NATIVE_READER", "1") == "0":
        direct = direct & False
    ex = Extractor(eng, float(os.environ.get("FY_BATCH_SECONDS", "2400")), int(os.environ.get("FY_BATCH_UTTS", "4096")))
    dev = eng.device
    tm = {"decode_wait_s": 0.0, "gpu_events": []}

    def load_one(i):
        x, sr = read_wav(paths[i], keys[i])
        return Utt(x, sr)

    def start(b):
        """Begin loading one batch: the native reader fills a slot of the upload ring in a pool thread (one call per
        batch, GIL released), anything else is decoded per file by the Python reader."""
        if all(direct[i] for i in b):
            lens = [n16[i] for i in b]
            total = int(sum(lens))
            slot, view = ex.ring.acquire(total)
            return b, (slot, total, lens, pool.submit(read_s16_into, scan, b, view, n_io, keys))
        return b, [pool.submit(load_one, i) for i in b]

    def batches_of(todo):
        """(indices, device PCM, lengths) or (indices, decoded utterances, None) per batch, two batches ahead; the time
        this rank's main thread spends blocked on the readers is accumulated in tm (the GPU keeps running the previous
        batch meanwhile: nothing in the loop synchronises)."""
        local_dur = [dur[i] for i in todo]
        plan = [[todo[j] for j in b] for b in plan_batches(local_dur, ex.max_batch_seconds, ex.max_batch_utts)]
        bar = _progress(range(len(plan)), len(plan), rank, f"stage {stage}")
        it = iter(plan)
        pending = collections.deque(start(b) for b in itertools.islice(it, 2))
        for _ in bar:
            b, job = pending.popleft()
            t = time.perf_counter()
            if isinstance(job, tuple):
                slot, total, lens, fut = job
                fut.result()                           # a read error propagates here with the utterance key
                item = (b, ex.ring.submit(slot, total), lens)
            else:
                item = (b, [f.result() for f in job], None)
            tm["decode_wait_s"] += time.perf_counter() - t
            nxt = next(it, None)
            if nxt is not None:
                pending.append(start(nxt))
            a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            yield item
            z.record()
            tm["gpu_events"].append((a, z))

    def to_dev(values, dtype):
        return torch.from_numpy(np.asarray(values, dtype=dtype)).to(dev, non_blocking=True)

    t_loop = time.perf_counter()
    if stage == 1:
        utt2spk = kaldi_io.read_table(os.path.join(args.dir, "utt2spk"))
        missing = [k for k in keys if k not in utt2spk]
        if missing:
            raise KeyError(f"utt2spk lacks {len(missing)} utterances of wav.scp, e.g. {missing[:3]}")
        spk_names = list(dict.fromkeys(utt2spk[k] for k in keys))
        spk_index = {s: j for j, s in enumerate(spk_names)}
        spk_of = np.fromiter((spk_index[utt2spk[k]] for k in keys), dtype=np.int32, count=len(keys))
        for i in mine:
            if kaldi_num_frames(n16[i]) < 3:
                raise ValueError(f"utterance {keys[i]}: {n16[i]} samples at 16 kHz is too short for a speaker embedding")
        ssum = torch.zeros(len(spk_names), 192, device=dev)
        scnt = torch.zeros(len(spk_names), dtype=torch.int32, device=dev)
        order, emb_parts = [], []
        for b, x, lens in batches_of(mine):
            emb = ex.embed_batch(x) if lens is None else ex.embed_pcm(x, lens)
            eng.speaker_sum(emb, to_dev(spk_of[b], np.int32), len(spk_names), ssum, scnt)
            order.extend(b)
            emb_parts.append(emb)
        index = to_dev(order, np.int64)
        emb = torch.cat(emb_parts) if emb_parts else torch.zeros(0, 192, device=dev)
        torch.cuda.synchronize()
        t_loop = time.perf_counter() - t_loop
        t0 = time.perf_counter()
        if warm is not None:
            warm.join()
        t_warm = time.perf_counter() - t0
        full = D.gather_stage1(index, emb, len(keys))
        D.reduce_speakers(ssum, scnt)
        mean = eng.speaker_mean(ssum, scnt)
        torch.cuda.synchronize()
        t_gather = time.perf_counter() - t0
        t0 = time.perf_counter()
        if rank == 0:
            full_h, mean_h = full.cpu().numpy(), mean.cpu().numpy()          # one device-to-host copy each
            # same files as torch.save(dict of lists), written from the arrays (kaldi_io.save_embeddings_atomic)
            kaldi_io.save_embeddings_atomic(keys, full_h, os.path.join(args.dir, "utt2embedding.pt"))
            kaldi_io.save_embeddings_atomic(spk_names, mean_h, os.path.join(args.dir, "spk2embedding.pt"))
            if os.environ.get("FY_SIDECAR", "1") != "0":      # compact copies for the Arrow stage-3 fast path
                parquet_pack.write_embedding_sidecar(os.path.join(args.dir, parquet_pack.EMB_SIDECAR), keys, full_h)
                parquet_pack.write_embedding_sidecar(os.path.join(args.dir, parquet_pack.SPK_SIDECAR), spk_names, mean_h)
        t_save = time.perf_counter() - t0
    else:
        gated = [i for i in mine if n16[i] / SR > 30]          # extract_speech_token.py:34-36: key kept, value []
        for i in gated:
            logging.warning("do not support extract speech token for audio longer than 30s (%s)", keys[i])
        todo = [i for i in mine if not n16[i] / SR > 30 and n16[i] >= 160]
        order, cnt_all, tok_parts = [], [], []
        for b, x, lens in batches_of(todo):
            tok, counts = ex.tokenize_batch(x) if lens is None else ex.tokenize_pcm(x, lens)
            order.extend(b)
            cnt_all.extend(counts)
            tok_parts.append(tok)
        index, counts = to_dev(order, np.int64), to_dev(cnt_all, np.int64)
        toks = torch.cat(tok_parts) if tok_parts else torch.zeros(0, dtype=torch.int32, device=dev)
        torch.cuda.synchronize()
        t_loop = time.perf_counter() - t_loop
        t0 = time.perf_counter()
        if warm is not None:
            warm.join()
        t_warm = time.perf_counter() - t0
        full_cnt, flat = D.gather_stage2(index, counts, toks, len(keys))
        torch.cuda.synchronize()
        t_gather = time.perf_counter() - t0
        t0 = time.perf_counter()
        if rank == 0:
            cnt_h, flat_h = full_cnt.cpu().numpy(), flat.cpu().numpy()       # one device-to-host copy each
            kaldi_io.save_tokens_atomic(keys, flat_h, cnt_h, os.path.join(args.dir, "utt2speech_token.pt"))
            if os.environ.get("FY_SIDECAR", "1") != "0":
                parquet_pack.write_token_sidecar_flat(os.path.join(args.dir, parquet_pack.TOK_SIDECAR), keys, flat_h, cnt_h)
        t_save = time.perf_counter() - t0
    if rank == 0:
        wall = time.perf_counter() - t_start
        gpu_s = sum(a.elapsed_time(z) for a, z in tm["gpu_events"]) / 1e3
        sys.stderr.write(json.dumps({
            "stage": stage, "utts": len(keys), "audio_hours": sum(dur) / 3600.0, "gpus": world, "wall_s": wall,
            "precision": eng.precision, "audio_hours_per_s": sum(dur) / 3600.0 / wall,
            # rank 0's view; decode+stage and GPU overlap inside loop_s (reader pool + upload ring run ahead of the kernels)
            "native_reader_utts": int(direct.sum()),
            # before_worker = interpreter start + imports of this process; process_group = rendezvous (communicator
            # built later by a side thread: communicator_wait is what of it was still exposed in front of the gather)
            "split_s": {"before_worker": since_exec, "process_group": t_init, "communicator_wait": t_warm,
                        "engine_and_weights": t_weights, "header_scan": t_scan, "header_scan_exposed": t_scan_wait,
                        "loop": t_loop,
                        "loop_decode_wait": tm["decode_wait_s"], "loop_pack_upload": ex.ring.pack_seconds,
                        "loop_gpu_busy": gpu_s, "gather": t_gather, "tolist_save": t_save}}) + "\n")
    pool.shutdown()
    if world > 1:
        import torch.distributed as dist
        dist.barrier()                      # rank 0 has every row (and has written the files) before anybody leaves
    if _FAST_EXIT:
        return                              # run_tool ends the process right away: the driver reclaims GPU state
    eng.close()
    if world > 1:
        dist.destroy_process_group()


def _progress(it, total, rank, desc):
    if rank != 0:
        return it
    try:
        from tqdm import tqdm
        return tqdm(it, total=total, desc=desc, unit="batch")
    except ImportError:
        return it


def run_stage(stage: int, argv=None, prog: str = "extract") -> int:
    args = parse_args(argv, prog)
    logging.basicConfig(level=logging.INFO, format="%(asctime)s %(levelname)s %(message)s")
    for name in ("wav.scp",) + (("utt2spk",) if stage == 1 else ()):
        if not os.path.exists(os.path.join(args.dir, name)):
            raise FileNotFoundError(os.path.join(args.dir, name))
    from . import _lib
    _lib.load()                     # fail before any work when the CUDA library is not built
    env = _dist_env()
    if env is not None:             # already under torchrun: one rank per process
        _worker(env[2], env[1], stage, args, int(os.environ.get("MASTER_PORT", "29531")), False)
        return 0
    fast = _visible_gpus_fast()
    world = max(1, min(fast, int(os.environ.get("FY_GPUS", fast)))) if fast else _n_gpus()
    if world == 1:
        _n_gpus()                   # raises when there is no CUDA device at all
        _worker(0, 1, stage, args, 0, True)
        return 0
    return _run_ranks(stage, args, world, prog)


def _run_ranks(stage: int, args, world: int, prog: str) -> int:
    """One worker per GPU without editing run.sh: this process becomes rank 0 and starts ranks 1..N-1 as plain
    subprocesses of the same command line (environment RANK / WORLD_SIZE / LOCAL_RANK / MASTER_*), BEFORE it imports
    torch itself, so the interpreters start, import and create their CUDA contexts side by side.  Only rank 0 writes
    files (atomically, after the gather), so the all-or-nothing contract needs one rule: if any rank dies, everybody
    dies before rank 0 reaches its write."""
    import signal
    import subprocess
    import threading
    port = 29500 + os.getpid() % 2000
    tool = {"extract_embedding.py": "extract_embedding.py", "extract_speech_token.py": "extract_speech_token.py"}.get(prog, prog)
    script = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools", tool)
    cmd = [sys.executable, script, "--dir", args.dir, "--onnx_path", args.onnx_path, "--num_thread", str(args.num_thread)]
    base = dict(os.environ, WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), FY_CHILD_RANK="1")
    kids = [subprocess.Popen(cmd, env=dict(base, RANK=str(r), LOCAL_RANK=str(r))) for r in range(1, world)]
    done = threading.Event()

    def watch():
        while not done.is_set():
            for k in kids:
                rc = k.poll()
                if rc not in (None, 0):
                    sys.stderr.write(f"fangyan_b200: rank process {k.pid} exited with status {rc}; aborting, nothing written\n")
                    for o in kids:
                        if o.poll() is None:
                            o.kill()
                    os._exit(1)             # rank 0 may be blocked inside a collective: no exception can reach it
            done.wait(0.2)

    threading.Thread(target=watch, daemon=True).start()
    os.environ.update(RANK="0", WORLD_SIZE=str(world), LOCAL_RANK="0", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    try:
        _worker(0, world, stage, args, port, False)
    except BaseException:
        done.set()
        for k in kids:
            if k.poll() is None:
                k.send_signal(signal.SIGTERM)
        raise
    done.set()
    rc = 0
    for k in kids:
        rc = rc or k.wait()
    return 1 if rc else 0


_FAST_EXIT = False


def run_tool(main) -> None:
    """Entry of the tools/*.py scripts: run `main`, then leave with os._exit once the outputs are on disk.  Orderly
    teardown of torch + the CUDA context + the NCCL communicator costs 2-3 s per process (profiles/r02_cli_e2e_n1.json:
    process wall - in-process wall), i.e. more than stage 1 of a 100-hour job needs on the GPU; nothing is left to flush
    or release that the operating system and the driver do not reclaim at exit.  FY_FAST_EXIT=0 restores sys.exit.
    Errors still surface the ordinary way (traceback, status 1)."""
    global _FAST_EXIT
    _FAST_EXIT = os.environ.get("FY_FAST_EXIT", "1") != "0"
    rc = main()
    if not _FAST_EXIT:
        sys.exit(rc)
    logging.shutdown()
    sys.stdout.flush()
    sys.stderr.flush()
    os._exit(int(rc or 0))


def main_embedding(argv=None) -> int:
    return run_stage(1, argv, "extract_embedding.py")


def main_speech_token(argv=None) -> int:
    return run_stage(2, argv, "extract_speech_token.py")
