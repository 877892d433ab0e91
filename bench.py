#!/usr/bin/env python3
"""Benchmark of the stage 1+2 extraction hot path (BASELINE.json metric: audio-hours/sec).

A "step" is one pass of stage 1 (fbank+CMN -> CAMPPlus -> speaker mean) and stage 2 (log-mel -> S3 encoder -> FSQ)
over a fixed set of synthetic utterances (default 1,000 utterances of 3-10 s, the size of BASELINE configs[1];
configs[2]'s 1,000 hours is this step repeated).  `value` times the step with PCM already resident in HBM;
`e2e` times it through the public API (fangyan_tts_b200.pipeline.Extractor) from pinned host PCM to host results.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "audio_hours_per_sec_stage1+2"
UNIT = "audio-hours/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--utts", type=int, default=int(os.environ.get("FY_BENCH_UTTS", "1000")))
    ap.add_argument("--precision", default=os.environ.get("FY_PRECISION", "bf16x3"), choices=["bf16", "bf16x3"])
    ap.add_argument("--batch-seconds", type=float, default=float(os.environ.get("FY_BATCH_SECONDS", "2400")))
    ap.add_argument("--cpu-sample", type=int, default=96, help="utterances in the bounded CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-bf16-mode", action="store_true")
    ap.add_argument("--stage3-pack", type=int, default=0, metavar="N",
                    help="row N1 instead of the headline step: pack N synthetic utterances into parquet shards with "
                         "the Arrow-direct writer and with the reference's pandas path (CPU only), one JSON line")
    ap.add_argument("--frontend-sweep", action="store_true",
                    help="BASELINE configs[4] instead of the headline step: front-end kernels only, batch 1-1024 of "
                         "10 s clips against the HBM roofline, with the CPU restatement timed beside them")
    return ap.parse_args()


def workload(n_utts: int, seed: int):
    """Seeded synthetic utterances: a pool of distinct waveforms, cut to U(3, 10) s lengths."""
    from fangyan_tts_b200 import synth
    rng = np.random.default_rng(seed)
    pool = [synth.utterance(rng, 10.0) for _ in range(min(n_utts, 48))]
    utts = []
    for i in range(n_utts):
        n = int(rng.uniform(3.0, 10.0) * 16000)
        base = pool[i % len(pool)]
        off = int(rng.integers(0, len(base) - n + 1))
        utts.append(np.ascontiguousarray(base[off:off + n]))
    return utts


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return d.get("hbm_gbs", 6650.0), d.get("bf16_tflops", 1590.0), d.get("bf16_tflops_sustained", 1400.0), "measured"
    return 6650.0, 1590.0, 1400.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        busy = sorted(sm)[len(sm) // 2:] if sm else []
        return {"sm_mhz": float(np.median(busy)) if busy else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ CPU arm


_CPU_MODELS = []


def cpu_reference_rate(utts, threads: int):
    """The reference's per-utterance path restated on the CPU (oracle/; torch fp32, batch 1, true length), driven the
    way the reference drives it: a ThreadPoolExecutor with one utterance per worker thread and intra-op parallelism 1
    (extract_embedding.py:81-82,100).  Returns (audio-hours/s, audio seconds processed, wall seconds)."""
    import torch
    from concurrent.futures import ThreadPoolExecutor
    from fangyan_tts_b200 import synth_weights
    from oracle import campplus as CP, frontend as F, s3tokenizer as S3
    torch.set_num_threads(1)
    if not _CPU_MODELS:
        cam, s3 = CP.CamPlus(), S3.S3Tokenizer()                 # same random-init tensors as the GPU arm
        cam.load_state_dict({k: torch.from_numpy(v) for k, v in synth_weights.campplus_state().items()}, strict=False)
        s3.load_state_dict({k: torch.from_numpy(v) for k, v in synth_weights.s3tok_state().items()})
        _CPU_MODELS.extend([cam.eval(), s3.eval()])
    cam, s3 = _CPU_MODELS

    def single_job(u):
        x = torch.from_numpy(u.astype(np.float32) / 32768.0)[None]
        CP.embed_one(cam, F.cmn(F.kaldi_fbank(x)))
        if F.stage2_gate(x.shape[1]):
            S3.tokenize_one(s3, F.whisper_log_mel(x))
        return len(u) / 16000.0

    with ThreadPoolExecutor(max_workers=threads) as ex:
        t0 = time.perf_counter()
        secs = sum(ex.map(single_job, utts))
        wall = time.perf_counter() - t0
    return secs / 3600.0 / wall, secs, wall


# ------------------------------------------------------------------------------------------------ GPU arm


def run_ours(args):
    import torch
    import torch.distributed as dist
    from fangyan_tts_b200.engine import Engine
    from fangyan_tts_b200.pipeline import Extractor, Utt, plan_batches
    from fangyan_tts_b200 import synth_weights
    from fangyan_tts_b200.configs import CamPlusConfig, S3Config

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = Engine(local, args.precision)
    cam_cfg, s3_cfg = CamPlusConfig(), S3Config()
    cam_state, s3_state = synth_weights.campplus_state(cam_cfg), synth_weights.s3tok_state(s3_cfg)
    eng.load_campplus(cam_state, cam_cfg)                                # random-init weights of the named architectures
    eng.load_s3tok(s3_state, s3_cfg)

    utts = workload(args.utts, 1986 + 2 + rank)          # weak scaling: every rank owns its own shard
    audio_s = sum(len(u) for u in utts) / 16000.0
    n_spk = 11
    dur = [len(u) / 16000.0 for u in utts]
    batches = plan_batches(dur, args.batch_seconds, 4096)
    dev = eng.device
    resident = []
    for idx in batches:
        pcm = torch.from_numpy(np.concatenate([utts[i] for i in idx])).to(dev)
        spk = torch.tensor([i % n_spk for i in idx], dtype=torch.int32, device=dev)
        resident.append((pcm, [len(utts[i]) for i in idx], spk))
    ssum = torch.zeros(n_spk, 192, device=dev)
    scnt = torch.zeros(n_spk, dtype=torch.int32, device=dev)

    engines = [eng]

    def step_device():
        eng = engines[0]
        ssum.zero_()
        scnt.zero_()
        out = None
        for pcm, lens, spk in resident:
            feat, frames = eng.fbank_cmn(pcm, lens)
            emb = eng.campplus(feat, frames)
            eng.speaker_sum(emb, spk, n_spk, ssum, scnt)
            mel, mframes = eng.logmel(pcm, lens)
            tok, counts = eng.s3tok(mel, mframes)
            out = (emb, tok)
        return eng.speaker_mean(ssum, scnt), out

    ex = Extractor(eng, max_batch_seconds=args.batch_seconds, max_batch_utts=4096)
    host_utts = [Utt(u.reshape(-1, 1), 16000) for u in utts]
    h2d = sum(len(u) for u in utts) * 2
    d2h = [0]

    def step_e2e():
        emb, toks = ex.extract(host_utts)                 # public API: host PCM16 in, host embeddings + token lists out
        d2h[0] = emb.nbytes + sum(len(t) for t in toks) * 4
        return emb, toks

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(steps):
            fn()
        b.record()
        barrier()
        ms = a.elapsed_time(b)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    for _ in range(max(args.warmup, 3)):
        step_device()
    prof = (C.c_double * 10)()
    eng.lib.fy_profile_read(eng.h, prof, 10)          # clear counters
    eng.lib.fy_profile_enable(eng.h, 1)
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ms = timed(step_device, args.steps)
    clk = clocks.stop() if rank == 0 else None
    eng.lib.fy_profile_read(eng.h, prof, 10)
    eng.lib.fy_profile_enable(eng.h, 0)
    p = list(prof)

    step_e2e()                                            # warm the pinned buffers
    t0 = time.perf_counter()
    e2e_steps = max(1, min(args.steps, 3))
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_e2e()
    barrier()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())

    # secondary figure: the single-pass bf16 encoder (BASELINE configs[3]); it does NOT meet the token-parity gate
    bf16_mode = None
    if world == 1 and args.precision == "bf16x3" and not args.no_bf16_mode:
        eng.close()
        e2 = Engine(local, "bf16")
        e2.load_campplus(cam_state, cam_cfg)
        e2.load_s3tok(s3_state, s3_cfg)
        engines[0] = e2
        for _ in range(3):
            step_device()
        ms2 = timed(step_device, max(1, min(args.steps, 3))) / max(1, min(args.steps, 3))
        bf16_mode = {"value": audio_s / 3600.0 / (ms2 / 1e3), "unit": UNIT, "ms_per_step": ms2,
                     "note": "single bf16 pass in the S3 encoder: 96 % token match, fails the 99.9 % parity gate; "
                             "reported for BASELINE configs[3] only"}
    del cam_state, s3_state

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    hbm, tf_burst, tf_sust, which = peaks()
    ms_per_step = ms / args.steps
    value = world * audio_s / 3600.0 / (ms_per_step / 1e3)
    gemm_ms, gemm_flops, gemm_n = p[0], p[3], p[6]
    achieved = gemm_flops / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else 0.0
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16x3 (split-bf16 tensor-core products, fp32 accumulate)" if args.precision == "bf16x3" else "bf16",
        "data": "synthetic",
        "config": {"workload": f"stage1+2 over {args.utts} synthetic 16 kHz utterances of 3-10 s per GPU "
                               f"({audio_s / 3600.0:.3f} audio-hours; BASELINE configs[1] size, both stages)",
                   "batches_per_step": len(batches), "batch_seconds": args.batch_seconds,
                   "weights": "random-init CAMPPlus + S3 tokenizer v3 (12 layers)", "precision": args.precision,
                   "l2": "per-step inputs and workspaces (GBs) exceed the 126 MB L2", "parallelism": f"dp{world}"},
        "e2e": {"value": world * audio_s / 3600.0 / (e2e_s / e2e_steps), "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h[0], "steps": e2e_steps},
        "gpu_launches": int(p[9]),
        "roofline": {"bound": "tensor", "kernel": "gemm_kernel<BN, OP, NCTA> (tcgen05; all launches: S3 encoder "
                                                  "bf16x3 CTA-pair tiles + CAMPPlus fp16 tiles)",
                     "achieved": achieved, "peak": tf_sust, "unit": "TFLOP/s", "frac": achieved / tf_sust,
                     # ncu dram__bytes_read+write per launch, mean of the four S3 layer GEMMs at 16.4 k rows
                     # (profiles/r01_ncu_gemm_pair_bf16x3.md); the bench batches are 3.7x larger
                     "traffic": 4.11e8, "traffic_unit": "bytes/launch at the profiled size",
                     "peak_source": f"{which} sustained cuBLAS bf16",
                     "executed_tflops": achieved * (3.0 if args.precision == "bf16x3" else 1.0),
                     "executed_frac": achieved * (3.0 if args.precision == "bf16x3" else 1.0) / tf_sust,
                     "note": "achieved counts each MAC once; split-bf16 executes three tensor-core products per MAC "
                             "(ceiling 1/3), executed_frac counts them (CAMPPlus launches, 1 product, are <4% of the FLOPs)",
                     "launches": int(gemm_n), "avg_launch_ms": gemm_ms / max(gemm_n, 1),
                     "gemm_share_of_step": gemm_ms / ms, "attention_ms_per_step": p[1] / args.steps},
        "clocks": clk,
    }
    if bf16_mode:
        line["bf16_mode"] = bf16_mode
    if not args.no_cpu_baseline and world == 1:
        threads = len(os.sched_getaffinity(0))
        sample = utts[:args.cpu_sample]
        v, secs, wall = cpu_reference_rate(sample, threads)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": f"first {len(sample)} utterances of the workload ({secs:.1f} audio-s, {wall:.1f} s wall), "
                                          "oracle = torch-cpu restatement (onnxruntime unavailable)"}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = len(os.sched_getaffinity(0))
    utts = workload(args.utts, 1986 + 2)
    sample = utts[:args.cpu_sample]
    for _ in range(min(args.warmup, 1)):
        cpu_reference_rate(sample[:1], threads)
    rates, wall_tot = [], 0.0
    for _ in range(max(1, min(args.steps, 3))):
        v, secs, wall = cpu_reference_rate(sample, threads)
        rates.append(v)
        wall_tot += wall
    v = float(np.median(rates))
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * wall_tot / len(rates), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"stage1+2, bounded sample: first {len(sample)} of {args.utts} synthetic utterances, "
                               "batch 1 at true length as the reference does"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"{len(sample)} utterances per step; torch-cpu restatement of the reference path "
                                   "(onnxruntime / whisper / the .onnx files are not installable offline)"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def run_frontend_sweep(args):
    """One JSON line per batch size (tools/bench_frontend.py does the GPU side) and one `cpu_baseline` line: the oracle
    front-ends (the reference's torchaudio / whisper / matcha calls restated) on 16 ten-second clips with every host
    thread torch will use."""
    import torch
    sys.path.insert(0, str(Path(__file__).resolve().parent / "tools"))
    import bench_frontend
    bench_frontend.main()
    from fangyan_tts_b200 import synth
    from oracle import frontend as F
    rng = np.random.default_rng(1986 + 4)
    x16 = torch.from_numpy(synth.utterance(rng, 10.0).astype(np.float32) / 32768.0)[None]
    x24 = F.resample(x16, 16000, 24000)
    n, res = 16, {"config": "front-end sweep", "cpu_baseline": {"kind": "port", "cores": torch.get_num_threads(),
                                                                   "sample": "16 x 10 s clips, batch 1 each"}}
    for name, fn, x in (("fbank_cmn", lambda v: F.cmn(F.kaldi_fbank(v)), x16), ("logmel", F.whisper_log_mel, x16),
                        ("matcha_mel_24k", F.matcha_mel, x24)):
        fn(x)
        t0 = time.perf_counter()
        for _ in range(n):
            fn(x)
        dt = time.perf_counter() - t0
        res["cpu_baseline"][name] = {"ms_per_clip": 1e3 * dt / n, "audio_hours_per_s": n * 10.0 / 3600.0 / dt}
    print(json.dumps(res), flush=True)


def run_stage3_pack(args):
    """Row N1: wall time from the stage-1/2 outputs on disk to parquet shards, reference path (torch.load of the list
    pickles + pandas frames of Python lists, oracle/stage3.py) against parquet_pack (sidecars + Arrow arrays)."""
    import tempfile
    import wave
    from fangyan_tts_b200 import kaldi_io, parquet_pack
    from oracle import stage3
    n = args.stage3_pack
    rng = np.random.default_rng(7)
    with tempfile.TemporaryDirectory() as tmp:
        src = Path(tmp) / "src"
        src.mkdir()
        keys = [f"utt{i:07d}" for i in range(n)]
        wavp = src / "clip.wav"                               # one small clip shared by every row: IO is not the subject
        with wave.open(str(wavp), "wb") as w:
            w.setnchannels(1); w.setsampwidth(2); w.setframerate(16000)
            w.writeframes(rng.integers(-3000, 3000, 16000, dtype=np.int16).tobytes())
        utt2spk = {k: f"spk{i % 200:03d}" for i, k in enumerate(keys)}
        with open(src / "wav.scp", "w") as f1, open(src / "text", "w") as f2, open(src / "utt2spk", "w") as f3:
            for k in keys:
                f1.write(f"{k} {wavp}\n"); f2.write(f"{k} some text of {k}\n"); f3.write(f"{k} {utt2spk[k]}\n")
        emb = rng.standard_normal((n, 192)).astype(np.float32)
        spk_names = sorted(set(utt2spk.values()))
        semb = rng.standard_normal((len(spk_names), 192)).astype(np.float32)
        toks = [rng.integers(0, 6561, int(rng.integers(75, 250))).astype(np.int32) for _ in range(n)]
        kaldi_io.save_atomic(kaldi_io.embeddings_to_dict(keys, emb), src / "utt2embedding.pt")
        kaldi_io.save_atomic(kaldi_io.speaker_dict(spk_names, semb), src / "spk2embedding.pt")
        kaldi_io.save_atomic({k: t.tolist() for k, t in zip(keys, toks)}, src / "utt2speech_token.pt")
        parquet_pack.write_embedding_sidecar(src / parquet_pack.EMB_SIDECAR, keys, emb)
        parquet_pack.write_embedding_sidecar(src / parquet_pack.SPK_SIDECAR, spk_names, semb)
        parquet_pack.write_token_sidecar(src / parquet_pack.TOK_SIDECAR, keys, toks)
        t0 = time.perf_counter()
        parquet_pack.make_parquet_list(str(src), str(Path(tmp) / "ours"), 1000, 1)
        ours = time.perf_counter() - t0
        t0 = time.perf_counter()
        utt2wav = kaldi_io.read_table(str(src / "wav.scp"))
        utt2text = {k: f"some text of {k}" for k in keys}
        e, s, k = stage3.load_pt(str(src))
        (Path(tmp) / "ref").mkdir()
        for i, j in enumerate(range(0, n, 1000)):
            stage3.job(keys[j:j + 1000], str(Path(tmp) / "ref" / f"parquet_{i:09d}.tar"), utt2wav, utt2text, utt2spk, e, s, k)
        ref = time.perf_counter() - t0
    print(json.dumps({"metric": "stage3_pack_utts_per_sec", "unit": "utterances/s", "utts": n, "workers": 1,
                      "value": n / ours, "seconds": ours,
                      "cpu_baseline": {"kind": "port", "value": n / ref, "seconds": ref, "cores": 1,
                                       "sample": f"{n} synthetic utterances, 1,000 per shard"}}), flush=True)


if __name__ == "__main__":
    a = parse_args()
    if a.stage3_pack:
        run_stage3_pack(a)
    elif a.frontend_sweep:
        run_frontend_sweep(a)
    elif a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
