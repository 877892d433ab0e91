#!/usr/bin/env python3
"""Benchmark of the stage 1+2 extraction hot path (BASELINE.json metric: audio-hours/sec).

One JOB is `--utts` synthetic utterances of 3-10 s per GPU (default 1,000 = BASELINE configs[1] size; configs[2]'s
1,000 hours is this job repeated), generated identically on every rank and sharded by duration with
pipeline.shard_assignment, as the drop-in CLIs do.  A "step" is one pass of stage 1 (fbank+CMN -> CAMPPlus -> speaker
sums) and stage 2 (log-mel -> S3 encoder -> FSQ) over this rank's shard, FOLLOWED BY the job's only communication:
the gather-v of embeddings / token ids to rank 0 and the all_reduce of the speaker sums (distributed.py), so `value`
at N GPUs contains the exchange.  `value` times the step with PCM already resident in HBM; `e2e` times it through the
public API (pipeline.Extractor.extract) from host PCM16 through pinned uploads to host results on rank 0; `parity`
compares the e2e outputs of the benchmarked job with the CPU oracle on a sample; `gather` times the exchange alone.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python bench.py --cli-e2e 40000        # drop-in CLIs, WAV files -> .pt files (SURVEY 8(d) regions ii, iii)
    python bench.py --long-stress          # BASELINE configs[3]
    python bench.py --frontend-sweep       # BASELINE configs[4]
    python bench.py --stage3-pack 20000    # row N1
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
N_SPK = 11              # the dialect data has ~11 synthetic speakers (generate_dialect_dataset.py:86-146)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--utts", type=int, default=int(os.environ.get("FY_BENCH_UTTS", "1000")), help="utterances per GPU")
    ap.add_argument("--precision", default=os.environ.get("FY_PRECISION", "bf16x3"), choices=["bf16", "bf16x3"])
    ap.add_argument("--batch-seconds", type=float, default=float(os.environ.get("FY_BATCH_SECONDS", "2400")))
    ap.add_argument("--cpu-sample", type=int, default=96, help="utterances in the bounded CPU-baseline / parity sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-bf16-mode", action="store_true")
    ap.add_argument("--no-frontend", action="store_true", help="skip the front-end figures at batch 1,024")
    ap.add_argument("--stage3-pack", type=int, default=0, metavar="N",
                    help="row N1 instead of the headline step: pack N synthetic utterances into parquet shards with "
                         "the Arrow-direct writer and with the reference's pandas path (CPU only), one JSON line")
    ap.add_argument("--frontend-sweep", action="store_true",
                    help="BASELINE configs[4] instead of the headline step: front-end kernels only, batch 1-1024 of "
                         "10 s clips against the HBM roofline, with the CPU restatement timed beside them")
    ap.add_argument("--cli-e2e", type=int, default=0, metavar="N",
                    help="time the drop-in CLIs over N utterances: WAV files on tmpfs -> the three .pt files")
    ap.add_argument("--cli-files", type=int, default=4000, help="distinct WAV files behind the --cli-e2e keys")
    ap.add_argument("--serialise", type=int, default=0, metavar="N",
                    help="SURVEY 8(d) region (iii) alone, CPU only: N result rows (192-float embeddings, 75-250 token ids) -> "
                         "the three .pt files, array-direct writers against tolist() + torch.save")
    ap.add_argument("--long-stress", action="store_true",
                    help="BASELINE configs[3]: 256 clips of 1-30 s (two at exactly 30.0 s), 12 layers, bf16x3 and bf16")
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


_CPU_MODELS = {}


def _cpu_models(n_layer=None):
    import torch
    from fangyan_tts_b200 import synth_weights
    from fangyan_tts_b200.configs import S3Config
    from oracle import campplus as CP, s3tokenizer as S3
    if n_layer not in _CPU_MODELS:
        cfg = S3Config() if n_layer is None else S3Config(n_layer=n_layer)
        cam, s3 = CP.CamPlus(), S3.S3Tokenizer(S3.S3Config(n_layer=cfg.n_layer))   # same random-init tensors as the GPU arm
        cam.load_state_dict({k: torch.from_numpy(v) for k, v in synth_weights.campplus_state().items()}, strict=False)
        s3.load_state_dict({k: torch.from_numpy(v) for k, v in synth_weights.s3tok_state(cfg).items()})
        _CPU_MODELS[n_layer] = (cam.eval(), s3.eval())
    return _CPU_MODELS[n_layer]


def literal_reference_available():
    """SURVEY.md 8(c) runtime upgrade path: the reference's own stack (onnx + onnxruntime + openai-whisper).  None of
    them is installable offline; when a box has them the CPU arm exports the random-init networks to ONNX and drives
    onnxruntime's CPUExecutionProvider the way single_job does."""
    try:
        import onnx  # noqa: F401
        import onnxruntime  # noqa: F401
        import whisper  # noqa: F401
        return True
    except Exception:
        return False


class _OrtSessions:
    """The literal path (only when literal_reference_available()): torch.onnx.export of the oracle modules with a
    dynamic time axis, then InferenceSession(..., providers=["CPUExecutionProvider"]) with intra/inter-op threads 1
    (tools/extract_embedding.py:79-94, tools/extract_speech_token.py:75-90)."""

    def __init__(self, cam, s3, tmp):
        import onnxruntime
        import torch

        class S3Ids(torch.nn.Module):
            def __init__(self, m):
                super().__init__()
                self.m = m

            def forward(self, feat, feat_len):
                return self.m.quantize(feat, feat_len.long())[0]

        cam_p, s3_p = os.path.join(tmp, "campplus.onnx"), os.path.join(tmp, "speech_tokenizer_v3.onnx")
        torch.onnx.export(cam, (torch.zeros(1, 200, 80),), cam_p, input_names=["input"], output_names=["output"],
                          dynamic_axes={"input": {1: "T"}}, opset_version=17, dynamo=False)
        torch.onnx.export(S3Ids(s3), (torch.zeros(1, 128, 200), torch.tensor([200], dtype=torch.int32)), s3_p,
                          input_names=["feat", "feat_len"], output_names=["ids"],
                          dynamic_axes={"feat": {2: "T"}, "ids": {1: "L"}}, opset_version=17, dynamo=False)
        opt = onnxruntime.SessionOptions()
        opt.graph_optimization_level = onnxruntime.GraphOptimizationLevel.ORT_ENABLE_ALL
        opt.intra_op_num_threads = 1
        opt.inter_op_num_threads = 1
        self.cam = onnxruntime.InferenceSession(cam_p, sess_options=opt, providers=["CPUExecutionProvider"])
        self.s3 = onnxruntime.InferenceSession(s3_p, sess_options=opt, providers=["CPUExecutionProvider"])


def cpu_reference_run(utts, threads: int, n_layer=None, want_outputs: bool = False):
    """The reference's per-utterance path on the CPU, driven the way the reference drives it: a ThreadPoolExecutor with
    one utterance per worker thread and intra-op parallelism 1 (extract_embedding.py:81-82,100).  `kind` is "port" (the
    torch-cpu restatement under oracle/, batch 1, true length) unless onnxruntime + whisper exist on the box.
    Returns {rate, audio_s, wall, kind, out}; out[i] = (embedding [192], ids list, fsq_pre [L, 8]) when asked for."""
    import torch
    from concurrent.futures import ThreadPoolExecutor
    from oracle import campplus as CP, frontend as F, s3tokenizer as S3
    torch.set_num_threads(1)
    cam, s3 = _cpu_models(n_layer)
    ort, kind = None, "port"
    if literal_reference_available() and not want_outputs:
        try:
            import tempfile
            ort = _CPU_MODELS.get("ort") or _OrtSessions(cam, s3, tempfile.mkdtemp(prefix="fy_ort_"))
            _CPU_MODELS["ort"] = ort
            kind = "reference"
        except Exception as e:                                  # an exporter / ORT version that cannot take the modules
            sys.stderr.write(f"literal onnxruntime path unavailable ({type(e).__name__}: {e}); using the port\n")
            ort = None

    def single_job(u):
        x = torch.from_numpy(u.astype(np.float32) / 32768.0)[None]
        feat = F.cmn(F.kaldi_fbank(x))
        res = None
        with torch.no_grad():
            if ort is not None:
                import whisper
                ort.cam.run(None, {"input": feat.unsqueeze(0).numpy()})
                if F.stage2_gate(x.shape[1]):
                    mel = whisper.log_mel_spectrogram(x, n_mels=128)
                    ort.s3.run(None, {"feat": mel.numpy(), "feat_len": np.array([mel.shape[2]], dtype=np.int32)})
            else:
                emb = CP.embed_one(cam, feat)
                ids, pre = [], None
                if F.stage2_gate(x.shape[1]):
                    mel = F.whisper_log_mel(x)
                    if want_outputs:
                        hid, ln = s3.encode(mel, torch.tensor([mel.shape[2]]))
                        pre = s3.fsq_pre(hid)[0, : int(ln[0])]
                        h = pre.round() + 1.0
                        pw = 3.0 ** torch.arange(8, dtype=torch.float32)
                        ids = (h * pw).sum(-1).int().tolist()
                    else:
                        ids = S3.tokenize_one(s3, mel)
                if want_outputs:
                    res = (emb.numpy(), ids, None if pre is None else pre.numpy())
        return len(u) / 16000.0, res

    with ThreadPoolExecutor(max_workers=threads) as ex:
        t0 = time.perf_counter()
        got = list(ex.map(single_job, utts))
        wall = time.perf_counter() - t0
    secs = sum(g[0] for g in got)
    return {"rate": secs / 3600.0 / wall, "audio_s": secs, "wall": wall, "kind": kind, "out": [g[1] for g in got]}


def parity_report(ref_out, emb, toks):
    """north_star gates on the benchmarked job: embeddings cosine >= 0.9999, token ids >= 99.9 % equal with every
    mismatch within 1e-3 of an FSQ rounding boundary, token-sequence lengths bit-exact.  ref_out from
    cpu_reference_run(want_outputs=True) for utterances 0..n-1 of the job; emb / toks = the GPU arm's results."""
    cos_min, n_tok, n_bad, off_boundary, worst, len_exact = 1.0, 0, 0, 0, 0.0, True
    for i, (e_ref, ids, pre) in enumerate(ref_out):
        a, b = np.asarray(emb[i], np.float64), np.asarray(e_ref, np.float64)
        cos_min = min(cos_min, float(a @ b / (np.linalg.norm(a) * np.linalg.norm(b) + 1e-30)))
        got = list(toks[i])
        if len(got) != len(ids):
            len_exact = False
            continue
        n_tok += len(ids)
        for j, (x, y) in enumerate(zip(got, ids)):
            if x != y:
                n_bad += 1
                d = float(np.min(np.abs(np.abs(pre[j]) - 0.5)))      # distance of the closest FSQ dim to +-0.5
                worst = max(worst, d)
                off_boundary += d >= 1e-3
    match = 1.0 - n_bad / max(n_tok, 1)
    ok = cos_min >= 0.9999 and match >= 0.999 and off_boundary == 0 and len_exact
    return {"utts": len(ref_out), "tokens": n_tok, "token_match": match, "mismatches": n_bad,
            "off_boundary_mismatches": int(off_boundary), "worst_mismatch_boundary_distance": worst,
            "cos_min": cos_min, "len_exact": len_exact, "pass": bool(ok),
            "oracle": "torch-cpu restatement (oracle/), batch 1 at true length; parity unpinned for the two networks",
            "gates": "cos >= 0.9999, token match >= 0.999, every mismatch < 1e-3 from an FSQ boundary, lengths exact"}


# ------------------------------------------------------------------------------------------------ GPU arm


def frontend_figures(eng, hbm, batch=1024):
    """BASELINE configs[4] at its largest batch, inside the driver-run line: fbank+CMN and log-mel over `batch` 10 s
    clips (inputs rotated through > L2), achieved algorithmic GB/s against the measured HBM peak."""
    import torch
    from fangyan_tts_b200 import synth
    base = synth.utterance(np.random.default_rng(1986 + 4), 10.0)
    sets = [torch.from_numpy(np.tile(np.roll(base, 37 * k), batch)).cuda() for k in range(5)]     # 5 x 328 MB
    lens, secs, out = [len(base)] * batch, batch * 10.0, {"batch": batch, "clip_s": 10.0, "l2": "inputs rotate through 1.6 GB"}
    for name, fn, bps in (("fbank_cmn", eng.fbank_cmn, 64000.0), ("logmel", eng.logmel, 83200.0)):
        for k in range(3):
            fn(sets[k % 5], lens)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        for k in range(10):
            fn(sets[k % 5], lens)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 10
        gbs = bps * secs / (ms * 1e-3) / 1e9
        out[name] = {"ms": ms, "alg_bytes_per_audio_s": bps, "achieved_GBps": gbs, "frac_of_hbm_peak": gbs / hbm,
                     "audio_hours_per_s": secs / 3600.0 / (ms * 1e-3)}
    # row A2: 24 kHz -> 16 kHz polyphase resampler on the same number of 10 s clips (fp32 in, fp32 out: 160,000 B/audio-s)
    x24 = torch.randn(batch * 240000, device="cuda") * 0.1
    l24 = [240000] * batch
    for _ in range(3):
        eng.resample(x24, l24, 24000, 16000)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(5):
        eng.resample(x24, l24, 24000, 16000)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 5
    gbs = 160000.0 * secs / (ms * 1e-3) / 1e9
    out["resample_24k_16k"] = {"ms": ms, "alg_bytes_per_audio_s": 160000.0, "achieved_GBps": gbs, "frac_of_hbm_peak": gbs / hbm,
                               "audio_hours_per_s": secs / 3600.0 / (ms * 1e-3), "l2": "1 GB in + 0.66 GB out per call"}
    out["bound_note"] = ("fbank / log-mel are instruction bound, not HBM bound (profiles/r02_ncu_frontend.md): fbank2 moves exactly its "
                         "algorithmic bytes (4.4 % of the DRAM peak) but executes 1.1 G warp instructions per 1,024 x 10 s, which at a "
                         "perfect issue rate is 0.98 ms = 10 % of the HBM peak; it reaches 55 % of that (shared-memory bank conflicts in "
                         "the sparse mel gather are ~30 % of the kernel); see DESIGN.md section 4.4")
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist
    from fangyan_tts_b200 import distributed as D
    from fangyan_tts_b200 import synth_weights
    from fangyan_tts_b200.configs import CamPlusConfig, S3Config
    from fangyan_tts_b200.engine import Engine, token_len
    from fangyan_tts_b200.pipeline import Extractor, Utt, plan_batches, shard_assignment

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    t_start = time.perf_counter()
    watchdog = float(os.environ.get("FY_BENCH_WATCHDOG", "240"))

    def mark(stage, extra_s=0.0):
        """One stderr line per phase and rank, and a watchdog: a phase that takes longer than FY_BENCH_WATCHDOG
        (+ extra_s) seconds dumps every thread's stack and ends the process (a hung rank must not hold an 8-GPU box)."""
        import faulthandler
        sys.stderr.write(f"[bench r{rank}/{world} +{time.perf_counter() - t_start:6.1f}s] {stage}\n")
        sys.stderr.flush()
        faulthandler.cancel_dump_traceback_later()
        if watchdog > 0 and stage != "done":
            faulthandler.dump_traceback_later(watchdog + extra_s, exit=True)

    mark("start")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    mark("process group up")
    eng = Engine(local, args.precision)
    cam_cfg, s3_cfg = CamPlusConfig(), S3Config()
    cam_state, s3_state = synth_weights.campplus_state(cam_cfg), synth_weights.s3tok_state(s3_cfg)
    eng.load_campplus(cam_state, cam_cfg)                                # random-init weights of the named architectures
    eng.load_s3tok(s3_state, s3_cfg)

    mark("weights on the device")
    # ---- the job: one utterance list, identical on every rank, sharded by duration like the CLIs shard wav.scp
    n_total = world * args.utts
    utts = workload(n_total, 1986 + 2)
    dur = [len(u) / 16000.0 for u in utts]
    audio_s = float(sum(dur))
    rank_of = shard_assignment(dur, world)
    mine = np.nonzero(rank_of == rank)[0].tolist()
    rows = [int((rank_of == r).sum()) for r in range(world)]
    tok_rows = [int(sum(token_len(len(utts[i]) // 160) for i in np.nonzero(rank_of == r)[0])) for r in range(world)]
    batches = [[mine[j] for j in b] for b in plan_batches([dur[i] for i in mine], args.batch_seconds, 4096)]
    dev = eng.device
    resident, order, cnts = [], [], []
    for idx in batches:
        pcm = torch.from_numpy(np.concatenate([utts[i] for i in idx])).to(dev)
        spk = torch.tensor([i % N_SPK for i in idx], dtype=torch.int32, device=dev)
        resident.append((pcm, [len(utts[i]) for i in idx], spk))
        order.extend(idx)
        cnts.extend(token_len(len(utts[i]) // 160) for i in idx)
    gidx = torch.tensor(order, dtype=torch.int64, device=dev)
    gcnt = torch.tensor(cnts, dtype=torch.int64, device=dev)
    ssum = torch.zeros(N_SPK, 192, device=dev)
    scnt = torch.zeros(N_SPK, dtype=torch.int32, device=dev)
    engines = [eng]
    last = {}

    def exchange(emb, tok):
        """The job's only communication: results to rank 0 in job order, speaker sums reduced."""
        full = D.gather_stage1(gidx, emb, n_total, rows=rows)
        cnt_full, flat = D.gather_stage2(gidx, gcnt, tok, n_total, rows=rows, tok_rows=tok_rows)
        D.reduce_speakers(ssum, scnt)
        return full, cnt_full, flat, engines[0].speaker_mean(ssum, scnt)

    def step_device():
        e = engines[0]
        ssum.zero_()
        scnt.zero_()
        embs, toks = [], []
        for pcm, lens, spk in resident:
            feat, frames = e.fbank_cmn(pcm, lens)
            emb = e.campplus(feat, frames)
            e.speaker_sum(emb, spk, N_SPK, ssum, scnt)
            mel, mframes = e.logmel(pcm, lens)
            tok, _ = e.s3tok(mel, mframes)
            embs.append(emb)
            toks.append(tok)
        last["emb"], last["tok"] = torch.cat(embs), torch.cat(toks)
        last["out"] = exchange(last["emb"], last["tok"])
        return last["out"]

    ex = Extractor(eng, max_batch_seconds=args.batch_seconds, max_batch_utts=4096)
    host_utts = [Utt(utts[i].reshape(-1, 1), 16000) for i in mine]
    h2d = sum(len(u) for u in utts) * 2                    # whole job, all ranks
    res_e2e = {}

    def step_e2e():
        # public API: host PCM16 in, host embeddings + token lists out (on rank 0 when sharded)
        emb, toks = ex.extract(host_utts, shard=(mine, n_total)) if world > 1 else ex.extract(host_utts)
        res_e2e["emb"], res_e2e["toks"] = emb, toks

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world > 1:
            t = torch.tensor([x], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return x

    def timed(fn, steps):
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(steps):
            fn()
        b.record()
        barrier()
        return max_over_ranks(a.elapsed_time(b))

    warm = max(args.warmup, 3)
    mark("shard resident", extra_s=2.0 * warm)
    for k in range(warm):
        step_device()
        if k == 0:
            torch.cuda.synchronize()
            mark("first step (with the first exchange) done")
    torch.cuda.synchronize()
    mark("warm-up done", extra_s=2.0 * args.steps)           # the timed region grows with --steps
    # ---- the exchange delivered exactly what the ranks computed (fingerprints of index x row bits, all-reduced)
    fp = torch.stack([D.checksum(gidx, last["emb"]), D.checksum_ragged(gidx, gcnt, last["tok"])])
    if world > 1:
        dist.all_reduce(fp)
    gather_ok = None
    if rank == 0:
        full, cnt_full, flat, _ = last["out"]
        everyone = torch.arange(n_total, device=dev)
        gather_ok = bool(int(fp[0]) == int(D.checksum(everyone, full)) and
                         int(fp[1]) == int(D.checksum_ragged(everyone, cnt_full, flat)) and
                         int(cnt_full.sum()) == sum(tok_rows))
    prof = (C.c_double * 10)()
    eng.lib.fy_profile_read(eng.h, prof, 10)          # clear counters
    eng.lib.fy_profile_enable(eng.h, 1)
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ms = timed(step_device, args.steps)
    clk = clocks.stop() if rank == 0 else None
    mark("timed steps done")
    eng.lib.fy_profile_read(eng.h, prof, 10)
    eng.lib.fy_profile_enable(eng.h, 0)
    p = list(prof)

    # ---- the exchange alone (device-timed, max over ranks)
    gather = None
    if world > 1:
        reps = 10
        g_ms = timed(lambda: exchange(last["emb"], last["tok"]), reps) / reps
        peer = [r for r in range(world) if r != 0]
        g_bytes = sum(rows[r] * (8 + 768 + 8 + 8) + tok_rows[r] * 2 for r in peer) + 2 * (N_SPK * 192 * 4 + N_SPK * 4)
        gather = {"ms": g_ms, "bytes": g_bytes, "GBps": g_bytes / (g_ms * 1e-3) / 1e9, "inside_value": True,
                  "what": "gather-v to rank 0 of (index, embedding[192] fp32, token count, int16 token ids) as grouped "
                          "NCCL send/recv over NVLink + all_reduce(SUM) of the speaker sums/counts; bytes = received by "
                          "rank 0 + all_reduce payload", "verified": gather_ok,
                  "share_of_step": g_ms / (ms / args.steps)}

    mark("exchange timed")
    step_e2e()                                            # warm the pinned buffers
    mark("first e2e step done")
    e2e_steps = max(1, min(args.steps, 3))
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_e2e()
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    mark("e2e timed")

    hbm, tf_burst, tf_sust, which = peaks()
    fe = None
    if rank == 0 and world == 1 and not args.no_frontend:
        fe = frontend_figures(eng, hbm)

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

    # Every GPU number is in hand: leave the process group NOW, so that the other ranks exit instead of spinning in a
    # barrier next to rank 0's CPU leg (oracle for `parity`).
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    mark("process group closed", extra_s=600.0)             # rank 0's CPU leg follows (oracle / literal reference stack)
    if rank != 0:
        mark("done")
        return 0
    ms_per_step = ms / args.steps
    value = audio_s / 3600.0 / (ms_per_step / 1e3)
    gemm_ms, gemm_flops, gemm_n = p[0], p[3], p[6]
    achieved = gemm_flops / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else 0.0
    emb_h, toks_h = res_e2e["emb"], res_e2e["toks"]
    d2h = int(emb_h.nbytes + sum(len(t) for t in toks_h) * 4 + 8 * n_total)
    x3 = 3.0 if args.precision == "bf16x3" else 1.0
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warm,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16x3 (split-bf16 tensor-core products, fp32 accumulate)" if args.precision == "bf16x3" else "bf16",
        "data": "synthetic",
        "config": {"workload": f"stage1+2 over one job of {n_total} synthetic 16 kHz utterances of 3-10 s "
                               f"({audio_s / 3600.0:.3f} audio-hours; {args.utts} per GPU = BASELINE configs[1] size, both "
                               f"stages), sharded by duration over {world} rank(s); every step ends with the result "
                               "gather to rank 0 + speaker all_reduce",
                   "utts_per_rank": rows, "batches_per_step": len(batches), "batch_seconds": args.batch_seconds,
                   "weights": "random-init CAMPPlus + S3 tokenizer v3 (12 layers)", "precision": args.precision,
                   "l2": "per-step inputs and workspaces (GBs) exceed the 126 MB L2", "parallelism": f"dp{world}"},
        "e2e": {"value": audio_s / 3600.0 / (e2e_s / e2e_steps), "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "steps": e2e_steps,
                "api": "pipeline.Extractor.extract: host PCM16 -> pinned upload ring -> kernels -> "
                       + ("NCCL gather to rank 0 -> " if world > 1 else "") + "host embeddings + token lists"},
        "gpu_launches": int(p[9]),
        "roofline": {"bound": "tensor", "kernel": "gemm_kernel<BN, OP, NCTA> (tcgen05; all launches: S3 encoder "
                                                  "bf16x3 CTA-pair tiles + CAMPPlus fp16 tiles)",
                     "achieved": achieved, "peak": tf_sust, "unit": "TFLOP/s", "frac": achieved / tf_sust,
                     # dram__bytes_read + write per launch, ncu --set full at THIS batch size (2,400 audio-s per batch):
                     # mean of the four S3 layer GEMMs, > 96 % of the FLOPs counted in `achieved`
                     "traffic": 1.755e9 if args.batch_seconds == 2400.0 and args.precision == "bf16x3" else None,
                     "traffic_source": "profiles/r02_ncu_gemm_pair_bench.md (ncu --set full of QKV / O / fc1 / fc2 at 370 "
                                       "bench utterances = one 2,400 audio-s batch: 1.34 / 0.92 / 1.79 / 2.97 GB per launch)",
                     "peak_source": f"{which} sustained cuBLAS bf16",
                     "executed_tflops": achieved * x3, "executed_frac": achieved * x3 / tf_sust,
                     "note": "achieved counts each MAC once; split-bf16 executes three tensor-core products per MAC "
                             "(ceiling 1/3), executed_frac counts them (CAMPPlus launches, 1 product, are <4% of the FLOPs)",
                     "launches": int(gemm_n), "avg_launch_ms": gemm_ms / max(gemm_n, 1),
                     "gemm_share_of_step": gemm_ms / ms, "attention_ms_per_step": p[1] / args.steps},
        "clocks": clk,
    }
    if gather:
        line["gather"] = gather
    if bf16_mode:
        line["bf16_mode"] = bf16_mode
    if fe:
        line["frontend_batch1024"] = fe
    rc = 0
    if not args.no_cpu_baseline:
        threads = len(os.sched_getaffinity(0))
        n = min(args.cpu_sample if world == 1 else min(args.cpu_sample, 32), n_total)
        sample = utts[:n]
        r = cpu_reference_run(sample, threads, want_outputs=True)
        line["parity"] = parity_report(r["out"], emb_h, toks_h)
        line["parity"]["sample"] = f"utterances 0..{n - 1} of the benchmarked job, GPU results taken from the e2e arm"
        if world == 1:
            note = "oracle = torch-cpu restatement (onnxruntime / whisper not importable)"
            if literal_reference_available():              # time the literal stack instead (outputs not needed here)
                r = cpu_reference_run(sample, threads)
                note = "onnxruntime CPUExecutionProvider on the exported random-init networks + whisper + torchaudio"
            line["cpu_baseline"] = {"value": r["rate"], "unit": UNIT, "cores": threads, "kind": r["kind"],
                                    "sample": f"first {n} utterances of the job ({r['audio_s']:.1f} audio-s, "
                                              f"{r['wall']:.1f} s wall), " + note}
        if not line["parity"]["pass"]:
            rc = 3
    print(json.dumps(line), flush=True)
    mark("done")
    if gather_ok is False:
        sys.stderr.write("bench: the gathered results do not match the per-rank fingerprints\n")
        rc = 4
    if rc == 3:
        sys.stderr.write("bench: a north_star parity gate failed, see the parity object\n")
    return rc


def run_reference(args):
    """Reference arm: the reference's CPU implementation of the path on the host cores (literal onnxruntime stack when
    the box has it, else the oracle port), EXACTLY --steps timed passes after --warmup untimed ones over a bounded
    sample sized from a probe so the whole run stays within a few minutes."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    threads = len(os.sched_getaffinity(0))
    utts = workload(args.utts, 1986 + 2)
    probe = cpu_reference_run(utts[:threads], threads)                    # also loads the models
    per_utt = probe["wall"] / max(len(utts[:threads]), 1)
    budget = 150.0                                                       # seconds for all (warmup + steps) passes
    n = int(budget / max(per_utt * (args.steps + args.warmup), 1e-6))
    n = max(min(n, args.cpu_sample, len(utts)), min(threads, len(utts)))
    sample = utts[:n]
    for _ in range(args.warmup):
        cpu_reference_run(sample, threads)
    rates, wall_tot, kind = [], 0.0, "port"
    for _ in range(args.steps):
        r = cpu_reference_run(sample, threads)
        rates.append(r["rate"])
        wall_tot += r["wall"]
        kind = r["kind"]
    v = float(np.median(rates))
    what = ("onnxruntime CPUExecutionProvider + whisper + torchaudio, as single_job runs them" if kind == "reference" else
            "torch-cpu restatement of the reference path (onnxruntime / whisper / the .onnx files are not installable offline)")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * wall_tot / max(len(rates), 1), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"stage1+2, bounded sample: first {n} of {args.utts} synthetic utterances of 3-10 s per step, "
                               "batch 1 at true length as the reference does", "threads": threads},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": kind,
                         "sample": f"{n} utterances per step, {args.steps} timed steps; " + what},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)
    return 0


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
    return 0


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
    return 0


def run_serialise(args):
    """SURVEY.md 8(d) region (iii), serialisation only (row A11): from the gathered result arrays on rank 0's host to
    utt2embedding.pt / utt2speech_token.pt.  `value` = the array-direct writers the CLIs use (kaldi_io.save_*_atomic);
    `cpu_baseline` = what the reference does at extract_embedding.py:41,57-58 / extract_speech_token.py:48,58: a
    Python list per key (tolist) and torch.save of the dict.  Both outputs are loaded back and compared.  No GPU."""
    import tempfile
    import torch
    from fangyan_tts_b200 import kaldi_io
    from fangyan_tts_b200.pipeline import split_tokens
    n = args.serialise
    rng = np.random.default_rng(1986 + 6)
    keys = [f"utt_{i:07d}" for i in range(n)]
    emb = rng.standard_normal((n, 192)).astype(np.float32)
    counts = rng.integers(75, 251, n)
    flat = rng.integers(0, 6561, int(counts.sum())).astype(np.int32)
    with tempfile.TemporaryDirectory() as tmp:
        t0 = time.perf_counter()
        kaldi_io.save_embeddings_atomic(keys, emb, os.path.join(tmp, "utt2embedding.pt"))
        t_e = time.perf_counter() - t0
        t0 = time.perf_counter()
        kaldi_io.save_tokens_atomic(keys, flat, counts, os.path.join(tmp, "utt2speech_token.pt"))
        t_t = time.perf_counter() - t0
        t0 = time.perf_counter()
        kaldi_io.save_atomic(kaldi_io.embeddings_to_dict(keys, emb), os.path.join(tmp, "ref_e.pt"))
        r_e = time.perf_counter() - t0
        t0 = time.perf_counter()
        kaldi_io.save_atomic(dict(zip(keys, split_tokens(flat, counts))), os.path.join(tmp, "ref_t.pt"))
        r_t = time.perf_counter() - t0
        sizes = {f: os.path.getsize(os.path.join(tmp, f)) for f in ("utt2embedding.pt", "utt2speech_token.pt", "ref_e.pt", "ref_t.pt")}
        t0 = time.perf_counter()
        same = (torch.load(os.path.join(tmp, "utt2embedding.pt")) == torch.load(os.path.join(tmp, "ref_e.pt")) and
                torch.load(os.path.join(tmp, "utt2speech_token.pt")) == torch.load(os.path.join(tmp, "ref_t.pt")))
        t_load = time.perf_counter() - t0
    hours = n * 6.5 / 3600.0
    print(json.dumps({
        "metric": "serialise_rows_per_sec", "unit": "utterances/s", "utts": n, "approx_audio_hours": hours,
        "value": n / (t_e + t_t), "seconds": {"utt2embedding.pt": t_e, "utt2speech_token.pt": t_t},
        "cpu_baseline": {"kind": "port", "value": n / (r_e + r_t), "cores": 1,
                         "seconds": {"utt2embedding.pt": r_e, "utt2speech_token.pt": r_t},
                         "sample": "tolist() per key + torch.save(dict of lists), as the reference's loops do"},
        "identical_after_torch_load": bool(same), "load_back_seconds": t_load, "bytes": sizes,
        "host": {"cores": len(os.sched_getaffinity(0))}}), flush=True)
    return 0 if same else 5


def run_cli_e2e(args):
    """SURVEY.md 8(d) regions (ii) and (iii) through the shipped command lines: N keys over `--cli-files` distinct
    16 kHz PCM16 WAVs on tmpfs -> tools/extract_embedding.py, then tools/extract_speech_token.py (separate processes,
    as run.sh:23-41 runs them) -> the three .pt files.  Each CLI prints its own time split; this adds the process wall
    (interpreter start, imports, weight upload included) and the device-resident rate of the same job for scale."""
    import shutil
    import tempfile
    from fangyan_tts_b200 import synth, synth_weights
    from fangyan_tts_b200.configs import S3Config
    n, n_files = args.cli_e2e, max(1, min(args.cli_files, args.cli_e2e))
    base = "/dev/shm" if os.path.isdir("/dev/shm") and shutil.disk_usage("/dev/shm").free > 6e9 else None
    tmp = Path(tempfile.mkdtemp(prefix="fy_cli_", dir=base))
    try:
        clips = workload(n_files, 1986 + 3)
        (tmp / "wav").mkdir()
        for i, u in enumerate(clips):
            synth.write_wav(tmp / "wav" / f"c{i:05d}.wav", u)
        audio_s = 0.0
        with open(tmp / "wav.scp", "w") as scp, open(tmp / "utt2spk", "w") as u2s:
            for i in range(n):
                j = i % n_files
                scp.write(f"utt_{i:07d} {tmp / 'wav' / f'c{j:05d}.wav'}\n")
                u2s.write(f"utt_{i:07d} spk{i % N_SPK:02d}\n")
                audio_s += len(clips[j]) / 16000.0
        np.savez(tmp / "campplus.npz", **synth_weights.campplus_state())
        np.savez(tmp / "s3tok.npz", **synth_weights.s3tok_state(S3Config()))
        env = dict(os.environ)
        if args.gpus:
            env["FY_GPUS"] = str(args.gpus)
        out = {"metric": "cli_audio_hours_per_sec_files_to_pt", "unit": UNIT, "n_gpus": args.gpus, "utts": n,
               "distinct_wav_files": n_files, "audio_hours": audio_s / 3600.0, "tmp": str(tmp),
               "config": {"workload": f"{n} wav.scp keys over {n_files} distinct 3-10 s PCM16 WAVs on "
                                      f"{'tmpfs' if base else 'the scratch disk'}; stage 1 then stage 2 as separate processes"}}
        total = 0.0
        for tool, weights, key in (("extract_embedding.py", "campplus.npz", "stage1"),
                                   ("extract_speech_token.py", "s3tok.npz", "stage2")):
            t0 = time.perf_counter()
            r = subprocess.run([sys.executable, str(ROOT / "tools" / tool), "--dir", str(tmp), "--onnx_path",
                                str(tmp / weights), "--num_thread", str(min(32, len(os.sched_getaffinity(0))))],
                               capture_output=True, text=True, env=env)
            wall = time.perf_counter() - t0
            if r.returncode != 0:
                sys.stderr.write(r.stderr[-4000:])
                return 1
            split = [json.loads(l) for l in r.stderr.splitlines() if l.startswith('{"stage"')]
            out[key] = {"process_wall_s": wall, **(split[-1] if split else {})}
            total += wall
        sizes = {f: os.path.getsize(tmp / f) for f in ("utt2embedding.pt", "spk2embedding.pt", "utt2speech_token.pt")}
        out["value"] = audio_s / 3600.0 / total
        out["total_wall_s"] = total
        out["in_process_value"] = audio_s / 3600.0 / (out["stage1"].get("wall_s", 0) + out["stage2"].get("wall_s", 0) or 1)
        out["output_bytes"] = sizes
        print(json.dumps(out), flush=True)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return 0


def run_long_stress(args):
    """BASELINE configs[3]: 256 clips of U(1, 30) s plus two at exactly 30.0 s, length-sorted packed batches, all 12
    encoder layers, in both encoder precisions; token parity of each mode against the oracle on a sample."""
    import torch
    from fangyan_tts_b200 import synth, synth_weights
    from fangyan_tts_b200.configs import CamPlusConfig, S3Config
    from fangyan_tts_b200.engine import Engine
    from fangyan_tts_b200.pipeline import plan_batches
    rng = np.random.default_rng(1986 + 3)
    pool = [synth.utterance(rng, 30.0) for _ in range(6)]
    utts = [np.ascontiguousarray(pool[i % 6][: (480000 if i in (0, 255) else int(rng.uniform(1.0, 30.0) * 16000))])
            for i in range(256)]
    dur = [len(u) / 16000.0 for u in utts]
    audio_s = float(sum(dur))
    sample = [0, 17, 64, 128, 200, 255]                                  # includes both 30.0 s clips
    threads = len(os.sched_getaffinity(0))
    ref = cpu_reference_run([utts[i] for i in sample], threads, want_outputs=True)
    cam_state, s3_state = synth_weights.campplus_state(), synth_weights.s3tok_state(S3Config())
    hbm, tf_burst, tf_sust, which = peaks()
    res = {"metric": METRIC, "unit": UNIT, "config": {"workload": "BASELINE configs[3]: 256 clips of U(1,30) s incl. two at "
           "exactly 30.0 s, one length-sorted packed batch per <= 2,400 audio-s, 12 encoder layers",
           "audio_hours": audio_s / 3600.0}, "n_gpus": 1, "data": "synthetic",
           "cpu_oracle_sample": {"utts": len(sample), "wall_s": ref["wall"], "cores": threads}}
    for prec in ("bf16x3", "bf16"):
        eng = Engine(0, prec)
        eng.load_campplus(cam_state, CamPlusConfig())
        eng.load_s3tok(s3_state, S3Config())
        dev = eng.device
        batches = plan_batches(dur, args.batch_seconds, 256)
        resident = [(torch.from_numpy(np.concatenate([utts[i] for i in b])).to(dev), [len(utts[i]) for i in b]) for b in batches]

        def step():
            outs = []
            for pcm, lens in resident:
                feat, frames = eng.fbank_cmn(pcm, lens)
                emb = eng.campplus(feat, frames)
                mel, mframes = eng.logmel(pcm, lens)
                tok, counts = eng.s3tok(mel, mframes)
                outs.append((emb, tok, counts))
            return outs

        for _ in range(3):
            outs = step()
        prof = (C.c_double * 10)()
        eng.lib.fy_profile_read(eng.h, prof, 10)
        eng.lib.fy_profile_enable(eng.h, 1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        for _ in range(args.steps):
            outs = step()
        b.record()
        torch.cuda.synchronize()
        eng.lib.fy_profile_read(eng.h, prof, 10)
        eng.lib.fy_profile_enable(eng.h, 0)
        ms = a.elapsed_time(b) / args.steps
        emb = np.zeros((256, 192), np.float32)
        toks = [None] * 256
        for bi, (e, t, counts) in zip(batches, outs):
            emb[bi] = e.cpu().numpy()
            th, o = t.cpu().numpy(), 0
            for i, c in zip(bi, counts):
                toks[i] = th[o:o + c].tolist()
                o += c
        par = parity_report(ref["out"], emb[sample], [toks[i] for i in sample])
        p = list(prof)
        res[prec] = {"value": audio_s / 3600.0 / (ms * 1e-3), "ms_per_step": ms, "batches": len(batches),
                     "attention_ms_per_step": p[1] / args.steps, "attention_share": p[1] / args.steps / ms,
                     "gemm_ms_per_step": p[0] / args.steps, "gemm_tflops_algorithmic": p[3] / (p[0] * 1e-3) / 1e12 if p[0] else 0,
                     "len_750_at_30s": len(toks[0]) == 750 and len(toks[255]) == 750, "parity": par}
        eng.close()
    print(json.dumps(res), flush=True)
    return 0


if __name__ == "__main__":
    a = parse_args()
    if a.stage3_pack:
        rc = run_stage3_pack(a)
    elif a.serialise:
        rc = run_serialise(a)
    elif a.frontend_sweep:
        rc = run_frontend_sweep(a)
    elif a.cli_e2e:
        rc = run_cli_e2e(a)
    elif a.long_stress:
        rc = run_long_stress(a)
    elif a.impl == "reference":
        rc = run_reference(a)
    else:
        rc = run_ours(a)
    sys.exit(rc or 0)
