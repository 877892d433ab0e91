"""GPU end-to-end test of the drop-in boundary B1/B2: the two CLIs over a Kaldi data dir, and the front-end twin.
Checks the file contract the consumer relies on (make_parquet_list.py:35-37,107-109) and parity with the oracle."""
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

from fangyan_tts_b200 import synth, synth_weights
from fangyan_tts_b200.configs import CamPlusConfig, S3Config
from oracle import campplus as CP
from oracle import frontend as F
from oracle import s3tokenizer as S3

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent
N_LAYER = 2


@pytest.fixture(scope="module")
def data_dir(tmp_path_factory):
    d = tmp_path_factory.mktemp("kaldi")
    utts = synth.make_set(9, 1986, 0.5, 4.0)
    utts.append(synth.make_set(1, 7, 30.2, 30.3)[0])              # > 30 s: tokens must be []
    utts.append(synth.utterance(np.random.default_rng(3), 30.0))   # exactly 30.0 s: tokenised
    keys = synth.write_kaldi_dir(d, utts, n_spk=3)
    # one 24 kHz stereo file: exercises resampling and the channel rules of the two stages
    rng = np.random.default_rng(5)
    st = np.stack([synth.utterance(rng, 1.5, 24000), synth.utterance(rng, 1.5, 24000)], axis=1)
    synth.write_wav(d / "wav" / "stereo24.wav", st, 24000)
    with open(d / "wav.scp", "a") as f:
        f.write(f"stereo24 {d / 'wav' / 'stereo24.wav'}\n")
    with open(d / "utt2spk", "a") as f:
        f.write("stereo24 spk00\n")
    np.savez(d / "campplus.npz", **synth_weights.campplus_state(seed=11))
    cfg = d / "s3.json"
    cfg.write_text('{"n_layer": %d}' % N_LAYER)
    np.savez(d / "s3tok.npz", **synth_weights.s3tok_state(S3Config(n_layer=N_LAYER), seed=12))
    return d, keys + ["stereo24"], utts, st


def _run(tool, d, weights, env_extra=None):
    import os
    env = dict(os.environ, **{"FY_GPUS": "1", **(env_extra or {})})
    r = subprocess.run([sys.executable, str(ROOT / "tools" / tool), "--dir", str(d), "--onnx_path", str(weights),
                        "--num_thread", "4"], capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stderr[-2000:]
    return r


def test_extract_embedding_cli(data_dir):
    d, keys, utts, st = data_dir
    _run("extract_embedding.py", d, d / "campplus.npz")
    u2e, s2e = torch.load(d / "utt2embedding.pt"), torch.load(d / "spk2embedding.pt")
    assert type(u2e) is dict and sorted(u2e) == sorted(keys)       # every wav.scp key exactly once
    assert all(type(v) is list and len(v) == 192 and type(v[0]) is float for v in u2e.values())
    assert sorted(s2e) == ["spk00", "spk01", "spk02"]
    cam = CP.CamPlus()
    cam.load_state_dict({k: torch.from_numpy(v) for k, v in synth_weights.campplus_state(seed=11).items()}, strict=False)
    cam.eval()
    refs = {}
    for k, u in zip(keys, utts):
        refs[k] = CP.embed_one(cam, F.cmn(F.kaldi_fbank(torch.from_numpy(u.astype(np.float32) / 32768.0)[None])))
    w = torch.from_numpy(st.astype(np.float32).T / 32768.0)
    refs["stereo24"] = CP.embed_one(cam, F.cmn(F.kaldi_fbank(F.resample(w, 24000))))    # kaldi.fbank takes channel 0
    for k in keys:
        cos = float(torch.nn.functional.cosine_similarity(torch.tensor(u2e[k]), refs[k], dim=0))
        assert cos >= 0.9999, (k, cos)
    spk = dict(l.split() for l in open(d / "utt2spk"))
    for s, v in s2e.items():
        ref = torch.stack([torch.tensor(u2e[k]) for k in keys if spk[k] == s]).mean(0)   # extract_embedding.py:55-56
        assert torch.allclose(torch.tensor(v), ref, atol=1e-5)


def test_extract_speech_token_cli(data_dir):
    d, keys, utts, st = data_dir
    _run("extract_speech_token.py", d, d / "s3tok.npz", {"FY_S3TOK_CONFIG": str(d / "s3.json")})
    u2t = torch.load(d / "utt2speech_token.pt")
    assert type(u2t) is dict and sorted(u2t) == sorted(keys)
    cfg = S3.S3Config(n_layer=N_LAYER)
    m = S3.S3Tokenizer(cfg)
    m.load_state_dict({k: torch.from_numpy(v) for k, v in synth_weights.s3tok_state(S3Config(n_layer=N_LAYER), seed=12).items()})
    m.eval()
    waves = {k: torch.from_numpy(u.astype(np.float32) / 32768.0)[None] for k, u in zip(keys, utts)}
    w = torch.from_numpy(st.astype(np.float32).T / 32768.0)
    waves["stereo24"] = F.resample(w, 24000).mean(dim=0, keepdim=True)           # resample then mono (:29-33)
    total = bad = 0
    for k in keys:
        n = waves[k].shape[1]
        if not F.stage2_gate(n):
            assert u2t[k] == []                                                  # key kept, empty list (:34-36)
            continue
        mel = F.whisper_log_mel(waves[k])
        ref = S3.tokenize_one(m, mel)
        assert type(u2t[k]) is list and len(u2t[k]) == len(ref) == S3.token_len(n // 160)
        assert all(type(t) is int and 0 <= t < 6561 for t in u2t[k])
        total += len(ref)
        bad += sum(a != b for a, b in zip(u2t[k], ref))
    assert len(u2t[keys[-2]]) == 750                                             # exactly 30.0 s is tokenised
    assert bad / total <= 1e-3, (bad, total)


def test_cli_outputs_feed_stage3(data_dir, tmp_path):
    """Row N1: the sidecars the CLIs leave hold exactly the .pt values, and the Arrow stage-3 writer fed by them
    produces the shard the reference's pandas job produces from the .pt files."""
    import pyarrow.parquet as pq
    from fangyan_tts_b200 import parquet_pack
    from oracle import stage3
    d, keys, utts, st = data_dir
    if not (d / "utt2embedding.pt").exists():
        _run("extract_embedding.py", d, d / "campplus.npz")
    if not (d / "utt2speech_token.pt").exists():
        _run("extract_speech_token.py", d, d / "s3tok.npz", {"FY_S3TOK_CONFIG": str(d / "s3.json")})
    e, sp, tk = stage3.load_pt(str(d))
    z = np.load(d / parquet_pack.EMB_SIDECAR)
    assert [str(k) for k in z["keys"]] == list(e) and np.array_equal(z["emb"], np.array([e[k] for k in e], dtype=np.float32))
    z = np.load(d / parquet_pack.TOK_SIDECAR)
    assert [str(k) for k in z["keys"]] == list(tk)
    assert z["flat"].tolist() == [t for k in tk for t in tk[k]]
    with open(d / "text", "w") as f:
        for k in keys:
            f.write(f"{k} text of {k}\n")
    parquet_pack.make_parquet_list(str(d), str(tmp_path / "pq"), 1000, 2)
    rd = lambda name: {l.split()[0]: " ".join(l.split()[1:]) for l in open(d / name)}
    stage3.job(list(rd("wav.scp")), tmp_path / "ref.parquet", rd("wav.scp"), rd("text"), rd("utt2spk"), e, sp, tk)
    a = pq.read_table(tmp_path / "pq" / "parquet_000000000.tar")
    b = pq.read_table(tmp_path / "ref.parquet").replace_schema_metadata(None)
    assert a.equals(b.cast(a.schema))


def test_cli_two_gpus_matches_one(data_dir, tmp_path):
    """Self-spawned multi-GPU run (one worker per GPU, NCCL gather) writes the same files as the single-GPU run."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import shutil
    d, keys, utts, st = data_dir
    for name in ("wav.scp", "utt2spk"):
        shutil.copy(d / name, tmp_path / name)
    _run("extract_embedding.py", tmp_path, d / "campplus.npz", {"FY_GPUS": "2"})
    _run("extract_speech_token.py", tmp_path, d / "s3tok.npz", {"FY_GPUS": "2", "FY_S3TOK_CONFIG": str(d / "s3.json")})
    if not (d / "utt2embedding.pt").exists():
        _run("extract_embedding.py", d, d / "campplus.npz")
    if not (d / "utt2speech_token.pt").exists():
        _run("extract_speech_token.py", d, d / "s3tok.npz", {"FY_S3TOK_CONFIG": str(d / "s3.json")})
    a, b = torch.load(d / "utt2embedding.pt"), torch.load(tmp_path / "utt2embedding.pt")
    assert sorted(a) == sorted(b)
    for k in a:     # batches differ between 1 and 2 ranks, results must not (each utterance is computed alone)
        assert a[k] == b[k], k
    assert torch.load(d / "utt2speech_token.pt") == torch.load(tmp_path / "utt2speech_token.pt")
    sa, sb = torch.load(d / "spk2embedding.pt"), torch.load(tmp_path / "spk2embedding.pt")
    for k in sa:
        assert torch.allclose(torch.tensor(sa[k]), torch.tensor(sb[k]), atol=1e-5)


def test_cli_is_all_or_nothing(data_dir, tmp_path):
    d, keys, utts, st = data_dir
    synth.write_kaldi_dir(tmp_path, utts[:2])
    with open(tmp_path / "wav.scp", "a") as f:
        f.write(f"broken {tmp_path / 'wav' / 'nope.wav'}\n")
    with open(tmp_path / "utt2spk", "a") as f:
        f.write("broken spk00\n")
    import os
    r = subprocess.run([sys.executable, str(ROOT / "tools" / "extract_embedding.py"), "--dir", str(tmp_path), "--onnx_path",
                        str(d / "campplus.npz")], capture_output=True, text=True, env=dict(os.environ, FY_GPUS="1"))
    assert r.returncode != 0 and "broken" in r.stderr
    assert not (tmp_path / "utt2embedding.pt").exists() and not (tmp_path / "spk2embedding.pt").exists()


def test_frontend_twin(data_dir):
    """B2: CosyVoiceFrontEnd._extract_spk_embedding / _extract_speech_token semantics (frontend.py:95-118)."""
    import os
    d, keys, utts, st = data_dir
    os.environ["FY_S3TOK_CONFIG"] = str(d / "s3.json")
    try:
        from fangyan_tts_b200.frontend import CosyVoiceFrontEndB200
        fe = CosyVoiceFrontEndB200(str(d / "campplus.npz"), str(d / "s3tok.npz"))
    finally:
        del os.environ["FY_S3TOK_CONFIG"]
    wav = str(d / "wav" / f"{keys[0]}.wav")
    emb = fe._extract_spk_embedding(wav)
    tok, tok_len = fe._extract_speech_token(wav)
    assert emb.shape == (1, 192) and emb.dtype == torch.float32 and emb.is_cuda
    assert tok.dtype == torch.int32 and tok.dim() == 2 and tok_len.tolist() == [tok.shape[1]]
    assert tok.shape[1] == S3.token_len(len(utts[0]) // 160)
    with pytest.raises(AssertionError):
        fe._extract_speech_token(str(d / "wav" / f"{keys[9]}.wav"))               # > 30 s asserts in the twin
    # stereo 24 kHz: the twin takes the channel MEAN, then resamples (file_utils.py:46-49)
    cam = CP.CamPlus()
    cam.load_state_dict({k: torch.from_numpy(v) for k, v in synth_weights.campplus_state(seed=11).items()}, strict=False)
    cam.eval()
    w = torch.from_numpy(st.astype(np.float32).T / 32768.0).mean(dim=0, keepdim=True)
    ref = CP.embed_one(cam, F.cmn(F.kaldi_fbank(F.resample(w, 24000))))
    got = fe._extract_spk_embedding(str(d / "wav" / "stereo24.wav"))[0].cpu()
    assert float(torch.nn.functional.cosine_similarity(got, ref, dim=0)) >= 0.9999


def test_twin_loads_exporter_style_onnx(data_dir, tmp_path):
    """the real --onnx_path files carry neither our names nor our layouts.  Graphs written with the torch
    exporter's conventions (tests/onnx_writer.py: BatchNorm folded into Conv, MatMul [in, out], foreign names) load
    through the topology mapping and give the same embedding / tokens as the .npz state they were written from."""
    import os
    import onnx_writer as OW
    from fangyan_tts_b200.frontend import CosyVoiceFrontEndB200
    d, keys, utts, st = data_dir
    s3_cfg = S3Config(n_layer=N_LAYER)
    OW.write_campplus(tmp_path / "campplus.onnx", synth_weights.campplus_state(seed=11), CamPlusConfig())
    OW.write_s3tok(tmp_path / "speech_tokenizer_v3.onnx", synth_weights.s3tok_state(s3_cfg, seed=12), s3_cfg)
    os.environ["FY_S3TOK_CONFIG"] = str(d / "s3.json")
    try:
        a = CosyVoiceFrontEndB200(str(d / "campplus.npz"), str(d / "s3tok.npz"))
        b = CosyVoiceFrontEndB200(str(tmp_path / "campplus.onnx"), str(tmp_path / "speech_tokenizer_v3.onnx"))
    finally:
        del os.environ["FY_S3TOK_CONFIG"]
    wav = str(d / "wav" / f"{keys[1]}.wav")
    ea, eb = a._extract_spk_embedding(wav)[0].cpu(), b._extract_spk_embedding(wav)[0].cpu()
    assert float(torch.nn.functional.cosine_similarity(ea, eb, dim=0)) >= 0.99999     # folding moves rounding only
    ta, tb = a._extract_speech_token(wav)[0].cpu(), b._extract_speech_token(wav)[0].cpu()
    assert torch.equal(ta, tb)                                                         # same weights bit for bit


def test_twin_cuda_graph_replay_is_bit_identical(data_dir, tmp_path):
    """B2 latency path: the captured launch sequence of a prompt length gives the same bits as the eager calls, for the
    prompt it was captured on and for a different prompt of the same length (replay with new input)."""
    import os
    from fangyan_tts_b200.frontend import CosyVoiceFrontEndB200
    d, keys, utts, st = data_dir
    rng = np.random.default_rng(8)
    a, b = synth.utterance(rng, 2.5), synth.utterance(rng, 2.5)
    synth.write_wav(tmp_path / "a.wav", a)
    synth.write_wav(tmp_path / "b.wav", b)
    os.environ["FY_S3TOK_CONFIG"] = str(d / "s3.json")
    try:
        eager = CosyVoiceFrontEndB200(str(d / "campplus.npz"), str(d / "s3tok.npz"), use_graphs=False)
        graph = CosyVoiceFrontEndB200(str(d / "campplus.npz"), str(d / "s3tok.npz"), use_graphs=True)
    finally:
        del os.environ["FY_S3TOK_CONFIG"]
    for name in ("a", "b", "a", "b"):                 # first a: eager warm-up + capture; then three replays
        wav = str(tmp_path / f"{name}.wav")
        assert torch.equal(graph._extract_spk_embedding(wav), eager._extract_spk_embedding(wav)), name
        tg, lg = graph._extract_speech_token(wav)
        te, le = eager._extract_speech_token(wav)
        assert torch.equal(tg, te) and torch.equal(lg, le), name
    assert len(graph._graphs) == 2                    # one graph per (method, length)
