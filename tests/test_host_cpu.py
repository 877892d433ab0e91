"""CPU tests of the host side: batching / sharding policy, WAV decode (row A1), Kaldi tables, the .pt output
contract (B1), the ONNX initialiser reader and the synthetic-weight tables."""
import os
import struct
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

from fangyan_tts_b200 import audio_io, kaldi_io, synth, synth_weights, weights
from fangyan_tts_b200.cli import len16k, prefetch
from fangyan_tts_b200.configs import CamPlusConfig, S3Config
from fangyan_tts_b200.engine import kaldi_num_frames, token_len, whisper_num_frames
from fangyan_tts_b200.pipeline import plan_batches, shard_indices

ROOT = Path(__file__).resolve().parent.parent


# ---------------------------------------------------------------------------------------- batching / sharding
def test_plan_batches_covers_everything_once_and_respects_limits():
    rng = np.random.default_rng(0)
    dur = rng.uniform(0.5, 30.0, 500).tolist()
    batches = plan_batches(dur, 600.0, 64)
    flat = [i for b in batches for i in b]
    assert sorted(flat) == list(range(500))
    for b in batches:
        assert len(b) <= 64
        assert sum(dur[i] for i in b) <= 600.0 or len(b) == 1
    # sorted by length: padding-free packing still wants similar lengths per batch (attention tiles, CMN blocks)
    firsts = [dur[b[0]] for b in batches]
    assert firsts == sorted(firsts)


def test_plan_batches_empty_and_single_long():
    assert plan_batches([], 10.0, 4) == []
    assert plan_batches([50.0], 10.0, 4) == [[0]]


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_shard_indices_partition_and_balance(world):
    rng = np.random.default_rng(world)
    dur = rng.uniform(3.0, 10.0, 1000).tolist()
    shards = [shard_indices(dur, r, world) for r in range(world)]
    assert sorted(i for s in shards for i in s) == list(range(1000))       # no duplication, nothing dropped
    loads = [sum(dur[i] for i in s) for s in shards]
    assert max(loads) - min(loads) <= 10.0                                  # within one utterance of each other


def test_length_formulas():
    assert [kaldi_num_frames(n) for n in (0, 399, 400, 559, 560, 48000)] == [0, 0, 1, 1, 2, 298]
    assert [whisper_num_frames(n) for n in (0, 159, 160, 48000)] == [0, 0, 1, 300]
    assert [token_len(t) for t in (1, 2, 3, 4, 5, 300, 1000, 3000)] == [1, 1, 1, 1, 2, 75, 250, 750]
    assert len16k(24000, 24000) == 16000 and len16k(44100, 44100) == 16000 and len16k(7, 44100) == 3
    assert len16k(12345, 16000) == 12345


def test_prefetch_preserves_order_and_propagates_errors():
    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(4) as pool:
        got = list(prefetch([[0, 1], [2], [3, 4, 5]], lambda i: i * i, pool))
        assert got == [([0, 1], [0, 1]), ([2], [4]), ([3, 4, 5], [9, 16, 25])]

        def boom(i):
            if i == 2:
                raise audio_io.AudioDecodeError("utterance k2: bad")
            return i
        with pytest.raises(audio_io.AudioDecodeError, match="k2"):
            list(prefetch([[0, 1], [2]], boom, pool))


# ---------------------------------------------------------------------------------------- A1 decode
def _riff(fmt_tag, ch, sr, bits, payload, extensible=False):
    if extensible:
        fmt = struct.pack("<HHIIHHHHIH14s", 0xFFFE, ch, sr, sr * ch * bits // 8, ch * bits // 8, bits, 22, bits, 0, fmt_tag,
                          b"\x00\x00\x00\x00\x10\x00\x80\x00\x00\xaa\x00\x38\x9b\x71")
    else:
        fmt = struct.pack("<HHIIHH", fmt_tag, ch, sr, sr * ch * bits // 8, ch * bits // 8, bits)
    junk = b"LIST" + struct.pack("<I", 5) + b"abcde\x00"           # odd-sized chunk + pad byte before fmt
    body = b"WAVE" + junk + b"fmt " + struct.pack("<I", len(fmt)) + fmt + b"data" + struct.pack("<I", len(payload)) + payload
    return b"RIFF" + struct.pack("<I", len(body)) + body


def test_read_wav_pcm16_mono_and_stereo(tmp_path):
    pcm = synth.make_set(1, 3, 0.2, 0.3)[0]
    synth.write_wav(tmp_path / "m.wav", pcm)
    x, sr = audio_io.read_wav(str(tmp_path / "m.wav"))
    assert sr == 16000 and x.dtype == np.int16 and x.shape == (len(pcm), 1) and np.array_equal(x[:, 0], pcm)
    st = np.stack([pcm, -pcm], axis=1)
    synth.write_wav(tmp_path / "s.wav", st, 24000)
    x, sr = audio_io.read_wav(str(tmp_path / "s.wav"))
    assert sr == 24000 and x.shape == st.shape and np.array_equal(x, st)
    assert kaldi_io.wav_header(str(tmp_path / "s.wav")) == (len(pcm), 2, 24000)


def test_read_wav_other_encodings(tmp_path):
    v = np.array([0, 1000, -1000, 32767, -32768], dtype=np.int16)
    cases = {
        "u8": (1, 8, ((v.astype(np.int32) >> 8) + 128).astype(np.uint8).tobytes(), (v >> 8).astype(np.float32) / 128.0),
        "s24": (1, 24, b"".join(struct.pack("<i", int(s) << 8)[:3] for s in v), v.astype(np.float32) / 32768.0),
        "s32": (1, 32, (v.astype(np.int32) << 16).astype("<i4").tobytes(), v.astype(np.float32) / 32768.0),
        "f32": (3, 32, (v.astype(np.float32) / 32768.0).astype("<f4").tobytes(), v.astype(np.float32) / 32768.0),
    }
    for name, (tag, bits, payload, want) in cases.items():
        p = tmp_path / f"{name}.wav"
        p.write_bytes(_riff(tag, 1, 16000, bits, payload))
        x, sr = audio_io.read_wav(str(p), name)
        assert sr == 16000 and x.dtype == np.float32 and np.allclose(x[:, 0], want, atol=1e-7), name
    p = tmp_path / "ext.wav"
    p.write_bytes(_riff(1, 1, 16000, 16, v.tobytes(), extensible=True))
    x, _ = audio_io.read_wav(str(p))
    assert x.dtype == np.int16 and np.array_equal(x[:, 0], v)


def test_read_wav_fails_loudly_with_key(tmp_path):
    p = tmp_path / "x.mp3"
    p.write_bytes(b"ID3\x03" + b"\x00" * 64)
    with pytest.raises(audio_io.AudioDecodeError, match="utt_42"):
        audio_io.read_wav(str(p), "utt_42")
    with pytest.raises(audio_io.AudioDecodeError, match="utt_43"):
        audio_io.read_wav(str(tmp_path / "missing.wav"), "utt_43")
    with pytest.raises(audio_io.AudioDecodeError):
        kaldi_io.wav_header(str(p), "utt_42")


def test_native_scan_and_direct_read_agree_with_the_python_reader(tmp_path):
    """Rows A1 / N2: fy_wav_scan + fy_wav_read_s16 (host threads in the library) against audio_io.read_wav, over the
    header variants of _riff (junk chunk before fmt, extensible), other encodings (flagged not-simple) and errors."""
    pcm = synth.make_set(6, 4, 0.05, 0.4)
    keys = synth.write_kaldi_dir(tmp_path, pcm)
    paths = [str(tmp_path / "wav" / f"{k}.wav") for k in keys]
    v = np.array([0, 1000, -1000, 32767, -32768], dtype=np.int16)
    extra = {"ext": _riff(1, 1, 16000, 16, v.tobytes(), extensible=True),
             "f32": _riff(3, 1, 16000, 32, (v.astype(np.float32) / 32768.0).astype("<f4").tobytes()),
             "st24k": _riff(1, 2, 24000, 16, np.stack([v, -v], 1).tobytes()),
             "cut": _riff(1, 1, 16000, 16, v.tobytes())[:-4]}                 # data chunk longer than the file
    for name, blob in extra.items():
        (tmp_path / f"{name}.wav").write_bytes(blob)
        keys.append(name)
        paths.append(str(tmp_path / f"{name}.wav"))
    sc = audio_io.scan_wavs(keys, paths, 3)
    for i, (k, p) in enumerate(zip(keys, paths)):
        x, sr = audio_io.read_wav(p, k)
        assert (int(sc["frames"][i]), int(sc["channels"][i]), int(sc["sample_rate"][i])) == (x.shape[0], x.shape[1], sr), k
        assert kaldi_io.wav_header(p, k) == (x.shape[0], x.shape[1], sr)
    assert sc["simple"].tolist() == [True] * 6 + [True, False, True, True]
    idx = [8, 3, 6, 0, 9]                                                    # stereo file: interleaved as stored
    dst = np.zeros(int((sc["frames"][idx] * sc["channels"][idx]).sum()), np.int16)
    audio_io.read_s16_into(sc, idx, dst, 2, keys)
    want = np.concatenate([audio_io.read_wav(paths[i])[0].reshape(-1) for i in idx])
    assert np.array_equal(dst, want)
    (tmp_path / "bad.wav").write_bytes(b"ID3" + b"\x00" * 80)
    with pytest.raises(audio_io.AudioDecodeError, match="k_bad"):
        audio_io.scan_wavs(["k_bad"], [str(tmp_path / "bad.wav")], 1)
    with pytest.raises(audio_io.AudioDecodeError, match="k_gone"):
        audio_io.scan_wavs(["k_gone"], [str(tmp_path / "gone.wav")], 1)
    big = _riff(1, 1, 16000, 16, v.tobytes())                                # data chunk beyond the first 4 KiB
    big = big[:12] + b"LIST" + struct.pack("<I", 5000) + b"\x00" * 5000 + big[12:]
    (tmp_path / "long.wav").write_bytes(big)
    sl = audio_io.scan_wavs(["long"], [str(tmp_path / "long.wav")], 1)
    assert int(sl["frames"][0]) == 5 and not sl["simple"][0]                # parsed by the Python reader, not direct


# ---------------------------------------------------------------------------------------- Kaldi tables + outputs
def test_kaldi_dir_round_trip_and_table_parsing(tmp_path):
    utts = synth.make_set(5, 9, 0.1, 0.3)
    keys = synth.write_kaldi_dir(tmp_path, utts, n_spk=2)
    scp = kaldi_io.read_table(tmp_path / "wav.scp")
    u2s = kaldi_io.read_table(tmp_path / "utt2spk")
    assert list(scp) == keys and set(u2s.values()) == {"spk00", "spk01"}
    (tmp_path / "t").write_text("a  x  extra\n\nb\ty\n")
    assert kaldi_io.read_table(tmp_path / "t") == {"a": "x", "b": "y"}
    (tmp_path / "bad").write_text("lonely\n")
    with pytest.raises(ValueError):
        kaldi_io.read_table(tmp_path / "bad")


def test_outputs_have_the_reference_format(tmp_path):
    """dict[str, list[float]] / dict[str, list[int]], readable with a plain torch.load as make_parquet_list.py:107-109
    does, floats holding exact fp32 values."""
    emb = np.random.default_rng(1).standard_normal((3, 192)).astype(np.float32)
    d = kaldi_io.embeddings_to_dict(["a", "b", "c"], emb)
    kaldi_io.save_atomic(d, tmp_path / "utt2embedding.pt")
    back = torch.load(tmp_path / "utt2embedding.pt")
    assert type(back) is dict and list(back) == ["a", "b", "c"]
    assert type(back["a"]) is list and len(back["a"]) == 192 and type(back["a"][0]) is float
    assert np.array_equal(np.array(back["b"], dtype=np.float32), emb[1])
    kaldi_io.save_atomic({"a": [1, 2, 3], "b": []}, tmp_path / "utt2speech_token.pt")
    tok = torch.load(tmp_path / "utt2speech_token.pt")
    assert tok == {"a": [1, 2, 3], "b": []} and type(tok["a"][0]) is int
    assert sorted(p.name for p in tmp_path.iterdir()) == ["utt2embedding.pt", "utt2speech_token.pt"]   # no temp left


def test_array_direct_writers_produce_what_torch_save_would(tmp_path):
    """save_embeddings_atomic / save_tokens_atomic (what the CLIs call at 10^5+ keys) lay the pickle out from the arrays;
    torch.load (both modes) and plain pickle must return the very dict torch.save(dict of lists) round-trips to."""
    import pickle
    import zipfile
    rng = np.random.default_rng(4)
    n = 20_011                                               # several chunks, last one ragged
    keys = [f"spk{i % 7:02d}_utt_{i:06d}" for i in range(n)]
    keys[3] = "方言_0003"                                     # non-ASCII key: BINUNICODE length is in bytes
    emb = rng.standard_normal((n, 192)).astype(np.float32)
    emb[0, :4] = [np.float32(1e-40), -0.0, np.float32(3.4e38), np.float32(-1.17549435e-38)]
    kaldi_io.save_embeddings_atomic(keys, emb, tmp_path / "utt2embedding.pt")
    kaldi_io.save_atomic(kaldi_io.embeddings_to_dict(keys, emb), tmp_path / "ref_e.pt")
    ref = torch.load(tmp_path / "ref_e.pt")
    for kw in ({}, {"weights_only": False}, {"weights_only": True}):
        got = torch.load(tmp_path / "utt2embedding.pt", **kw)
        assert type(got) is dict and list(got) == keys and got == ref
        assert type(got[keys[0]]) is list and type(got[keys[0]][0]) is float
    assert np.array_equal(np.array(got[keys[0]], dtype=np.float32).view(np.int32), emb[0].view(np.int32))   # -0.0 too
    counts = rng.integers(0, 300, n)
    counts[[0, 5, n - 1]] = 0                                # > 30 s utterances: key kept, empty list
    flat = rng.integers(0, 6561, int(counts.sum())).astype(np.int32)
    flat[:3] = [0, 255, 256]
    kaldi_io.save_tokens_atomic(keys, flat, counts, tmp_path / "utt2speech_token.pt")
    ref_t = dict(zip(keys, (flat[o - c:o].tolist() for o, c in zip(np.cumsum(counts), counts))))
    for kw in ({}, {"weights_only": False}):
        got = torch.load(tmp_path / "utt2speech_token.pt", **kw)
        assert list(got) == keys and got == ref_t and got[keys[0]] == [] and type(got[keys[1]][0]) is int
    with zipfile.ZipFile(tmp_path / "utt2speech_token.pt") as z:
        assert pickle.loads(z.read("utt2speech_token/data.pkl")) == ref_t
    kaldi_io.save_embeddings_atomic([], np.zeros((0, 192), np.float32), tmp_path / "empty.pt")
    assert torch.load(tmp_path / "empty.pt") == {}
    with pytest.raises(ValueError):
        kaldi_io.save_tokens_atomic(["a"], np.array([70000]), np.array([1]), tmp_path / "bad.pt")
    assert not any(p.name.startswith(".") for p in tmp_path.iterdir())                      # no temp left


def test_save_atomic_leaves_nothing_on_failure(tmp_path):
    class Unpicklable:
        def __reduce__(self):
            raise RuntimeError("nope")
    with pytest.raises(Exception):
        kaldi_io.save_atomic({"a": Unpicklable()}, tmp_path / "out.pt")
    assert list(tmp_path.iterdir()) == []


# ---------------------------------------------------------------------------------------- weights
def _pb_varint(v):
    out = b""
    while True:
        b = v & 0x7F
        v >>= 7
        out += bytes([b | (0x80 if v else 0)])
        if not v:
            return out


def _pb_field(num, wt, payload):
    return _pb_varint(num << 3 | wt) + (_pb_varint(len(payload)) + payload if wt == 2 else payload)


def test_onnx_initializer_reader(tmp_path):
    a = np.arange(24, dtype=np.float32).reshape(2, 3, 4)
    b = np.array([1.5, -2.5], dtype=np.float32)
    t_a = _pb_field(1, 2, b"".join(_pb_varint(d) for d in a.shape)) + _pb_field(2, 0, _pb_varint(1)) + \
        _pb_field(8, 2, b"conv.weight") + _pb_field(9, 2, a.tobytes())
    t_b = _pb_field(1, 0, _pb_varint(2)) + _pb_field(2, 0, _pb_varint(1)) + _pb_field(4, 2, b.tobytes()) + \
        _pb_field(8, 2, b"conv.bias")
    graph = _pb_field(2, 2, b"g") + _pb_field(5, 2, t_a) + _pb_field(5, 2, t_b)
    model = _pb_field(1, 0, _pb_varint(8)) + _pb_field(7, 2, graph)
    p = tmp_path / "m.onnx"
    p.write_bytes(model)
    st = weights.load_state(p)
    assert set(st) == {"conv.weight", "conv.bias"}
    assert np.array_equal(st["conv.weight"], a) and np.array_equal(st["conv.bias"], b)
    with pytest.raises(KeyError, match="lacks 1 of 2"):
        weights.as_torch_state(st, ["conv.weight", "missing.tensor"])
    (tmp_path / "empty.onnx").write_bytes(_pb_field(1, 0, _pb_varint(8)))
    with pytest.raises(ValueError):
        weights.load_state(tmp_path / "empty.onnx")


def test_state_file_formats(tmp_path):
    st = {"w": np.ones((2, 2), np.float32)}
    np.savez(tmp_path / "w.npz", **st)
    torch.save({k: torch.from_numpy(v) for k, v in st.items()}, tmp_path / "w.pt")
    for f in ("w.npz", "w.pt"):
        got = weights.load_state(tmp_path / f)
        assert np.array_equal(got["w"], st["w"])
    with pytest.raises(ValueError):
        weights.load_state(tmp_path / "w.bogus")


def test_synth_weights_match_oracle_modules():
    from oracle import campplus as CP, s3tokenizer as S3
    cam = CP.CamPlus()
    st = synth_weights.campplus_state(CamPlusConfig(), 3)
    want = {k: tuple(v.shape) for k, v in cam.state_dict().items() if "num_batches" not in k}
    assert {k: v.shape for k, v in st.items()} == want
    cfg = S3Config(n_layer=1)
    s3 = S3.S3Tokenizer(S3.S3Config(n_layer=1))
    st = synth_weights.s3tok_state(cfg, 3)
    assert {k: v.shape for k, v in st.items()} == {k: tuple(v.shape) for k, v in s3.state_dict().items()}
    assert synth_weights.s3tok_state(cfg, 3)["fc1.weight" if False else "blocks.0.fc1.weight"].dtype == np.float32


# ---------------------------------------------------------------------------------------- CLI without a GPU
@pytest.mark.parametrize("tool", ["extract_embedding.py", "extract_speech_token.py"])
def test_cli_fails_loudly_without_gpu_and_writes_nothing(tmp_path, tool):
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    synth.write_kaldi_dir(tmp_path, synth.make_set(2, 1, 0.2, 0.3))
    np.savez(tmp_path / "w.npz", w=np.zeros(1, np.float32))
    before = sorted(p.name for p in tmp_path.iterdir())
    r = subprocess.run([sys.executable, str(ROOT / "tools" / tool), "--dir", str(tmp_path), "--onnx_path",
                        str(tmp_path / "w.npz")], capture_output=True, text=True)
    assert r.returncode != 0
    assert "no CUDA device" in r.stderr or "FyError" in r.stderr
    assert sorted(p.name for p in tmp_path.iterdir()) == before


def test_run_tool_exit_paths(tmp_path):
    """cli.run_tool (entry of the two tool scripts): status of `main` becomes the process status, buffered output is
    flushed before os._exit, an exception still ends with a traceback and status 1, FY_FAST_EXIT=0 takes sys.exit."""
    prog = ("import sys; sys.path.insert(0, %r)\n"
            "from fangyan_tts_b200 import cli\n"
            "def main():\n"
            "    print('out-line'); sys.stderr.write('err-line\\n')\n"
            "    if sys.argv[1] == 'raise': raise RuntimeError('boom')\n"
            "    return int(sys.argv[1])\n"
            "cli.run_tool(main)\n") % str(ROOT)
    for env_extra in ({}, {"FY_FAST_EXIT": "0"}):
        env = dict(os.environ, **env_extra)
        ok = subprocess.run([sys.executable, "-c", prog, "0"], capture_output=True, text=True, env=env)
        assert ok.returncode == 0 and "out-line" in ok.stdout and "err-line" in ok.stderr
        bad = subprocess.run([sys.executable, "-c", prog, "3"], capture_output=True, text=True, env=env)
        assert bad.returncode == 3 and "out-line" in bad.stdout
        exc = subprocess.run([sys.executable, "-c", prog, "raise"], capture_output=True, text=True, env=env)
        assert exc.returncode == 1 and "RuntimeError: boom" in exc.stderr


def test_seconds_since_exec_is_this_process_age():
    import time
    from fangyan_tts_b200.cli import _seconds_since_exec
    a = _seconds_since_exec()
    time.sleep(0.3)
    b = _seconds_since_exec()
    assert 0.0 <= a < 3600.0 and 0.2 <= b - a <= 1.0
