"""Row N1 (stage 3 packing): the Arrow-direct writer against the reference's pandas job restated here as the checker
(tools/make_parquet_list.py:26-62 of the reference): same files, same schema, same values -- from the sidecars and
from the .pt fall-back."""
import json
import os
import wave

import numpy as np
import pyarrow.parquet as pq
import pytest
import torch

from fangyan_tts_b200 import kaldi_io, parquet_pack
from oracle import stage3


reference_job = stage3.job      # make_parquet_list.py:26-57 restated (oracle/stage3.py)


@pytest.fixture()
def corpus(tmp_path):
    rng = np.random.default_rng(3)
    src = tmp_path / "src"
    src.mkdir()
    keys = [f"utt{i:03d}" for i in range(23)]
    spk = {k: f"spk{i % 4}" for i, k in enumerate(keys)}
    with open(src / "wav.scp", "w") as f1, open(src / "text", "w") as f2, open(src / "utt2spk", "w") as f3, \
            open(src / "instruct", "w") as f4:
        for i, k in enumerate(keys):
            path = src / f"{k}.wav"
            with wave.open(str(path), "wb") as w:
                w.setnchannels(1); w.setsampwidth(2); w.setframerate(16000)
                w.writeframes(rng.integers(-3000, 3000, 800 + 37 * i, dtype=np.int16).tobytes())
            f1.write(f"{k} {path}\n")
            f2.write(f"{k} 你好 world {i}\n")
            f3.write(f"{k} {spk[k]}\n")
            f4.write(f"{k} 用粤语说 <|endofprompt|>\n")
    emb = rng.standard_normal((len(keys), 192)).astype(np.float32)
    spk_names = sorted(set(spk.values()))
    semb = rng.standard_normal((len(spk_names), 192)).astype(np.float32)
    toks = [None if i == 5 else rng.integers(0, 6561, 10 + 3 * i).astype(np.int32) for i in range(len(keys))]
    kaldi_io.save_atomic(kaldi_io.embeddings_to_dict(keys, emb), src / "utt2embedding.pt")
    kaldi_io.save_atomic(kaldi_io.speaker_dict(spk_names, semb), src / "spk2embedding.pt")
    kaldi_io.save_atomic({k: ([] if t is None else t.tolist()) for k, t in zip(keys, toks)}, src / "utt2speech_token.pt")
    return dict(src=src, keys=keys, emb=emb, spk_names=spk_names, semb=semb, toks=toks, tmp=tmp_path)


def check_against_reference(c, des, instruct):
    src = c["src"]
    t = lambda name: {l.split()[0]: " ".join(l.split()[1:]) for l in open(src / name)}
    utt2wav, utt2text, utt2spk = t("wav.scp"), t("text"), t("utt2spk")
    ref_dir = c["tmp"] / "ref"
    ref_dir.mkdir(exist_ok=True)
    e, s, k = (torch.load(src / n) for n in ("utt2embedding.pt", "spk2embedding.pt", "utt2speech_token.pt"))
    files = [l.strip() for l in open(des / "data.list")]
    assert [os.path.basename(f) for f in files] == [f"parquet_{i:09d}.tar" for i in range(3)]      # 23 utts / 10
    for i, f in enumerate(files):
        utts = c["keys"][10 * i:10 * i + 10]
        ref = ref_dir / f"ref_{i}.parquet"
        reference_job(utts, ref, utt2wav, utt2text, utt2spk, e, s, k, t("instruct") if instruct else None)
        a, b = pq.read_table(f), pq.read_table(ref).replace_schema_metadata(None)
        assert a.schema.names == b.schema.names
        for fa, fb in zip(a.schema, b.schema):
            # pandas >= 3 stores str columns as Arrow large_string, the pandas the reference pins as string: both are
            # the same parquet BYTE_ARRAY/UTF8 column, so they are compared as string; everything else must be equal
            ta = "string" if str(fa.type) in ("string", "large_string") else str(fa.type)
            tb = "string" if str(fb.type) in ("string", "large_string") else str(fb.type)
            assert ta == tb, (fa, fb)
        assert a.equals(b.cast(a.schema))
        pa_, pb = a.to_pandas(), b.to_pandas()
        assert list(pa_.columns) == list(pb.columns)
        for col in pa_.columns:
            for x, y in zip(pa_[col], pb[col]):
                assert np.array_equal(np.asarray(x), np.asarray(y))
        m = json.load(open(des / f"utt2parquet_{i:09d}.json"))
        assert list(m) == utts and set(m.values()) == {f}
        ms = json.load(open(des / f"spk2parquet_{i:09d}.json"))
        assert set(ms) == {utt2spk[u] for u in utts}
    assert len(open(des / "utt2data.list").read().split()) == len(open(des / "spk2data.list").read().split()) == 3


@pytest.mark.parametrize("instruct", [False, True])
def test_pt_fallback_matches_reference(corpus, instruct):
    des = corpus["tmp"] / "out_pt"
    parquet_pack.make_parquet_list(str(corpus["src"]), str(des), 10, 2, instruct=instruct)
    check_against_reference(corpus, des, instruct)


def test_sidecars_match_reference(corpus):
    src = corpus["src"]
    parquet_pack.write_embedding_sidecar(src / parquet_pack.EMB_SIDECAR, corpus["keys"], corpus["emb"])
    parquet_pack.write_embedding_sidecar(src / parquet_pack.SPK_SIDECAR, corpus["spk_names"], corpus["semb"])
    parquet_pack.write_token_sidecar(src / parquet_pack.TOK_SIDECAR, corpus["keys"], corpus["toks"])
    # poison the .pt files' availability: the sidecars alone must be enough for the numeric columns
    os.rename(src / "utt2embedding.pt", src / "keep_e.pt"); os.rename(src / "spk2embedding.pt", src / "keep_s.pt")
    os.rename(src / "utt2speech_token.pt", src / "keep_t.pt")
    des = corpus["tmp"] / "out_sc"
    assert parquet_pack.main(["--src_dir", str(src), "--des_dir", str(des), "--num_utts_per_parquet", "10",
                              "--num_processes", "3"]) == 0
    os.rename(src / "keep_e.pt", src / "utt2embedding.pt"); os.rename(src / "keep_s.pt", src / "spk2embedding.pt")
    os.rename(src / "keep_t.pt", src / "utt2speech_token.pt")
    check_against_reference(corpus, des, False)


def test_missing_embedding_fails(corpus):
    src = corpus["src"]
    parquet_pack.write_embedding_sidecar(src / parquet_pack.EMB_SIDECAR, corpus["keys"][:-1], corpus["emb"][:-1])
    os.utime(src / parquet_pack.EMB_SIDECAR, None)
    with pytest.raises(KeyError):
        parquet_pack.make_parquet_list(str(src), str(corpus["tmp"] / "bad"), 10, 1)
