"""Stage 3 fast path: the parquet shards of tools/make_parquet_list.py written from
Arrow arrays instead of a pandas frame of Python lists.

Same files, names and schema as the reference (make_parquet_list.py:26-62,117-141): `parquet_%09d.tar` holding columns
utt, wav, audio_data, text, spk, utt_embedding (list<double>), spk_embedding (list<double>), speech_token (list<int64>)
[, reject_speech_token][, instruct]; `utt2parquet_%09d.json`, `spk2parquet_%09d.json`; data.list / utt2data.list /
spk2data.list.  What changes is how the numeric columns get there: the stage-1/2 CLIs of this repo leave compact
sidecars next to the three `.pt` files (keys + one float32 matrix, keys + flat int32 ids + offsets) and this module
turns them into Arrow list arrays by buffer (no per-element Python objects).  Without sidecars it falls back to
`torch.load` of the `.pt` dicts, still skipping pandas.
"""
from __future__ import annotations

import json
import os
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path
from typing import Dict, List, Optional, Sequence

import numpy as np

from . import kaldi_io

EMB_SIDECAR = "utt2embedding.f32.npz"
SPK_SIDECAR = "spk2embedding.f32.npz"
TOK_SIDECAR = "utt2speech_token.i32.npz"


# ------------------------------------------------------------------------------------------------ sidecars
def _save_npz_atomic(path, **arrays) -> None:
    path = Path(path)
    tmp = path.with_name(f".{path.name}.tmp.{os.getpid()}.npz")
    try:
        np.savez(tmp, **arrays)
        os.replace(tmp, path)
    finally:
        if tmp.exists():
            tmp.unlink()


def write_embedding_sidecar(path, keys: Sequence[str], emb: np.ndarray) -> None:
    """keys[i] -> emb[i] (float32 [n, 192]); the same values utt2embedding.pt / spk2embedding.pt hold as lists."""
    _save_npz_atomic(path, keys=np.asarray(list(keys), dtype=np.str_), emb=np.ascontiguousarray(emb, dtype=np.float32))


def write_token_sidecar(path, keys: Sequence[str], tokens: Sequence[Optional[np.ndarray]]) -> None:
    """keys[i] -> tokens[i] (int ids; None or empty = the reference's [] for gated clips)."""
    counts = np.array([0 if t is None else len(t) for t in tokens], dtype=np.int64)
    off = np.zeros(len(counts) + 1, dtype=np.int64)
    np.cumsum(counts, out=off[1:])
    flat = np.zeros(int(off[-1]), dtype=np.int32)
    for i, t in enumerate(tokens):
        if counts[i]:
            flat[off[i]:off[i + 1]] = np.asarray(t, dtype=np.int32)
    _save_npz_atomic(path, keys=np.asarray(list(keys), dtype=np.str_), flat=flat, offsets=off)


def write_token_sidecar_flat(path, keys: Sequence[str], flat: np.ndarray, counts: np.ndarray) -> None:
    """The same file from already-packed ids (wav.scp order) + per-utterance counts: no per-utterance work."""
    off = np.zeros(len(counts) + 1, dtype=np.int64)
    np.cumsum(np.asarray(counts, dtype=np.int64), out=off[1:])
    _save_npz_atomic(path, keys=np.asarray(list(keys), dtype=np.str_), flat=np.asarray(flat, dtype=np.int32), offsets=off)


class _Rows:
    """key -> row view over a matrix or a ragged flat array, from a sidecar or from the reference's .pt dict."""

    def __init__(self, keys, values, offsets=None):
        self.index = {k: i for i, k in enumerate(keys)}
        self.values, self.offsets = values, offsets

    def gather(self, keys: Sequence[str], missing_ok: bool):
        """(flat values, int32 offsets[len(keys)+1]) for these keys, in order."""
        idx = np.fromiter((self.index.get(k, -1) for k in keys), dtype=np.int64, count=len(keys))
        if not missing_ok and (idx < 0).any():
            raise KeyError(keys[int(np.argmax(idx < 0))])
        if self.offsets is None:
            rows = self.values[np.maximum(idx, 0)]
            w = rows.shape[1]
            return rows.reshape(-1), np.arange(len(keys) + 1, dtype=np.int32) * w
        lo, hi = self.offsets[np.maximum(idx, 0)], self.offsets[np.maximum(idx, 0) + 1]
        n = np.where(idx >= 0, hi - lo, 0)
        off = np.zeros(len(keys) + 1, dtype=np.int64)
        np.cumsum(n, out=off[1:])
        take = np.repeat(lo - off[:-1], n) + np.arange(int(off[-1]), dtype=np.int64)
        return self.values[take], off.astype(np.int32)


def _load_matrix(src_dir, sidecar: str, pt: str) -> _Rows:
    p = Path(src_dir) / sidecar
    q = Path(src_dir) / pt
    if p.exists() and (not q.exists() or p.stat().st_mtime >= q.stat().st_mtime - 1.0):
        z = np.load(p)
        return _Rows([str(k) for k in z["keys"]], z["emb"])
    import torch
    d = torch.load(q)
    keys = list(d.keys())
    return _Rows(keys, np.asarray([d[k] for k in keys], dtype=np.float64).reshape(len(keys), -1))


def _load_tokens(src_dir, sidecar: str = TOK_SIDECAR, pt: str = "utt2speech_token.pt") -> _Rows:
    p = Path(src_dir) / sidecar
    q = Path(src_dir) / pt
    if p.exists() and (not q.exists() or p.stat().st_mtime >= q.stat().st_mtime - 1.0):
        z = np.load(p)
        return _Rows([str(k) for k in z["keys"]], z["flat"], z["offsets"])
    import torch
    d = torch.load(q)
    keys = list(d.keys())
    counts = np.array([len(d[k]) for k in keys], dtype=np.int64)
    off = np.zeros(len(keys) + 1, dtype=np.int64)
    np.cumsum(counts, out=off[1:])
    flat = np.fromiter((t for k in keys for t in d[k]), dtype=np.int64, count=int(off[-1]))
    return _Rows(keys, flat, off)


# ------------------------------------------------------------------------------------------------ one shard
def _list_array(values: np.ndarray, offsets: np.ndarray, dtype):
    import pyarrow as pa
    return pa.ListArray.from_arrays(pa.array(offsets, type=pa.int32()), pa.array(np.asarray(values, dtype=dtype)))


def write_shard(utts: Sequence[str], utt2wav: Dict[str, str], utt2text: Dict[str, str], utt2spk: Dict[str, str],
                utt_emb: _Rows, spk_emb: _Rows, tokens: _Rows, parquet_file, utt2parquet_file, spk2parquet_file,
                reject: Optional[_Rows] = None, utt2instruct: Optional[Dict[str, str]] = None) -> None:
    """One `job` of make_parquet_list.py:26-62."""
    import pyarrow as pa
    import pyarrow.parquet as pq
    utts = list(utts)
    data = []
    for u in utts:
        with open(utt2wav[u], "rb") as f:
            data.append(f.read())
    spks = [utt2spk[u] for u in utts]
    cols = {
        "utt": pa.array(utts, type=pa.string()),
        "wav": pa.array([utt2wav[u] for u in utts], type=pa.string()),
        "audio_data": pa.array(data, type=pa.binary()),
        "text": pa.array([utt2text[u] for u in utts], type=pa.string()),
        "spk": pa.array(spks, type=pa.string()),
        "utt_embedding": _list_array(*utt_emb.gather(utts, False), np.float64),
        "spk_embedding": _list_array(*spk_emb.gather(spks, False), np.float64),
        "speech_token": _list_array(*tokens.gather(utts, True), np.int64),      # .get(utt, []) in the reference
    }
    if reject is not None:
        cols["reject_speech_token"] = _list_array(*reject.gather(utts, False), np.int64)
    if utt2instruct is not None:
        cols["instruct"] = pa.array([utt2instruct[u] for u in utts], type=pa.string())
    pq.write_table(pa.table(cols), parquet_file)
    with open(utt2parquet_file, "w") as f:
        json.dump({k: str(parquet_file) for k in utts}, f, ensure_ascii=False, indent=2)
    with open(spk2parquet_file, "w") as f:
        json.dump({k: str(parquet_file) for k in list(set(spks))}, f, ensure_ascii=False, indent=2)


# ------------------------------------------------------------------------------------------------ whole directory
def _read_text_table(path) -> Dict[str, str]:
    out = {}
    with open(path) as f:
        for line in f:
            parts = line.replace("\n", "").split()
            out[parts[0]] = " ".join(parts[1:])
    return out


def make_parquet_list(src_dir, des_dir, num_utts_per_parquet: int = 1000, num_processes: int = 1,
                      instruct: bool = False, dpo: bool = False) -> List[str]:
    """The `__main__` of make_parquet_list.py:65-141.  Returns the parquet file names, in order."""
    utt2wav = kaldi_io.read_table(os.path.join(src_dir, "wav.scp"))
    utt2text = _read_text_table(os.path.join(src_dir, "text"))
    utt2spk = kaldi_io.read_table(os.path.join(src_dir, "utt2spk"))
    utt2instruct = _read_text_table(os.path.join(src_dir, "instruct")) if instruct else None
    utt_emb = _load_matrix(src_dir, EMB_SIDECAR, "utt2embedding.pt")
    spk_emb = _load_matrix(src_dir, SPK_SIDECAR, "spk2embedding.pt")
    tokens = _load_tokens(src_dir)
    reject = _load_tokens(f"{src_dir}_reject") if dpo else None
    utts = list(utt2wav.keys())
    os.makedirs(des_dir, exist_ok=True)
    names = [[], [], []]
    jobs = []
    for i, j in enumerate(range(0, len(utts), num_utts_per_parquet)):
        files = [os.path.join(des_dir, f"{stem}_{i:09d}.{ext}") for stem, ext in
                 (("parquet", "tar"), ("utt2parquet", "json"), ("spk2parquet", "json"))]
        for lst, name in zip(names, files):
            lst.append(name)
        jobs.append((utts[j:j + num_utts_per_parquet], *files))
    with ThreadPoolExecutor(max_workers=max(1, num_processes)) as pool:
        futs = [pool.submit(write_shard, u, utt2wav, utt2text, utt2spk, utt_emb, spk_emb, tokens, pf, uf, sf, reject,
                            utt2instruct) for u, pf, uf, sf in jobs]
        for f in futs:
            f.result()          # unlike apply_async in the reference, a failing shard fails the run
    for lst, fname in zip(names, ("data.list", "utt2data.list", "spk2data.list")):
        with open(os.path.join(des_dir, fname), "w", encoding="utf8") as f:
            for name in lst:
                f.write(name + "\n")
    return names[0]


def main(argv=None) -> int:
    import argparse
    ap = argparse.ArgumentParser(prog="make_parquet_list.py")
    ap.add_argument("--num_utts_per_parquet", type=int, default=1000, help="num utts per parquet")
    ap.add_argument("--num_processes", type=int, default=1, help="num workers for make parquets")
    ap.add_argument("--instruct", action="store_true", default=False, help="has instruct file or not")
    ap.add_argument("--src_dir", type=str)
    ap.add_argument("--des_dir", type=str)
    ap.add_argument("--dpo", action="store_true", default=False, help="Use Direct Preference Optimization")
    a = ap.parse_args(argv)
    make_parquet_list(a.src_dir, a.des_dir, a.num_utts_per_parquet, a.num_processes, a.instruct, a.dpo)
    return 0
