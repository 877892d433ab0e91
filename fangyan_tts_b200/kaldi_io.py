"""Kaldi-style inputs and the three `.pt` outputs of stages 1-2 (contract B1).

Parsing mirrors tools/extract_embedding.py:69-76 and tools/extract_speech_token.py:69-72 of the reference: each line
is split on whitespace, field 0 is the key and field 1 the value.  Outputs are plain `torch.save`d dicts of Python
lists (extract_embedding.py:57-58, extract_speech_token.py:58) written atomically: a temp file in the same directory,
renamed only on success, because run.sh:25,35 skips a stage whenever its output file exists.
"""
from __future__ import annotations

import os
import struct
from pathlib import Path
from typing import Dict, List, Sequence, Tuple

import numpy as np


def read_table(path) -> Dict[str, str]:
    """`key<ws>value ...` per line -> ordered dict key -> second field (later duplicates win, as in the reference)."""
    out: Dict[str, str] = {}
    with open(path) as f:
        for ln, line in enumerate(f, 1):
            parts = line.replace("\n", "").split()
            if not parts:
                continue
            if len(parts) < 2:
                raise ValueError(f"{path}:{ln}: expected 'key value', got {line!r}")
            out[parts[0]] = parts[1]
    return out


def wav_header(path: str, key: str = "?") -> Tuple[int, int, int]:
    """(frames, channels, sample_rate) from the RIFF header only -- used to plan batches without decoding."""
    from .audio_io import AudioDecodeError
    try:
        with open(path, "rb") as f:
            head = f.read(4096)
            size = os.fstat(f.fileno()).st_size
    except OSError as e:
        raise AudioDecodeError(f"utterance {key}: cannot read {path}: {e}") from e
    if len(head) < 12 or head[:4] != b"RIFF" or head[8:12] != b"WAVE":
        raise AudioDecodeError(f"utterance {key}: {path} is not a RIFF/WAVE file")
    pos, fmt = 12, None
    while pos + 8 <= len(head):
        cid, csz = head[pos:pos + 4], struct.unpack_from("<I", head, pos + 4)[0]
        if cid == b"fmt ":
            fmt = struct.unpack_from("<HHIIHH", head, pos + 8)
        elif cid == b"data":
            if fmt is None:
                break
            frame = fmt[1] * fmt[5] // 8
            nbytes = min(csz, size - (pos + 8))
            return nbytes // max(frame, 1), fmt[1], fmt[2]
        pos += 8 + csz + (csz & 1)
    # header longer than 4 KiB (rare LIST chunks): fall back to a full parse
    from .audio_io import read_wav
    x, sr = read_wav(path, key)
    return x.shape[0], x.shape[1], sr


def save_atomic(obj, path) -> None:
    """torch.save to `path` via a same-directory temp file + rename; nothing is left behind on failure."""
    import torch
    path = Path(path)
    tmp = path.with_name(f".{path.name}.tmp.{os.getpid()}")
    try:
        torch.save(obj, tmp)
        os.replace(tmp, path)
    finally:
        if tmp.exists():
            tmp.unlink()


# ---- the same files without a Python object per value -----------------------------------------------------------
# torch.save of a dict of lists is a zip archive whose `data.pkl` is an ordinary protocol-2 pickle.  At 10^5..10^6 keys
# building the lists (tolist) and pickling ~200 values per key costs more than the GPU spends on the audio, so the two
# large outputs are written as the SAME byte-level format straight from the arrays: per key BINUNICODE, EMPTY_LIST, MARK,
# the values as fixed-width opcodes laid out by numpy (BINFLOAT = 'G' + big-endian float64 of the fp32 value, BININT2 =
# 'M' + uint16), APPENDS; one MARK .. SETITEMS around all pairs.  torch.load (weights_only or not) and pickle.load
# return exactly the dict the reference's loop would have saved (memo opcodes are optional in the format and omitted).
_ROWS_PER_CHUNK = 8192


def _key_op(k: str) -> bytes:
    b = k.encode("utf-8")
    return b"X" + struct.pack("<I", len(b)) + b


def _float_row_chunks(keys: Sequence[str], emb: np.ndarray):
    emb = np.ascontiguousarray(emb, dtype=np.float32)
    assert emb.ndim == 2 and emb.shape[0] == len(keys)
    n, d = emb.shape
    for a in range(0, n, _ROWS_PER_CHUNK):
        z = min(n, a + _ROWS_PER_CHUNK)
        body = np.empty((z - a, d, 9), dtype=np.uint8)
        body[:, :, 0] = ord("G")
        body[:, :, 1:] = emb[a:z].astype(">f8").view(np.uint8).reshape(z - a, d, 8)
        parts = []
        for i in range(a, z):
            parts.append(_key_op(keys[i]))
            if d:
                parts.append(b"](")
                parts.append(body[i - a].tobytes())
                parts.append(b"e")
            else:
                parts.append(b"]")
        yield b"".join(parts)


def _int_row_chunks(keys: Sequence[str], flat: np.ndarray, counts: np.ndarray):
    flat = np.ascontiguousarray(flat).astype(np.int64, copy=False).ravel()
    counts = np.asarray(counts, dtype=np.int64).ravel()
    assert counts.shape[0] == len(keys) and int(counts.sum()) == flat.shape[0]
    if flat.size and (int(flat.min()) < 0 or int(flat.max()) > 0xFFFF):
        raise ValueError("token ids outside [0, 65535] cannot be written as BININT2")
    ends = np.cumsum(counts)
    n = len(keys)
    for a in range(0, n, _ROWS_PER_CHUNK):
        z = min(n, a + _ROWS_PER_CHUNK)
        t0, t1 = int(ends[a] - counts[a]), int(ends[z - 1])
        body = np.empty((t1 - t0, 3), dtype=np.uint8)
        body[:, 0] = ord("M")
        body[:, 1] = flat[t0:t1] & 0xFF
        body[:, 2] = flat[t0:t1] >> 8
        raw = body.tobytes()
        parts, o = [], 0
        for i in range(a, z):
            c = int(counts[i])
            parts.append(_key_op(keys[i]))
            if c:
                parts.append(b"](")
                parts.append(raw[3 * o:3 * (o + c)])
                parts.append(b"e")
            else:
                parts.append(b"]")
            o += c
        yield b"".join(parts)


def _save_pickled_dict_atomic(path, chunks) -> None:
    """Write {data.pkl = PROTO 2, EMPTY_DICT, MARK, <chunks>, SETITEMS, STOP} in torch.save's zip layout, atomically."""
    import zipfile
    path = Path(path)
    tmp = path.with_name(f".{path.name}.tmp.{os.getpid()}")
    top = path.stem
    try:
        with zipfile.ZipFile(tmp, "w", zipfile.ZIP_STORED, allowZip64=True) as z:
            with z.open(f"{top}/data.pkl", "w", force_zip64=True) as f:
                f.write(b"\x80\x02}(")
                for c in chunks:
                    f.write(c)
                f.write(b"u.")
            z.writestr(f"{top}/.format_version", b"1")
            z.writestr(f"{top}/.storage_alignment", b"64")
            z.writestr(f"{top}/byteorder", b"little")
            z.writestr(f"{top}/version", b"3\n")
        os.replace(tmp, path)
    finally:
        if tmp.exists():
            tmp.unlink()


def save_embeddings_atomic(keys: Sequence[str], emb: np.ndarray, path) -> None:
    """utt2embedding.pt / spk2embedding.pt: {key: [192 floats]} (extract_embedding.py:57-58) from the [n, 192] array."""
    _save_pickled_dict_atomic(path, _float_row_chunks(keys, emb))


def save_tokens_atomic(keys: Sequence[str], flat: np.ndarray, counts: np.ndarray, path) -> None:
    """utt2speech_token.pt: {key: [ids]} (extract_speech_token.py:58) from the packed ids + per-key counts."""
    _save_pickled_dict_atomic(path, _int_row_chunks(keys, flat, counts))


def embeddings_to_dict(keys: Sequence[str], emb: np.ndarray) -> Dict[str, List[float]]:
    """[n, 192] float32 -> {key: list of Python floats holding the exact fp32 values} (extract_embedding.py:41)."""
    rows = np.asarray(emb, dtype=np.float32).tolist()
    return dict(zip(keys, rows))


def speaker_dict(spk_names: Sequence[str], mean: np.ndarray) -> Dict[str, List[float]]:
    return dict(zip(spk_names, np.asarray(mean, dtype=np.float32).tolist()))
