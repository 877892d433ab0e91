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


def embeddings_to_dict(keys: Sequence[str], emb: np.ndarray) -> Dict[str, List[float]]:
    """[n, 192] float32 -> {key: list of Python floats holding the exact fp32 values} (extract_embedding.py:41)."""
    rows = np.asarray(emb, dtype=np.float32).tolist()
    return dict(zip(keys, rows))


def speaker_dict(spk_names: Sequence[str], mean: np.ndarray) -> Dict[str, List[float]]:
    return dict(zip(spk_names, np.asarray(mean, dtype=np.float32).tolist()))
