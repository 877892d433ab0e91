"""Weight loading for the drop-in CLIs: `--onnx_path` -> named fp32 tensors for fy_load_campplus / fy_load_s3tok.

The reference hands `--onnx_path` to onnxruntime.InferenceSession (tools/extract_embedding.py:94,
tools/extract_speech_token.py:90).  Neither `onnx` nor `onnxruntime` is installable offline, so this module reads
the ONNX protobuf directly (initialisers only; a ~100-line wire-format reader) and also accepts a torch state dict
(`.pt` / `.pth`) or an `.npz` with the tensor names of oracle/campplus.py and oracle/s3tokenizer.py.

UNVERIFIED: the real campplus.onnx / speech_tokenizer_v3.onnx are not available offline, so the mapping from their
initialiser names to ours cannot be checked.  Names are matched exactly first; a file whose names do not match fails
loudly listing what is missing instead of guessing.
"""
from __future__ import annotations

import struct
from pathlib import Path
from typing import Dict

import numpy as np

_ONNX_DTYPES = {1: np.float32, 10: np.float16, 11: np.float64, 6: np.int32, 7: np.int64}


def _varint(buf: memoryview, pos: int):
    val, shift = 0, 0
    while True:
        b = buf[pos]
        pos += 1
        val |= (b & 0x7F) << shift
        if not b & 0x80:
            return val, pos
        shift += 7


def _fields(buf: memoryview):
    """Yield (field_number, wire_type, value) over one protobuf message."""
    pos, n = 0, len(buf)
    while pos < n:
        key, pos = _varint(buf, pos)
        fnum, wt = key >> 3, key & 7
        if wt == 0:
            v, pos = _varint(buf, pos)
        elif wt == 1:
            v = buf[pos:pos + 8]
            pos += 8
        elif wt == 2:
            ln, pos = _varint(buf, pos)
            v = buf[pos:pos + ln]
            pos += ln
        elif wt == 5:
            v = buf[pos:pos + 4]
            pos += 4
        else:
            raise ValueError(f"unsupported protobuf wire type {wt}")
        yield fnum, wt, v


def _tensor(buf: memoryview):
    """onnx.TensorProto -> (name, ndarray) for the dtypes initialisers use."""
    dims, dtype, name, raw, floats, int64s = [], 1, "", None, None, None
    for f, wt, v in _fields(buf):
        if f == 1:      # dims (possibly packed)
            if wt == 0:
                dims.append(v)
            else:
                p = 0
                while p < len(v):
                    d, p = _varint(v, p)
                    dims.append(d)
        elif f == 2:
            dtype = v
        elif f == 8:
            name = bytes(v).decode()
        elif f == 9:
            raw = v
        elif f == 4:    # float_data, packed
            floats = np.frombuffer(v, dtype="<f4") if wt == 2 else np.array([struct.unpack("<f", v)[0]], np.float32)
        elif f == 7:    # int64_data, packed varints
            vals, p = [], 0
            while p < len(v):
                d, p = _varint(v, p)
                vals.append(d)
            int64s = np.array(vals, dtype=np.int64)
    if dtype not in _ONNX_DTYPES:
        return name, None
    if raw is not None:
        arr = np.frombuffer(raw, dtype=np.dtype(_ONNX_DTYPES[dtype]).newbyteorder("<"))
    elif floats is not None:
        arr = floats
    elif int64s is not None:
        arr = int64s
    else:
        arr = np.zeros(0, _ONNX_DTYPES[dtype])
    return name, arr.reshape(dims) if dims else arr


def read_onnx_initializers(path) -> Dict[str, np.ndarray]:
    """ModelProto.graph (field 7) -> GraphProto.initializer (field 5) -> TensorProto."""
    data = memoryview(Path(path).read_bytes())
    out: Dict[str, np.ndarray] = {}
    for f, wt, v in _fields(data):
        if f == 7 and wt == 2:
            for gf, gwt, gv in _fields(v):
                if gf == 5 and gwt == 2:
                    name, arr = _tensor(gv)
                    if arr is not None:
                        out[name] = arr
    if not out:
        raise ValueError(f"{path}: no initialisers found (not an ONNX model?)")
    return out


def load_state(path) -> Dict[str, np.ndarray]:
    """Named tensors from .onnx / .pt / .pth / .npz."""
    p = Path(path)
    suf = p.suffix.lower()
    if suf == ".onnx":
        return read_onnx_initializers(p)
    if suf == ".npz":
        with np.load(p) as z:
            return {k: z[k] for k in z.files}
    if suf in (".pt", ".pth", ".bin"):
        import torch
        sd = torch.load(p, map_location="cpu", weights_only=True)
        if isinstance(sd, dict) and "state_dict" in sd:
            sd = sd["state_dict"]
        return {k: v.detach().cpu().numpy() for k, v in sd.items()}
    raise ValueError(f"{path}: unknown weight file type '{suf}' (expected .onnx, .pt or .npz)")


def as_torch_state(state: Dict[str, np.ndarray], expected_names):
    """Check the names the loader needs are present and convert to torch tensors."""
    import torch
    missing = [n for n in expected_names if n not in state]
    if missing:
        raise KeyError(f"weight file lacks {len(missing)} of {len(expected_names)} expected tensors, e.g. {missing[:5]}; "
                       "the initialiser-name mapping for the real ONNX files is UNVERIFIED (see fangyan_tts_b200/weights.py)")
    return {n: torch.from_numpy(np.ascontiguousarray(state[n], dtype=np.float32)) for n in expected_names}
