"""Row A1: audio decode for the drop-in CLIs (host side, stdlib + numpy only).

The reference calls torchaudio.load (tools/extract_embedding.py:26, tools/extract_speech_token.py:28), which
returns float32 [C, N] in [-1, 1).  Its realistic input is the 16 kHz mono PCM16 WAV that
prepare_training_data.py:96-117 writes with ffmpeg.  Here the file is parsed on the host and the samples stay
int16 until they are on the GPU (the 1/32768 scaling is fused into the first kernel); other PCM widths are
converted to float32 with torchaudio's normalisation constants.  Compressed formats (mp3/flac) are out of scope and
fail loudly with the utterance key.
"""
from __future__ import annotations

import struct

import numpy as np


class AudioDecodeError(RuntimeError):
    pass


def read_wav(path: str, key: str = "?"):
    """Returns (samples, sample_rate): int16 [N, C] for PCM16, else float32 [N, C] already in [-1, 1)."""
    try:
        with open(path, "rb") as f:
            data = f.read()
    except OSError as e:
        raise AudioDecodeError(f"utterance {key}: cannot read {path}: {e}") from e
    if len(data) < 12 or data[:4] != b"RIFF" or data[8:12] != b"WAVE":
        raise AudioDecodeError(f"utterance {key}: {path} is not a RIFF/WAVE file (mp3/flac input is not supported; "
                               "convert with prepare_training_data.py first)")
    pos, fmt, pcm = 12, None, None
    while pos + 8 <= len(data):
        cid, size = data[pos:pos + 4], struct.unpack_from("<I", data, pos + 4)[0]
        body = pos + 8
        if cid == b"fmt ":
            fmt = struct.unpack_from("<HHIIHH", data, body)
            if fmt[0] == 0xFFFE and size >= 26:  # WAVE_FORMAT_EXTENSIBLE: real tag is the sub-format GUID's first word
                fmt = (struct.unpack_from("<H", data, body + 24)[0],) + fmt[1:]
        elif cid == b"data":
            pcm = data[body:min(body + size, len(data))]
            break
        pos = body + size + (size & 1)
    if fmt is None or pcm is None:
        raise AudioDecodeError(f"utterance {key}: {path} has no fmt/data chunk")
    tag, ch, sr, _, _, bits = fmt
    if ch < 1:
        raise AudioDecodeError(f"utterance {key}: {path} has {ch} channels")
    frame = ch * bits // 8
    n = len(pcm) // frame
    raw = memoryview(pcm)[:n * frame]
    if tag == 1 and bits == 16:
        return np.frombuffer(raw, dtype="<i2").reshape(n, ch), sr
    if tag == 1 and bits == 32:
        return (np.frombuffer(raw, dtype="<i4").reshape(n, ch).astype(np.float32) / 2147483648.0), sr
    if tag == 1 and bits == 8:
        return ((np.frombuffer(raw, dtype=np.uint8).reshape(n, ch).astype(np.float32) - 128.0) / 128.0), sr
    if tag == 1 and bits == 24:
        b = np.frombuffer(raw, dtype=np.uint8).reshape(n, ch, 3).astype(np.int32)
        v = (b[..., 0] | (b[..., 1] << 8) | (b[..., 2] << 16))
        v = np.where(v >= 1 << 23, v - (1 << 24), v)
        return (v.astype(np.float32) / 8388608.0), sr
    if tag == 3 and bits == 32:
        return np.frombuffer(raw, dtype="<f4").reshape(n, ch).copy(), sr
    raise AudioDecodeError(f"utterance {key}: {path}: unsupported WAV encoding (format tag {tag}, {bits} bits)")


# ---------------------------------------------------------------------------------------------- native fast path (N2)
def scan_wavs(keys, paths, n_threads: int = 8):
    """Header pass over every file with the library's thread pool (fy_wav_scan): returns a dict of numpy arrays
    frames / channels / sample_rate / data_offset / simple (plain PCM16, readable straight into pinned memory by
    read_s16_into).  Files with long headers are re-parsed by read_wav; unreadable or non-WAV files raise
    AudioDecodeError naming the utterance key, as the per-file reader does."""
    import ctypes as C
    from . import _lib
    lib = _lib.load()
    n = len(paths)
    arr = (C.c_char_p * n)(*[p.encode() for p in paths])
    info = (_lib.FyWavInfo * n)()
    _lib.check(lib.fy_wav_scan(arr, n, info, max(1, n_threads)))
    a = np.frombuffer(info, dtype=np.dtype([("frames", "<i8"), ("data_offset", "<i8"), ("channels", "<i4"),
                                            ("sample_rate", "<i4"), ("format", "<i4"), ("bits", "<i4"),
                                            ("status", "<i4"), ("reserved", "<i4")]), count=n).copy()
    for i in np.nonzero(a["status"] != 0)[0].tolist():
        if a["status"][i] == -3:                     # long header: full parse in Python
            x, sr = read_wav(paths[i], keys[i])
            a["frames"][i], a["channels"][i], a["sample_rate"][i], a["format"][i], a["bits"][i] = x.shape[0], x.shape[1], sr, 0, 0
            continue
        if a["status"][i] == -1:
            raise AudioDecodeError(f"utterance {keys[i]}: cannot read {paths[i]}")
        raise AudioDecodeError(f"utterance {keys[i]}: {paths[i]} is not a RIFF/WAVE file (mp3/flac input is not supported; "
                               "convert with prepare_training_data.py first)")
    simple = (a["status"] == 0) & (a["format"] == 1) & (a["bits"] == 16)
    return {"frames": a["frames"], "channels": a["channels"], "sample_rate": a["sample_rate"],
            "data_offset": a["data_offset"], "simple": simple, "_paths": arr}


def read_s16_into(scan, indices, dst: np.ndarray, n_threads: int = 8, keys=None):
    """Samples of the `simple` mono files `indices` -> consecutive slices of the int16 array `dst` (e.g. a view of a
    pinned staging buffer), read by the library's threads with the GIL released."""
    import ctypes as C
    from . import _lib
    lib = _lib.load()
    n = len(indices)
    idx = np.asarray(indices, dtype=np.int64)
    cnt = np.ascontiguousarray(scan["frames"][idx] * scan["channels"][idx], dtype=np.int64)
    off = np.ascontiguousarray(scan["data_offset"][idx], dtype=np.int64)
    starts = np.zeros(n, dtype=np.int64)
    np.cumsum(cnt[:-1], out=starts[1:])
    if int(starts[-1] + cnt[-1] if n else 0) > dst.size:
        raise ValueError("destination too small")
    base = dst.ctypes.data
    ptrs = (C.c_void_p * n)(*[base + 2 * int(s) for s in starts])
    paths = (C.c_char_p * n)(*[scan["_paths"][int(i)] for i in idx])
    bad = C.c_int32(-1)
    rc = lib.fy_wav_read_s16(paths, off.ctypes.data_as(C.POINTER(C.c_int64)), cnt.ctypes.data_as(C.POINTER(C.c_int64)),
                             ptrs, n, max(1, n_threads), C.byref(bad))
    if rc != 0:
        k = keys[int(idx[bad.value])] if keys is not None and bad.value >= 0 else "?"
        raise AudioDecodeError(f"utterance {k}: {lib.fy_last_error().decode()}")
    return cnt
