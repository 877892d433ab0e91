"""fy_fbank_cmn_static: the fbank + CMN entry point that allocates nothing (scratch in a static __device__ array inside
the library, include/fangyan_b200.h).  It has no counterpart in the reference; the check is against the existing entry
point fy_fbank_cmn on the same inputs, through the C-ABI, on two calls in a row that share the static scratch.

Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

REL_TOL = 1e-4        # the front-end tolerance of tests/test_gpu_frontend.py; the two entry points run the same kernels
STATIC_BYTES = 4 << 20


def _kaldi_frames(n):
    return 0 if n < 400 else 1 + (n - 400) // 160


def _batch(seed, lengths):
    rng = np.random.default_rng(seed)
    return [(rng.standard_normal(n) * 6000.0).clip(-32768, 32767).astype(np.int16) for n in lengths]


def _both(engine, utts):
    """(feat through fy_fbank_cmn with a caller workspace, feat through fy_fbank_cmn_static, free bytes before / after
    the static call): raw C-ABI calls on device pointers, nothing created between the two memory readings."""
    import torch
    from fangyan_tts_b200._lib import check
    lib, dev = engine.lib, engine.device
    lens = [len(u) for u in utts]
    frames = [_kaldi_frames(n) for n in lens]
    pcm_off = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    fr_off = np.concatenate([[0], np.cumsum(frames)]).astype(np.int32)
    total = int(fr_off[-1])
    pcm = torch.from_numpy(np.concatenate(utts)).to(dev)
    d_po, d_fo = torch.from_numpy(pcm_off).to(dev), torch.from_numpy(fr_off).to(dev)
    nb = lib.fy_fbank_workspace_bytes(len(utts), total)
    assert nb <= STATIC_BYTES
    ws = torch.empty(nb, dtype=torch.uint8, device=dev)
    a = torch.full((total, 80), float("nan"), device=dev)
    b = torch.full((total, 80), float("nan"), device=dev)
    P = lambda t: C.c_void_p(t.data_ptr())
    check(lib.fy_fbank_cmn(engine.h, P(pcm), 0 if pcm.dtype == torch.int16 else 1, P(d_po), P(d_fo), len(utts), total,
                           P(a), P(ws), nb))
    torch.cuda.synchronize(dev)
    free0 = torch.cuda.mem_get_info(dev)[0]
    check(lib.fy_fbank_cmn_static(engine.h, P(pcm), 0 if pcm.dtype == torch.int16 else 1, P(d_po), P(d_fo), len(utts),
                                  total, P(b)))
    torch.cuda.synchronize(dev)
    free1 = torch.cuda.mem_get_info(dev)[0]
    return a, b, free0, free1


def _assert_same(a, b):
    import torch
    assert a.shape == b.shape and a.numel() > 0
    assert bool(torch.isfinite(b).all()), "the static-scratch call left frames unwritten"
    rel = float((a - b).abs().max() / a.abs().max())
    assert rel <= REL_TOL, f"static-scratch fbank differs from fy_fbank_cmn: {rel:.3e} relative"
    assert torch.equal(a, b), "same kernels and launch order: the results must be bit-identical"


def test_static_scratch_fbank_matches_workspace_fbank_twice(engine):
    # call 1: a batch that fills a good part of the scratch (40 utterances, ragged, one of them a single frame)
    big = _batch(7, [400, 16000, 48000, 79999, 160000, 161, 23456] + [16000 * (3 + i % 8) + 37 * i for i in range(33)])
    big = [u for u in big if len(u) >= 400]
    a1, b1, _, _ = _both(engine, big)
    _assert_same(a1, b1)
    # call 2, right after: a smaller batch, so most of the scratch still holds call 1's sums and means
    small = _batch(8, [32000, 401, 5000])
    a2, b2, free0, free1 = _both(engine, small)
    _assert_same(a2, b2)
    assert free1 == free0, f"device memory changed across fy_fbank_cmn_static ({free0} -> {free1} bytes free)"
    # and call 1's inputs again give call 1's result
    a3, b3, free0, free1 = _both(engine, big)
    _assert_same(a3, b3)
    import torch
    assert torch.equal(b3, b1) and free1 == free0


def test_static_scratch_limit_is_an_error(engine):
    from fangyan_tts_b200._lib import FyError, check
    lib = engine.lib
    import torch
    n_utts, total = 4, 4 * 60000                      # 240,000 frames need ~4.8 MB of scratch
    assert lib.fy_fbank_workspace_bytes(n_utts, total) > STATIC_BYTES
    dummy = torch.zeros(16, dtype=torch.int32, device=engine.device)
    P = C.c_void_p(dummy.data_ptr())
    rc = lib.fy_fbank_cmn_static(engine.h, P, 0, P, P, n_utts, total, P)     # rejected before anything is launched
    assert rc != 0
    with pytest.raises(FyError, match="static array holds"):
        check(rc)
