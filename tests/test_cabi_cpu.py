"""CPU checks of the C-ABI boundary: the library builds for sm_100a, loads, and exports exactly what
include/fangyan_b200.h declares.  No compute call is made here (there is no GPU in the build container)."""
import ctypes as C
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
HEADER = ROOT / "include" / "fangyan_b200.h"


@pytest.fixture(scope="module")
def lib():
    from fangyan_tts_b200 import _lib, build
    build.build()
    return _lib.load()


def _declared():
    text = re.sub(r"/\*.*?\*/", "", HEADER.read_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(fy_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported(lib):
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"


def test_ctypes_prototypes_cover_header():
    from fangyan_tts_b200._lib import PROTOTYPES
    assert sorted(PROTOTYPES) == _declared()


def test_header_cites_reference_call_sites():
    text = HEADER.read_text()
    for cite in ("tools/extract_embedding.py:29-33", "tools/extract_speech_token.py:38", "tools/extract_embedding.py:36-41",
                 "tools/extract_speech_token.py:40-48", "tools/extract_embedding.py:50-56"):
        assert cite in text


def test_version_and_error_strings(lib):
    assert b"sm_100a" in lib.fy_version()
    h = C.c_void_p()
    rc = lib.fy_create(0, 7, C.byref(h))          # bad precision is rejected before any CUDA call
    assert rc == -1 and b"precision" in lib.fy_last_error()
    assert lib.fy_create(0, 1, None) == -1


def test_create_fails_loudly_without_gpu(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    h = C.c_void_p()
    rc = lib.fy_create(0, 1, C.byref(h))
    assert rc < 0 and not h.value and len(lib.fy_last_error()) > 0


def test_engine_has_no_cpu_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from fangyan_tts_b200._lib import FyError
    from fangyan_tts_b200.engine import Engine
    with pytest.raises(FyError):
        Engine(0)


def test_sass_uses_tcgen05_and_tma():
    """The GEMM must be a Blackwell-native kernel: UTC*MMA (tcgen05.mma), LDTM (tcgen05.ld), UTMALDG (TMA)."""
    import shutil
    import subprocess
    from fangyan_tts_b200 import build
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    obj = build.OBJ / "fy_gemm.o"
    sass = subprocess.run(["cuobjdump", "-sass", str(obj)], capture_output=True, text=True).stdout
    for mnemonic in ("UTCHMMA", "LDTM", "UTMALDG"):
        assert mnemonic in sass, mnemonic
