"""CPU tests of the C-ABI boundary: the library builds for sm_100a, loads, and exports every
symbol include/b200_hotpath.h declares.  No compute call is made (no GPU here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib_path():
    from sglang_security_b200 import build

    return build.build()


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200_hotpath.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported(lib_path):
    lib = ctypes.CDLL(str(lib_path))
    declared = _declared_symbols()
    assert len(declared) >= 8
    missing = [s for s in declared if not hasattr(lib, s)]
    assert not missing, f"declared in the header but not exported: {missing}"


def test_python_binding_covers_header(lib_path):
    from sglang_security_b200 import _lib

    assert sorted(_lib.exported_symbols()) == _declared_symbols()
    lib = _lib.load()
    assert lib.b200_version() >= 1
    assert lib.b200_launch_count() >= 0


def test_host_side_helpers_need_no_gpu(lib_path):
    from sglang_security_b200 import _lib

    lib = _lib.load()
    # pure host functions: workspace sizing and the split heuristic
    assert lib.b200_attn_decode_workspace_bytes(64, 32, 128, 1) == 0
    assert lib.b200_attn_decode_workspace_bytes(64, 32, 128, 4) == 64 * 32 * 4 * (128 + 1) * 4
    s = lib.b200_attn_decode_num_splits(64, 8, 4, 128, 4096, 148)
    assert s == 1   # 512 CTAs fit the 592 resident slots: one balanced wave, no merge
    # per-rank head counts of TP 2 / 4 / 8: the largest split count that still fits one wave
    assert lib.b200_attn_decode_num_splits(64, 4, 4, 128, 4096, 148) == 2
    assert lib.b200_attn_decode_num_splits(64, 2, 4, 128, 4096, 148) == 4
    assert lib.b200_attn_decode_num_splits(64, 1, 4, 128, 4096, 148) == 9
    assert lib.b200_attn_decode_num_splits(1, 8, 4, 128, 100, 148) == 1
    assert lib.b200_attn_decode_split_align(128, 4) == 48  # 3 warps x 16-token tiles


def test_noalloc_decode_rejects_oversize_requests_before_touching_the_gpu(lib_path):
    """b200_attn_decode_noalloc: the size check against the built-in 8 MiB scratch comes before any
    CUDA call, so it can be exercised here (pointers are never dereferenced)."""
    from sglang_security_b200 import _lib

    lib = _lib.load()
    assert lib.b200_attn_decode_workspace_bytes(64, 32, 128, 8) > (8 << 20)
    fake = 0x1000
    rc = lib.b200_attn_decode_noalloc(fake, 4096, fake, fake, 1024, fake, 4096, fake, 64, fake, fake,
                                      64, 32, 8, 128, 128, 0.088, 0.0, _lib.BF16, 8, None)
    assert rc == 1
    msg = lib.b200_last_error().decode()
    assert "8388608" in msg and "scratch" in msg
    assert lib.b200_attn_decode_noalloc(fake, 4096, fake, fake, 1024, fake, 4096, fake, 64, fake, fake,
                                        0, 32, 8, 128, 128, 0.088, 0.0, _lib.BF16, 8, None) == 0  # empty batch


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from sglang_security_b200 import _lib

    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "_LIB_PATH", tmp_path / "nope.so")
    with pytest.raises(_lib.B200Error):
        _lib.load()
