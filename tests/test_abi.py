"""The C-ABI library loads without a GPU and exports every symbol include/ssi_b200.h declares."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "ssi_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ssi_b200_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_path():
    syms = declared_symbols()
    for s in ("ssi_b200_lmhead_ce_fwd", "ssi_b200_lmhead_ce_bwd", "ssi_b200_embedding_grad_scatter",
              "ssi_b200_label_shift_count", "ssi_b200_workspace_bytes", "ssi_b200_version", "ssi_b200_last_error"):
        assert s in syms


def test_library_exports_every_declared_symbol():
    from ssi_b200 import _lib, build

    build.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for s in declared_symbols():
        assert hasattr(lib, s), f"{s} declared in ssi_b200.h but not exported"
    assert set(declared_symbols()) == set(_lib.SYMBOLS), "ctypes binding and header disagree"
    assert _lib.load().ssi_b200_version() == _lib.ABI_VERSION


def test_host_side_argument_errors_need_no_gpu():
    from ssi_b200 import _lib

    lib = _lib.load()
    f = ctypes.c_size_t(0)
    assert lib.ssi_b200_workspace_bytes(0, 10, 64, ctypes.byref(f), None, None) < 0
    assert b"bad shape" in lib.ssi_b200_last_error()
    assert lib.ssi_b200_workspace_bytes(128, 1000, 100, ctypes.byref(f), None, None) < 0  # D % 64
    with pytest.raises(RuntimeError):
        _lib.workspace_bytes(128, 1000, 100)
    fwd, bwd, sc = _lib.workspace_bytes(32768, 130306, 2048)
    assert fwd > 0 and bwd >= 32768 * 256 * 2 and sc > 32768 * 4 * 4
    # null pointers are rejected before any CUDA call
    assert lib.ssi_b200_lmhead_ce_fwd(None, None, None, 1, 1, 64, -100, None, None, None, None, 0, None) < 0
