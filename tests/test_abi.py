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


def test_data_parallel_and_next_row_entry_points_validate_on_the_host():
    """Size queries and argument errors of the exchange and of the §8(f) rows need no GPU."""
    from ctypes import POINTER, byref, c_int64, c_size_t

    from ssi_b200 import _lib

    lib = _lib.load()
    nbytes, off = c_size_t(0), c_size_t(0)
    V, D = 133258, 2048
    R = 16384  # token rows per rank of c5 at 8 GPUs
    assert lib.ssi_b200_dp_buffer_bytes(V, D, 8, R, byref(nbytes), byref(off)) == 0
    slice_rows = -(-V // 8)
    up = lambda x: -(-x // 256) * 256  # noqa: E731
    # [ctrl][dW_acc fp32][stage bf16 slice][slot u64 V][ids i32 R][rows bf16 R x D]
    # [own_flag u32 ceil(V/8)][own_rows bf16 ceil(V/8) x D]  (two-phase sparse exchange)
    assert off.value == 4096 and nbytes.value == (4096 + up(V * D * 4) + up(slice_rows * D * 2) + up(V * 8)
                                                  + up(R * 4) + up(R * D * 2) + up(slice_rows * 4)
                                                  + up(slice_rows * D * 2))
    assert lib.ssi_b200_dp_buffer_bytes(V, D, 9, R, byref(nbytes), byref(off)) < 0  # one NVSwitch node: <= 8 ranks
    assert b"world size" in lib.ssi_b200_last_error()
    assert lib.ssi_b200_dp_exchange(None, 0, 2, V, D, R, 1, None, None, None, 0, None, None) < 0
    assert lib.ssi_b200_dp_counts(None, 0, 2, 1, None, None, None, None, None) < 0
    assert lib.ssi_b200_dp_embedding_exchange(None, 8, V, D, None, 0, None, 0, 2, R, 1, None, None, None) < 0
    assert lib.ssi_b200_grad_add(None, None, 8, None) < 0 and lib.ssi_b200_count_valid(None, 8, -100, None, None) < 0
    assert lib.ssi_b200_lmhead_ce_fwd_grad_ex(None, None, None, None, None, 1, 1, 64, -100, None, None, None, None,
                                              None, None, 0, None, None) < 0
    b = c_size_t(0)
    assert lib.ssi_b200_rmsnorm_bwd_workspace_bytes(32768, D, byref(b)) == 0 and b.value >= 148 * 2 * 8 * D * 4
    assert lib.ssi_b200_rmsnorm_fwd(None, None, 1e-6, 8, 64, None, None, None) < 0
    assert lib.ssi_b200_grad_scale_sqnorm_workspace_bytes(byref(b)) == 0 and b.value > 0
    lo = (c_int64 * 16)(*range(16))
    rc = lib.ssi_b200_label_shift_count_types(1, 1, 2, 8, -100, lo, lo, 16, 0, 1, 1, 1, None)  # 16 ranges > 15
    assert rc < 0 and b"n_ranges" in lib.ssi_b200_last_error()
    with pytest.raises(RuntimeError):
        _lib.set_option("no_such_option", 1)
    for name, val in (("lockstep_k1", 1), ("tail_split", 1), ("wide_dw", 2), ("wide_dx", 2), ("g_cap_mib", 4608)):
        _lib.set_option(name, val)
