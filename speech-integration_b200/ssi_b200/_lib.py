"""ctypes binding of ``libssi_b200.so`` (C ABI declared in ``include/ssi_b200.h``).

The library is built in-tree by ``ssi_b200.build.build()`` (nvcc, sm_100a only).  There is deliberately no
fallback: if the shared object is missing or fails to load, every op raises ``RuntimeError``.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_int, c_int64, c_size_t, c_void_p, POINTER

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libssi_b200.so")
ABI_VERSION = 114

# name -> (restype, argtypes); must list every symbol include/ssi_b200.h declares
SYMBOLS = {
    "ssi_b200_version": (c_int, []),
    "ssi_b200_last_error": (c_char_p, []),
    "ssi_b200_launch_count": (c_int64, []),
    "ssi_b200_set_option": (c_int, [c_char_p, c_int64]),
    "ssi_b200_get_option": (c_int, [ctypes.c_char_p, POINTER(c_int64)]),
    "ssi_b200_workspace_bytes": (
        c_int,
        [c_int64, c_int64, c_int64, POINTER(c_size_t), POINTER(c_size_t), POINTER(c_size_t)],
    ),
    "ssi_b200_plan_info": (c_int, [c_int64, c_int64, c_int64, c_void_p]),
    "ssi_b200_lmhead_ce_fwd": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_void_p,
         c_size_t, c_void_p],
    ),
    "ssi_b200_lmhead_ce_bwd": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_void_p,
         c_void_p, c_int, c_void_p, c_size_t, c_void_p],
    ),
    "ssi_b200_single_pass_workspace_bytes": (c_int, [c_int64, c_int64, c_int64, POINTER(c_size_t)]),
    "ssi_b200_lmhead_ce_fwd_grad": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_void_p, c_void_p,
         c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "ssi_b200_lmhead_ce_fwd_grad_dp": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_void_p, c_void_p,
         c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p, c_int, c_void_p, c_void_p],
    ),
    "ssi_b200_lmhead_ce_fwd_grad_norm": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_void_p, c_void_p,
         c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p, c_int, c_void_p, c_void_p, c_void_p,
         ctypes.c_float, c_void_p],
    ),
    "ssi_b200_lmhead_ce_fwd_grad_ex": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_void_p, c_void_p,
         c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p, c_void_p],
    ),
    "ssi_b200_single_pass_lockstep_status": (c_int, [c_void_p, c_int64, c_int64, c_int64, POINTER(c_int)]),
    "ssi_b200_lmhead_ce_apply_grad": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_void_p, c_int, c_void_p],
    ),
    "ssi_b200_embedding_grad_scatter": (
        c_int,
        [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "ssi_b200_embedding_grad_scatter_f32": (
        c_int,
        [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "ssi_b200_embedding_grad_sort": (c_int, [c_void_p, c_int64, c_int64, c_int64, c_void_p, c_size_t, c_void_p]),
    "ssi_b200_embedding_grad_scatter_sorted_f32": (
        c_int,
        [c_void_p, c_int64, c_int64, c_int64, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "ssi_b200_embedding_grad_scatter_sorted": (
        c_int,
        [c_void_p, c_int64, c_int64, c_int64, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "ssi_b200_dp_buffer_bytes": (c_int, [c_int64, c_int64, c_int, c_int64, POINTER(c_size_t), POINTER(c_size_t)]),
    "ssi_b200_dp_alloc": (c_int, [c_size_t, POINTER(c_void_p), c_void_p]),
    "ssi_b200_dp_open": (c_int, [c_void_p, POINTER(c_void_p)]),
    "ssi_b200_dp_close": (c_int, [c_void_p]),
    "ssi_b200_dp_free": (c_int, [c_void_p]),
    "ssi_b200_dp_status": (c_int, [c_void_p, POINTER(c_int)]),
    "ssi_b200_dp_reset": (c_int, [c_void_p]),
    "ssi_b200_dp_counts": (
        c_int,
        [POINTER(c_void_p), c_int, c_int, ctypes.c_uint32, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "ssi_b200_dp_exchange": (
        c_int,
        [POINTER(c_void_p), c_int, c_int, c_int64, c_int64, c_int64, ctypes.c_uint32, c_void_p, c_void_p, c_void_p,
         c_int, c_void_p, c_void_p],
    ),
    "ssi_b200_dp_embedding_exchange": (
        c_int,
        [c_void_p, c_int64, c_int64, c_int64, c_void_p, c_size_t, POINTER(c_void_p), c_int, c_int, c_int64,
         ctypes.c_uint32, c_void_p, c_void_p, c_void_p],
    ),
    "ssi_b200_grad_add": (c_int, [c_void_p, c_void_p, c_int64, c_void_p]),
    "ssi_b200_label_shift_count": (c_int, [c_void_p, c_int64, c_int64, c_int64, c_void_p, c_void_p, c_void_p]),
    "ssi_b200_count_valid": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_void_p]),
    "ssi_b200_rmsnorm_fwd": (c_int, [c_void_p, c_void_p, ctypes.c_float, c_int64, c_int64, c_void_p, c_void_p, c_void_p]),
    "ssi_b200_rmsnorm_bwd_workspace_bytes": (c_int, [c_int64, c_int64, POINTER(c_size_t)]),
    "ssi_b200_rmsnorm_bwd": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "ssi_b200_rmsnorm_bwd_static": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p],
    ),
    "ssi_b200_label_shift_count_types": (
        c_int,
        [c_void_p, c_void_p, c_int64, c_int64, c_int64, POINTER(c_int64), POINTER(c_int64), c_int, c_int64, c_void_p,
         c_void_p, c_void_p, c_void_p],
    ),
    "ssi_b200_grad_scale_sqnorm_workspace_bytes": (c_int, [POINTER(c_size_t)]),
    "ssi_b200_grad_scale_sqnorm": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "ssi_b200_grad_fold_workspace_bytes": (c_int, [c_int64, c_int64, POINTER(c_size_t)]),
    "ssi_b200_grad_fold_sqnorm": (c_int, [c_void_p, c_size_t, c_void_p, c_void_p]),
    "ssi_b200_lmhead_ce_apply_grad_fold": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_void_p, c_int, c_void_p, c_void_p,
         c_size_t, c_void_p],
    ),
    "ssi_b200_embedding_grad_scatter_sorted_fold": (
        c_int,
        [c_void_p, c_int64, c_int64, c_int64, c_void_p, c_void_p, c_size_t, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "ssi_b200_dp_exchange_fold": (
        c_int,
        [POINTER(c_void_p), c_int, c_int, c_int64, c_int64, c_int64, ctypes.c_uint32, c_void_p, c_void_p, c_void_p,
         c_int, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "ssi_b200_dp_exchange_nvls": (
        c_int,
        [POINTER(c_void_p), c_void_p, c_int, c_int, c_int64, c_int64, c_int64, ctypes.c_uint32, c_void_p, c_void_p,
         c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "ssi_b200_dp_embedding_exchange_fold": (
        c_int,
        [c_void_p, c_int64, c_int64, c_int64, c_void_p, c_size_t, POINTER(c_void_p), c_int, c_int, c_int64,
         ctypes.c_uint32, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "ssi_b200_embedding_gather": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_void_p, c_void_p]),
    "ssi_b200_profile_begin": (c_int, []),
    "ssi_b200_profile_end": (c_int, [c_void_p, c_void_p]),
    "ssi_b200_debug_gemm": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int, c_int, c_int, c_void_p],
    ),
}



class FwdGradOpts(ctypes.Structure):
    """``ssi_b200_fwd_grad_opts`` of include/ssi_b200.h."""

    _fields_ = [
        ("accumulate_dW_acc", c_int),
        ("dw_acc_init_event", c_void_p),
        ("dw_ready_event", c_void_p),
        ("pre_grad_event", c_void_p),
        ("norm_scale", c_void_p),
        ("norm_eps", ctypes.c_float),
        ("dscale", c_void_p),
    ]


_lib = None


def load() -> ctypes.CDLL:
    """Load the shared library (once). Raises RuntimeError if it is missing or has the wrong ABI."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"ssi_b200: CUDA extension {LIB_PATH} is not built. Run `python -c 'import __graft_entry__ as g; "
            "g.build()'` (nvcc, sm_100a). There is no CPU or PyTorch fallback for this path."
        )
    try:
        lib = ctypes.CDLL(LIB_PATH)
    except OSError as e:  # pragma: no cover - depends on the box
        raise RuntimeError(f"ssi_b200: failed to load {LIB_PATH}: {e}") from e
    for name, (restype, argtypes) in SYMBOLS.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:
            raise RuntimeError(f"ssi_b200: {LIB_PATH} does not export {name}") from e
        fn.restype = restype
        fn.argtypes = argtypes
    v = lib.ssi_b200_version()
    if v != ABI_VERSION:
        raise RuntimeError(f"ssi_b200: ABI version mismatch: library {v}, binding {ABI_VERSION}")
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().ssi_b200_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"ssi_b200.{what} failed (rc={rc}): {msg}")


def workspace_bytes(N: int, V: int, D: int, which: str = "all") -> tuple[int, int, int]:
    """(forward, backward, scatter) workspace sizes in bytes; ``which`` = "all" | "lmhead" | "scatter"."""
    f, b, s = c_size_t(0), c_size_t(0), c_size_t(0)
    pf = ctypes.byref(f) if which in ("all", "lmhead") else None
    pb = ctypes.byref(b) if which in ("all", "lmhead") else None
    ps = ctypes.byref(s) if which in ("all", "scatter") else None
    check(load().ssi_b200_workspace_bytes(N, V, D, pf, pb, ps), "workspace_bytes")
    return f.value, b.value, s.value


def single_pass_workspace_bytes(N: int, V: int, D: int) -> int:
    b = c_size_t(0)
    check(load().ssi_b200_single_pass_workspace_bytes(N, V, D, ctypes.byref(b)), "single_pass_workspace_bytes")
    return b.value


def set_option(name: str, value: int) -> None:
    check(load().ssi_b200_set_option(name.encode(), int(value)), f"set_option({name})")


def get_option(name: str) -> int:
    v = c_int64(0)
    check(load().ssi_b200_get_option(name.encode(), ctypes.byref(v)), f"get_option({name})")
    return int(v.value)


def launch_count() -> int:
    return int(load().ssi_b200_launch_count())


PROFILE_CLASSES = ("lse_fwd", "dlogits", "dW", "dX", "scatter", "aux", "softmax_grad", "exchange")


def profile_begin() -> None:
    check(load().ssi_b200_profile_begin(), "profile_begin")


def profile_end() -> dict[str, tuple[float, int]]:
    """{class: (total ms, launches)} for the launches made since profile_begin (synchronises)."""
    ms = (ctypes.c_double * 8)()
    cnt = (ctypes.c_int64 * 8)()
    check(load().ssi_b200_profile_end(ctypes.cast(ms, c_void_p), ctypes.cast(cnt, c_void_p)), "profile_end")
    return {PROFILE_CLASSES[i]: (float(ms[i]), int(cnt[i])) for i in range(8)}


def plan_info(N: int, V: int, D: int) -> dict[str, int]:
    out = (ctypes.c_int64 * 8)()
    check(load().ssi_b200_plan_info(N, V, D, ctypes.cast(out, c_void_p)), "plan_info")
    keys = ("fwd_splits", "vtiles_per_split", "tiles_m", "vtiles_total", "bwd_blocks", "vtiles_per_block",
            "block_cols", "rows_padded")
    return {k: int(out[i]) for i, k in enumerate(keys)}
