"""Thin tensor-level wrappers over the C ABI: argument checking, workspace reuse, current-stream plumbing.

PyTorch is used only for device memory and streams; all arithmetic happens in libssi_b200.so.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib

_WS: dict[tuple, torch.Tensor] = {}


def _workspace(kind: str, nbytes: int, device: torch.device) -> torch.Tensor:
    """Grow-only per-(kind, device, stream) scratch buffer; the kernels of one call only use it in stream order."""
    key = (kind, device.index, torch.cuda.current_stream(device).cuda_stream)
    buf = _WS.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=device)
        _WS[key] = buf
    return buf


def free_workspaces() -> None:
    _WS.clear()


def single_pass_status(N: int, V: int, D: int, device: torch.device | None = None) -> int:
    """Status bits of the last :func:`lmhead_ce_fwd_grad` call of this shape on the current stream (synchronises;
    a debug query): bit 0 = a lockstep wait gave up (slower, still correct), bit 1 = the exp-stash form overflowed
    on some row (a logit > ~88 nats above the row's first 256 vocabulary columns): loss inf, gradients NaN -- rerun
    with ``_lib.set_option("exp_stash", 0)``."""
    import ctypes

    device = torch.device("cuda", torch.cuda.current_device()) if device is None else device
    ws = _WS.get(("single_pass", device.index, torch.cuda.current_stream(device).cuda_stream))
    if ws is None:
        raise RuntimeError("ssi_b200: no single-pass call has run on this device and stream yet")
    out = ctypes.c_int(0)
    with torch.cuda.device(device):
        rc = _lib.load().ssi_b200_single_pass_lockstep_status(ws.data_ptr(), N, V, D, ctypes.byref(out))
    _lib.check(rc, "single_pass_lockstep_status")
    return out.value


def _stream(device: torch.device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def _require_cuda(*ts: torch.Tensor) -> torch.device:
    dev = ts[0].device
    for t in ts:
        if t.device.type != "cuda":
            raise RuntimeError(
                "ssi_b200: tensors must live on a CUDA (sm_100a) device; there is no CPU fallback for this path"
            )
        if t.device != dev:
            raise RuntimeError("ssi_b200: all tensors must be on the same device")
    return dev


def _check_xw(X: torch.Tensor, W: torch.Tensor) -> tuple[int, int, int]:
    if X.dtype != torch.bfloat16 or W.dtype != torch.bfloat16:
        raise TypeError(f"ssi_b200: X and W must be bfloat16 (got {X.dtype}, {W.dtype})")
    if X.dim() != 2 or W.dim() != 2 or X.size(1) != W.size(1):
        raise ValueError(f"ssi_b200: expected X [N, D] and W [V, D], got {tuple(X.shape)} and {tuple(W.shape)}")
    if not X.is_contiguous() or not W.is_contiguous():
        raise ValueError("ssi_b200: X and W must be contiguous")
    return X.size(0), W.size(0), X.size(1)


def lmhead_ce_fwd(X: torch.Tensor, W: torch.Tensor, labels: torch.Tensor, ignore_index: int = -100):
    """Fused ``X @ W.T`` + cross entropy forward.

    Returns ``(loss_out, lse, n_valid)``: ``loss_out`` fp32 ``[2]`` = {sum of per-row losses, mean over valid
    rows}, ``lse`` fp32 ``[N]``, ``n_valid`` int64 ``[1]``.  Replaces chunked_output + CEWithChunkedOutputLoss
    (tt/torchtune/modules/transformer.py:499-502, tt/torchtune/modules/loss/ce_chunked_output_loss.py:45-83).
    """
    dev = _require_cuda(X, W, labels)
    N, V, D = _check_xw(X, W)
    if labels.dtype != torch.int64 or labels.numel() != N or not labels.is_contiguous():
        raise ValueError("ssi_b200: labels must be a contiguous int64 tensor with one entry per row of X")
    lib = _lib.load()
    fwd_b, _, _ = _lib.workspace_bytes(N, V, D, "lmhead")
    ws = _workspace("fwd", fwd_b, dev)
    lse = torch.empty(N, dtype=torch.float32, device=dev)
    loss_out = torch.empty(2, dtype=torch.float32, device=dev)
    n_valid = torch.empty(1, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        rc = lib.ssi_b200_lmhead_ce_fwd(X.data_ptr(), W.data_ptr(), labels.data_ptr(), N, V, D, int(ignore_index),
                                        lse.data_ptr(), loss_out.data_ptr(), n_valid.data_ptr(), ws.data_ptr(),
                                        ws.numel(), _stream(dev))
    _lib.check(rc, "lmhead_ce_fwd")
    return loss_out, lse, n_valid


def lmhead_ce_bwd(X: torch.Tensor, W: torch.Tensor, labels: torch.Tensor, lse: torch.Tensor,
                  grad_out: torch.Tensor, n_valid_total: torch.Tensor, ignore_index: int = -100,
                  dW: torch.Tensor | None = None, accumulate_dW: bool = False):
    """Backward of :func:`lmhead_ce_fwd`: returns ``(dX, dW)`` (bf16).

    ``grad_out``: fp32 device scalar (upstream gradient of the mean loss); ``n_valid_total``: int64 device
    scalar used as the mean's divisor.  If ``dW`` is given it is written in place (added to when
    ``accumulate_dW``), which is how the tied weight's ``.grad`` buffer is shared with the embedding gradient.
    """
    dev = _require_cuda(X, W, labels, lse, grad_out, n_valid_total)
    N, V, D = _check_xw(X, W)
    if grad_out.dtype != torch.float32 or grad_out.numel() != 1:
        raise ValueError("ssi_b200: grad_out must be a single fp32 value on the device")
    if n_valid_total.dtype != torch.int64 or n_valid_total.numel() != 1:
        raise ValueError("ssi_b200: n_valid_total must be a single int64 value on the device")
    if lse.dtype != torch.float32 or lse.numel() != N:
        raise ValueError("ssi_b200: lse must be fp32 [N]")
    if dW is None:
        if accumulate_dW:
            raise ValueError("ssi_b200: accumulate_dW needs an existing dW buffer")
        dW = torch.empty_like(W)
    elif dW.shape != W.shape or dW.dtype != torch.bfloat16 or not dW.is_contiguous() or dW.device != dev:
        raise ValueError("ssi_b200: dW must be a contiguous bf16 tensor shaped like W on the same device")
    lib = _lib.load()
    _, bwd_b, _ = _lib.workspace_bytes(N, V, D, "lmhead")
    ws = _workspace("bwd", bwd_b, dev)
    dX = torch.empty_like(X)
    with torch.cuda.device(dev):
        rc = lib.ssi_b200_lmhead_ce_bwd(X.data_ptr(), W.data_ptr(), labels.data_ptr(), lse.data_ptr(),
                                        grad_out.data_ptr(), n_valid_total.data_ptr(), N, V, D, int(ignore_index),
                                        dX.data_ptr(), dW.data_ptr(), 1 if accumulate_dW else 0, ws.data_ptr(),
                                        ws.numel(), _stream(dev))
    _lib.check(rc, "lmhead_ce_bwd")
    return dX, dW


def lmhead_ce_fwd_grad(X: torch.Tensor, W: torch.Tensor, labels: torch.Tensor, grad_hint: torch.Tensor,
                       n_valid_total: torch.Tensor | None = None, ignore_index: int = -100,
                       dW_acc: torch.Tensor | None = None, dw_ready_event: torch.cuda.Event | None = None,
                       accumulate_dW_acc: bool = False, dw_acc_init_event: torch.cuda.Event | None = None,
                       norm_scale: torch.Tensor | None = None, norm_eps: float = 1e-6,
                       pre_grad_event: torch.cuda.Event | None = None):
    """Single-pass mode: fused projection + CE forward that also forms the gradients (each logits tile is
    computed once). Returns ``(loss_out, lse, n_valid, dX, dW_acc)``; ``dX`` bf16 ``[N, D]`` and ``dW_acc`` fp32
    ``[V, D]`` are the gradients for an upstream gradient of ``grad_hint`` (fp32 device scalar);
    :func:`lmhead_ce_apply_grad` rescales them for the gradient autograd actually delivers.

    ``dW_acc``: optional existing fp32 ``[V, D]`` buffer to write into (data-parallel mode passes the view of its
    peer-visible exchange buffer); ``dw_ready_event``: recorded right after the last dW contraction;
    ``accumulate_dW_acc``: the caller initialised ``dW_acc`` and every dW contraction adds to it;
    ``dw_acc_init_event``: the stream waits for it right before the first dW contraction;
    ``pre_grad_event``: the stream waits for it right before the first softmax-grad pass, the first reader of
    ``grad_hint`` / ``n_valid_total`` (data parallel: both come from the counts exchange on another stream).

    ``norm_scale`` (bf16 ``[D]``): ``X`` holds the PRE-norm hidden states and the final RMSNorm
    (tt/torchtune/modules/rms_norm.py:38-43) is fused into the row gather / dX scatter of the path; ``dX`` is then
    the gradient of the pre-norm states and a sixth value ``dscale`` (fp32 ``[D]``) is returned.
    """
    dev = _require_cuda(X, W, labels, grad_hint)
    N, V, D = _check_xw(X, W)
    if labels.dtype != torch.int64 or labels.numel() != N or not labels.is_contiguous():
        raise ValueError("ssi_b200: labels must be a contiguous int64 tensor with one entry per row of X")
    if grad_hint.dtype != torch.float32 or grad_hint.numel() != 1:
        raise ValueError("ssi_b200: grad_hint must be a single fp32 value on the device")
    if n_valid_total is not None and (n_valid_total.dtype != torch.int64 or n_valid_total.numel() != 1
                                      or n_valid_total.device != dev):
        raise ValueError("ssi_b200: n_valid_total must be a single int64 value on the device")
    lib = _lib.load()
    ws = _workspace("single_pass", _lib.single_pass_workspace_bytes(N, V, D), dev)
    lse = torch.empty(N, dtype=torch.float32, device=dev)
    loss_out = torch.empty(2, dtype=torch.float32, device=dev)
    n_valid = torch.empty(1, dtype=torch.int64, device=dev)
    dX = torch.empty_like(X)
    if dW_acc is None:
        dW_acc = torch.empty(V, D, dtype=torch.float32, device=dev)
    elif (tuple(dW_acc.shape) != (V, D) or dW_acc.dtype != torch.float32 or not dW_acc.is_contiguous()
          or dW_acc.device != dev):
        raise ValueError("ssi_b200: dW_acc must be a contiguous fp32 [V, D] tensor on the same device")
    if accumulate_dW_acc and dW_acc is None:
        raise ValueError("ssi_b200: accumulate_dW_acc needs an initialised dW_acc buffer")
    evs = []
    for e in (dw_acc_init_event, dw_ready_event, pre_grad_event):
        h = None
        if e is not None:
            h = e.cuda_event
            if not h:
                raise ValueError("ssi_b200: event has no CUDA handle yet (record it once before first use)")
        evs.append(h)
    dscale = None
    if norm_scale is not None:
        if (norm_scale.dtype != torch.bfloat16 or norm_scale.numel() != D or not norm_scale.is_contiguous()
                or norm_scale.device != dev):
            raise ValueError("ssi_b200: norm_scale must be a contiguous bf16 [D] tensor on the same device")
        dscale = torch.empty(D, dtype=torch.float32, device=dev)
    import ctypes

    opts = _lib.FwdGradOpts(1 if accumulate_dW_acc else 0, evs[0], evs[1], evs[2],
                            norm_scale.data_ptr() if norm_scale is not None else None, float(norm_eps),
                            dscale.data_ptr() if dscale is not None else None)
    with torch.cuda.device(dev):
        rc = lib.ssi_b200_lmhead_ce_fwd_grad_ex(
            X.data_ptr(), W.data_ptr(), labels.data_ptr(), grad_hint.data_ptr(),
            n_valid_total.data_ptr() if n_valid_total is not None else None, N, V, D, int(ignore_index),
            lse.data_ptr(), loss_out.data_ptr(), n_valid.data_ptr(), dX.data_ptr(), dW_acc.data_ptr(),
            ws.data_ptr(), ws.numel(), _stream(dev), ctypes.byref(opts))
    _lib.check(rc, "lmhead_ce_fwd_grad")
    if norm_scale is not None:
        return loss_out, lse, n_valid, dX, dW_acc, dscale
    return loss_out, lse, n_valid, dX, dW_acc


def _fold_args(post_scale: torch.Tensor | None, fold_ws: torch.Tensor | None, dev: torch.device):
    """(post_scale ptr, fold_ws ptr, fold_ws bytes) of the ``_fold`` entry points (include/ssi_b200.h)."""
    if post_scale is not None and (post_scale.dtype != torch.float32 or post_scale.numel() != 1
                                   or post_scale.device != dev):
        raise ValueError("ssi_b200: post_scale must be a single fp32 value on the device")
    if fold_ws is not None and (fold_ws.dtype != torch.uint8 or fold_ws.device != dev or not fold_ws.is_contiguous()):
        raise ValueError("ssi_b200: fold_ws must be a contiguous uint8 tensor on the device (grad_fold_workspace)")
    return (post_scale.data_ptr() if post_scale is not None else None,
            fold_ws.data_ptr() if fold_ws is not None else None, fold_ws.numel() if fold_ws is not None else 0)


def grad_fold_workspace(max_rows: int, D: int, device: torch.device) -> torch.Tensor:
    """Scratch of the scale_grads / grad-norm fold for micro-batches of at most ``max_rows`` token rows."""
    b = ctypes.c_size_t(0)
    _lib.check(_lib.load().ssi_b200_grad_fold_workspace_bytes(int(max_rows), int(D), ctypes.byref(b)),
               "grad_fold_workspace_bytes")
    # not initialised here: the dense site of every fold zeroes it on ITS stream (an init kernel on the allocating
    # stream could land after a side stream's first partials)
    return torch.empty(b.value, dtype=torch.uint8, device=device)


def grad_fold_sqnorm(fold_ws: torch.Tensor) -> torch.Tensor:
    """fp32 ``[1]``: sum of squares of the whole tied gradient, from the partials the dense and the sparse site of
    this step left in ``fold_ws`` (no pass over the gradient)."""
    dev = _require_cuda(fold_ws)
    out = torch.empty(1, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.load().ssi_b200_grad_fold_sqnorm(fold_ws.data_ptr(), fold_ws.numel(), out.data_ptr(), _stream(dev))
    _lib.check(rc, "grad_fold_sqnorm")
    return out


def lmhead_ce_apply_grad(dX: torch.Tensor, dW_acc: torch.Tensor, grad_out: torch.Tensor, grad_hint: torch.Tensor,
                         dW: torch.Tensor | None = None, accumulate_dW: bool = False, skip_dW: bool = False,
                         post_scale: torch.Tensor | None = None, fold_ws: torch.Tensor | None = None):
    """Backward of single-pass mode: ``dX *= grad_out / grad_hint`` in place (a no-op on the device when the
    ratio is 1) and ``dW = bf16(post_scale * ((accumulate ? dW : 0) + ratio * dW_acc))``. Returns ``(dX, dW)``.
    ``skip_dW``: only rescale dX (data-parallel mode: the peer exchange rounds the SUMMED dW instead).
    ``post_scale`` / ``fold_ws``: the scale_grads / grad-norm fold (``ssi_b200_lmhead_ce_apply_grad_fold``)."""
    dev = _require_cuda(dX, dW_acc, grad_out, grad_hint)
    N, D = dX.shape
    V = dW_acc.size(0)
    if skip_dW:
        with torch.cuda.device(dev):
            rc = _lib.load().ssi_b200_lmhead_ce_apply_grad(dX.data_ptr(), dW_acc.data_ptr(), grad_out.data_ptr(),
                                                           grad_hint.data_ptr(), N, V, D, None, 0, _stream(dev))
        _lib.check(rc, "lmhead_ce_apply_grad")
        return dX, None
    if dW is None:
        if accumulate_dW:
            raise ValueError("ssi_b200: accumulate_dW needs an existing dW buffer")
        dW = torch.empty(V, D, dtype=torch.bfloat16, device=dev)
    elif tuple(dW.shape) != (V, D) or dW.dtype != torch.bfloat16 or not dW.is_contiguous() or dW.device != dev:
        raise ValueError("ssi_b200: dW must be a contiguous bf16 [V, D] tensor on the same device")
    lib = _lib.load()
    ps, fw, fwb = _fold_args(post_scale, fold_ws, dev)
    with torch.cuda.device(dev):
        rc = lib.ssi_b200_lmhead_ce_apply_grad_fold(dX.data_ptr(), dW_acc.data_ptr(), grad_out.data_ptr(),
                                                    grad_hint.data_ptr(), N, V, D, dW.data_ptr(),
                                                    1 if accumulate_dW else 0, ps, fw, fwb, _stream(dev))
    _lib.check(rc, "lmhead_ce_apply_grad")
    return dX, dW


def embedding_grad_scatter(dH: torch.Tensor, ids: torch.Tensor, dW: torch.Tensor) -> torch.Tensor:
    """``dW[ids[i]] += dH[i]`` in place (sorted-id warp-segmented reduction, deterministic). Returns ``dW``.

    Replaces the backward of the tied ``nn.Embedding`` (tt/torchtune/models/llama3_2/_component_builders.py:114,
    tt/torchtune/modules/transformer.py:628).
    """
    dev = _require_cuda(dH, ids, dW)
    if dH.dtype != torch.bfloat16 or dW.dtype != torch.bfloat16:
        raise TypeError("ssi_b200: dH and dW must be bfloat16")
    if dH.dim() != 2 or dW.dim() != 2 or dH.size(1) != dW.size(1):
        raise ValueError("ssi_b200: expected dH [N, D] and dW [V, D]")
    if ids.dtype != torch.int64 or ids.numel() != dH.size(0):
        raise ValueError("ssi_b200: ids must be int64 with one entry per row of dH")
    if not (dH.is_contiguous() and dW.is_contiguous() and ids.is_contiguous()):
        raise ValueError("ssi_b200: dH, ids and dW must be contiguous")
    N, D = dH.shape
    V = dW.size(0)
    lib = _lib.load()
    _, _, sc_b = _lib.workspace_bytes(N, V, D, "scatter")
    ws = _workspace("scatter", sc_b, dev)
    with torch.cuda.device(dev):
        rc = lib.ssi_b200_embedding_grad_scatter(dH.data_ptr(), ids.data_ptr(), N, V, D, dW.data_ptr(),
                                                 ws.data_ptr(), ws.numel(), _stream(dev))
    _lib.check(rc, "embedding_grad_scatter")
    return dW


def embedding_grad_scatter_f32(dH: torch.Tensor, ids: torch.Tensor, dW_acc: torch.Tensor) -> torch.Tensor:
    """``dW_acc[ids[i]] += dH[i]`` into an fp32 ``[V, D]`` buffer (data-parallel mode: the embedding gradient joins
    the fp32 LM-head sum before the peer exchange, so the tied gradient is rounded to bf16 once)."""
    dev = _require_cuda(dH, ids, dW_acc)
    if dH.dtype != torch.bfloat16 or dW_acc.dtype != torch.float32:
        raise TypeError("ssi_b200: dH must be bfloat16 and dW_acc float32")
    if dH.dim() != 2 or dW_acc.dim() != 2 or dH.size(1) != dW_acc.size(1):
        raise ValueError("ssi_b200: expected dH [N, D] and dW_acc [V, D]")
    if ids.dtype != torch.int64 or ids.numel() != dH.size(0):
        raise ValueError("ssi_b200: ids must be int64 with one entry per row of dH")
    if not (dH.is_contiguous() and dW_acc.is_contiguous() and ids.is_contiguous()):
        raise ValueError("ssi_b200: dH, ids and dW_acc must be contiguous")
    N, D = dH.shape
    V = dW_acc.size(0)
    lib = _lib.load()
    _, _, sc_b = _lib.workspace_bytes(N, V, D, "scatter")
    ws = _workspace("scatter", sc_b, dev)
    with torch.cuda.device(dev):
        rc = lib.ssi_b200_embedding_grad_scatter_f32(dH.data_ptr(), ids.data_ptr(), N, V, D, dW_acc.data_ptr(),
                                                     ws.data_ptr(), ws.numel(), _stream(dev))
    _lib.check(rc, "embedding_grad_scatter_f32")
    return dW_acc


def embedding_grad_sort(ids: torch.Tensor, V: int, D: int) -> torch.Tensor:
    """First half of the embedding-gradient scatter: radix-sort the token ids on the current stream (one launch; it
    needs only the ids, so the embedding FORWARD can start it).  Returns a fresh workspace holding the order; pass it
    to :func:`embedding_grad_scatter_sorted` / ``_sorted_f32`` / ``TiedGradExchange.embedding_exchange`` (any
    stream that is ordered after this call)."""
    dev = _require_cuda(ids)
    if ids.dtype != torch.int64 or not ids.is_contiguous():
        raise ValueError("ssi_b200: ids must be contiguous int64")
    N = ids.numel()
    lib = _lib.load()
    _, _, sc_b = _lib.workspace_bytes(N, V, D, "scatter")
    ws = torch.empty(max(sc_b, 256), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        rc = lib.ssi_b200_embedding_grad_sort(ids.data_ptr(), N, V, D, ws.data_ptr(), ws.numel(), _stream(dev))
    _lib.check(rc, "embedding_grad_sort")
    return ws


def embedding_grad_scatter_sorted_f32(dH: torch.Tensor, ws: torch.Tensor, dW_acc: torch.Tensor) -> torch.Tensor:
    """Second half: ``dW_acc[id] += sum of the dH rows with that id`` in the order :func:`embedding_grad_sort`
    left in ``ws`` (no shared memory: co-runs with a tensor-core contraction on another stream)."""
    dev = _require_cuda(dH, ws, dW_acc)
    if dH.dtype != torch.bfloat16 or dW_acc.dtype != torch.float32 or dH.dim() != 2 or dW_acc.dim() != 2:
        raise TypeError("ssi_b200: expected bf16 dH [N, D] and fp32 dW_acc [V, D]")
    if dH.size(1) != dW_acc.size(1) or not (dH.is_contiguous() and dW_acc.is_contiguous()):
        raise ValueError("ssi_b200: dH and dW_acc must be contiguous with the same row length")
    N, D = dH.shape
    V = dW_acc.size(0)
    with torch.cuda.device(dev):
        rc = _lib.load().ssi_b200_embedding_grad_scatter_sorted_f32(dH.data_ptr(), N, V, D, dW_acc.data_ptr(),
                                                                    ws.data_ptr(), ws.numel(), _stream(dev))
    _lib.check(rc, "embedding_grad_scatter_sorted_f32")
    return dW_acc


def embedding_grad_scatter_sorted(dH: torch.Tensor, ws: torch.Tensor, dW: torch.Tensor,
                                  post_scale: torch.Tensor | None = None,
                                  fold_ws: torch.Tensor | None = None) -> torch.Tensor:
    """``dW[id] += post_scale * sum of the dH rows with that id`` (bf16 ``[V, D]``, the tied weight's ``.grad``) in the
    order :func:`embedding_grad_sort` left in ``ws``; ``post_scale`` / ``fold_ws``: the scale_grads / grad-norm fold."""
    dev = _require_cuda(dH, ws, dW)
    if dH.dtype != torch.bfloat16 or dW.dtype != torch.bfloat16 or dH.dim() != 2 or dW.dim() != 2:
        raise TypeError("ssi_b200: expected bf16 dH [N, D] and bf16 dW [V, D]")
    if dH.size(1) != dW.size(1) or not (dH.is_contiguous() and dW.is_contiguous()):
        raise ValueError("ssi_b200: dH and dW must be contiguous with the same row length")
    N, D = dH.shape
    V = dW.size(0)
    ps, fw, fwb = _fold_args(post_scale, fold_ws, dev)
    with torch.cuda.device(dev):
        rc = _lib.load().ssi_b200_embedding_grad_scatter_sorted_fold(dH.data_ptr(), N, V, D, dW.data_ptr(),
                                                                     ws.data_ptr(), ws.numel(), ps, fw, fwb,
                                                                     _stream(dev))
    _lib.check(rc, "embedding_grad_scatter_sorted")
    return dW


def grad_add(dst: torch.Tensor, src: torch.Tensor) -> torch.Tensor:
    """``dst = bf16(float(dst) + float(src))`` in place for two contiguous bf16 tensors of the same shape."""
    dev = _require_cuda(dst, src)
    if dst.dtype != torch.bfloat16 or src.dtype != torch.bfloat16 or dst.shape != src.shape:
        raise ValueError("ssi_b200: dst and src must be bf16 tensors of the same shape")
    if not (dst.is_contiguous() and src.is_contiguous()) or dst.numel() % 8:
        raise ValueError("ssi_b200: dst and src must be contiguous with a multiple of 8 elements")
    with torch.cuda.device(dev):
        rc = _lib.load().ssi_b200_grad_add(dst.data_ptr(), src.data_ptr(), dst.numel(), _stream(dev))
    _lib.check(rc, "grad_add")
    return dst


def label_shift_count(labels: torch.Tensor, ignore_index: int = -100):
    """Shift labels left by one (ssi/loss.py:16) and count non-ignored tokens before/after the shift.

    Returns ``(shifted_labels [B, S], counts int64 [2])`` with counts[0] the unshifted count used by
    ssi/trainer.py:391 and counts[1] the shifted count the loss divides by.
    """
    dev = _require_cuda(labels)
    if labels.dtype != torch.int64 or labels.dim() != 2 or not labels.is_contiguous():
        raise ValueError("ssi_b200: labels must be a contiguous int64 [B, S] tensor")
    B, S = labels.shape
    out = torch.empty_like(labels)
    counts = torch.empty(2, dtype=torch.int64, device=dev)
    lib = _lib.load()
    with torch.cuda.device(dev):
        rc = lib.ssi_b200_label_shift_count(labels.data_ptr(), B, S, int(ignore_index), out.data_ptr(),
                                            counts.data_ptr(), _stream(dev))
    _lib.check(rc, "label_shift_count")
    return out, counts


def count_valid(labels: torch.Tensor, ignore_index: int = -100) -> torch.Tensor:
    """int64 ``[2]`` device tensor, both entries = number of labels != ignore_index (flat, already shifted labels)."""
    dev = _require_cuda(labels)
    if labels.dtype != torch.int64 or not labels.is_contiguous():
        raise ValueError("ssi_b200: labels must be a contiguous int64 tensor")
    counts = torch.empty(2, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.load().ssi_b200_count_valid(labels.data_ptr(), labels.numel(), int(ignore_index), counts.data_ptr(),
                                              _stream(dev))
    _lib.check(rc, "count_valid")
    return counts


def debug_gemm(A: torch.Tensor, B: torch.Tensor, a_mn_major: bool = False, b_mn_major: bool = False,
               C: torch.Tensor | None = None) -> torch.Tensor:
    """Test hook for the shared tcgen05 mainloop: fp32 ``C[M, N] (+)= op(A) @ op(B)^T`` (see ssi_b200.h)."""
    dev = _require_cuda(A, B)
    if A.dtype != torch.bfloat16 or B.dtype != torch.bfloat16 or not A.is_contiguous() or not B.is_contiguous():
        raise ValueError("ssi_b200: A and B must be contiguous bf16")
    M, K = (A.size(1), A.size(0)) if a_mn_major else (A.size(0), A.size(1))
    N, K2 = (B.size(1), B.size(0)) if b_mn_major else (B.size(0), B.size(1))
    if K != K2:
        raise ValueError("ssi_b200: contraction sizes differ")
    acc = C is not None
    if C is None:
        C = torch.empty(M, N, dtype=torch.float32, device=dev)
    lib = _lib.load()
    with torch.cuda.device(dev):
        rc = lib.ssi_b200_debug_gemm(A.data_ptr(), B.data_ptr(), C.data_ptr(), M, N, K, int(a_mn_major),
                                     int(b_mn_major), int(acc), _stream(dev))
    _lib.check(rc, "debug_gemm")
    return C


# ----------------------------------------------------------------------------------------------------------
# Rows SURVEY.md §8(f) marks "next": final RMSNorm, token-type histogram, scale_grads + grad norm, embedding gather
# ----------------------------------------------------------------------------------------------------------
def rmsnorm_fwd(x: torch.Tensor, scale: torch.Tensor, eps: float = 1e-6, want_rstd: bool = True):
    """Final RMSNorm (tt/torchtune/modules/rms_norm.py:38-43): ``x [N, D]`` bf16, ``scale [D]`` bf16 ->
    ``(y [N, D] bf16, rstd [N] fp32 or None)``, fp32 math, one rounding."""
    dev = _require_cuda(x, scale)
    if x.dtype != torch.bfloat16 or scale.dtype != torch.bfloat16:
        raise TypeError("ssi_b200: x and scale must be bfloat16")
    if x.dim() != 2 or scale.dim() != 1 or scale.numel() != x.size(1) or not (x.is_contiguous() and scale.is_contiguous()):
        raise ValueError("ssi_b200: expected contiguous x [N, D] and scale [D]")
    N, D = x.shape
    y = torch.empty_like(x)
    rstd = torch.empty(N, dtype=torch.float32, device=dev) if want_rstd else None
    with torch.cuda.device(dev):
        rc = _lib.load().ssi_b200_rmsnorm_fwd(x.data_ptr(), scale.data_ptr(), float(eps), N, D, y.data_ptr(),
                                              rstd.data_ptr() if rstd is not None else None, _stream(dev))
    _lib.check(rc, "rmsnorm_fwd")
    return y, rstd


def rmsnorm_bwd(dy: torch.Tensor, x: torch.Tensor, scale: torch.Tensor, rstd: torch.Tensor):
    """Backward of :func:`rmsnorm_fwd`: returns ``(dx [N, D] bf16, dscale [D] fp32)``."""
    dev = _require_cuda(dy, x, scale, rstd)
    if dy.dtype != torch.bfloat16 or x.dtype != torch.bfloat16 or scale.dtype != torch.bfloat16:
        raise TypeError("ssi_b200: dy, x and scale must be bfloat16")
    if dy.shape != x.shape or x.dim() != 2 or not (dy.is_contiguous() and x.is_contiguous()):
        raise ValueError("ssi_b200: dy and x must be contiguous [N, D] tensors of the same shape")
    if rstd.dtype != torch.float32 or rstd.numel() != x.size(0):
        raise ValueError("ssi_b200: rstd must be fp32 [N]")
    N, D = x.shape
    lib = _lib.load()
    import ctypes

    b = ctypes.c_size_t(0)
    _lib.check(lib.ssi_b200_rmsnorm_bwd_workspace_bytes(N, D, ctypes.byref(b)), "rmsnorm_bwd_workspace_bytes")
    ws = _workspace("rmsnorm_bwd", b.value, dev)
    dx = torch.empty_like(x)
    dscale = torch.empty(D, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        rc = lib.ssi_b200_rmsnorm_bwd(dy.data_ptr(), x.data_ptr(), scale.data_ptr(), rstd.data_ptr(), N, D,
                                      dx.data_ptr(), dscale.data_ptr(), ws.data_ptr(), ws.numel(), _stream(dev))
    _lib.check(rc, "rmsnorm_bwd")
    return dx, dscale


def label_shift_count_types(labels: torch.Tensor, tokens: torch.Tensor, ranges: dict, pad_idx: int,
                            ignore_index: int = -100):
    """Label shift + valid-token counts (as :func:`label_shift_count`) and the token-type histogram of
    ssi/train_utils.py:150-165 in one pass, everything left on the device (no host sync).

    ``ranges``: ``{name: (first_id, last_id)}`` inclusive, at most 15 entries.  Returns ``(shifted, counts int64
    [2], type_counts int64 [len(ranges) + 1])``; the last entry of ``type_counts`` is "total" = tokens != pad_idx,
    the others follow ``ranges``' order."""
    dev = _require_cuda(labels, tokens)
    if labels.dtype != torch.int64 or labels.dim() != 2 or not labels.is_contiguous():
        raise ValueError("ssi_b200: labels must be a contiguous int64 [B, S] tensor")
    if tokens.dtype != torch.int64 or tokens.shape != labels.shape or not tokens.is_contiguous():
        raise ValueError("ssi_b200: tokens must be a contiguous int64 tensor shaped like labels")
    import ctypes

    names = list(ranges)
    n = len(names)
    lo = (ctypes.c_int64 * max(n, 1))(*[int(ranges[k][0]) for k in names])
    hi = (ctypes.c_int64 * max(n, 1))(*[int(ranges[k][1]) for k in names])
    B, S = labels.shape
    out = torch.empty_like(labels)
    counts = torch.empty(2, dtype=torch.int64, device=dev)
    type_counts = torch.empty(n + 1, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.load().ssi_b200_label_shift_count_types(labels.data_ptr(), tokens.data_ptr(), B, S,
                                                          int(ignore_index), lo, hi, n, int(pad_idx), out.data_ptr(),
                                                          counts.data_ptr(), type_counts.data_ptr(), _stream(dev))
    _lib.check(rc, "label_shift_count_types")
    return out, counts, type_counts


def grad_scale_sqnorm(grad: torch.Tensor, scale: torch.Tensor) -> torch.Tensor:
    """``grad *= scale`` in place (tt/torchtune/training/_grad_scaler.py:25-31) and, in the same pass, the fp32
    sum of squares of the scaled bf16 values (this parameter's share of clip_grad_norm_'s total norm, ssi/
    trainer.py:405-408). ``scale``: fp32 device scalar. Returns the fp32 ``[1]`` sum of squares."""
    dev = _require_cuda(grad, scale)
    if grad.dtype != torch.bfloat16 or not grad.is_contiguous() or grad.numel() % 8:
        raise ValueError("ssi_b200: grad must be a contiguous bf16 tensor with a multiple of 8 elements")
    if scale.dtype != torch.float32 or scale.numel() != 1:
        raise ValueError("ssi_b200: scale must be a single fp32 value on the device")
    import ctypes

    lib = _lib.load()
    b = ctypes.c_size_t(0)
    _lib.check(lib.ssi_b200_grad_scale_sqnorm_workspace_bytes(ctypes.byref(b)), "grad_scale_sqnorm_workspace_bytes")
    ws = _workspace("sqnorm", b.value, dev)
    out = torch.empty(1, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        rc = lib.ssi_b200_grad_scale_sqnorm(grad.data_ptr(), grad.numel(), scale.data_ptr(), out.data_ptr(),
                                            ws.data_ptr(), ws.numel(), _stream(dev))
    _lib.check(rc, "grad_scale_sqnorm")
    return out


def embedding_gather(W: torch.Tensor, ids: torch.Tensor) -> torch.Tensor:
    """Forward of the tied ``nn.Embedding`` (tt/torchtune/modules/transformer.py:628): ``W[ids]`` as
    ``[*ids.shape, D]`` bf16, an exact copy of the rows."""
    dev = _require_cuda(W, ids)
    if W.dtype != torch.bfloat16 or W.dim() != 2 or not W.is_contiguous():
        raise ValueError("ssi_b200: W must be a contiguous bf16 [V, D] tensor")
    if ids.dtype != torch.int64:
        raise ValueError("ssi_b200: ids must be int64")
    flat = ids.reshape(-1)
    if not flat.is_contiguous():
        flat = flat.contiguous()
    V, D = W.shape
    out = torch.empty(flat.numel(), D, dtype=W.dtype, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.load().ssi_b200_embedding_gather(W.data_ptr(), flat.data_ptr(), flat.numel(), V, D, out.data_ptr(),
                                                   _stream(dev))
    _lib.check(rc, "embedding_gather")
    return out.reshape(*ids.shape, D)
