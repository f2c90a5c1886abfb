"""Peer-memory exchange of the tied LM-head / embedding gradient for one process per GPU on one
NVSwitch node.

Each rank owns one exchange buffer (ctrl flags + token counts | fp32 ``dW_acc [V, D]`` | bf16 stage slice | slot
table | compact embedding rows) allocated by the library and mapped into every peer through CUDA IPC
(``ssi_b200_dp_alloc`` / ``ssi_b200_dp_open``); the handles travel through ``torch.distributed`` (any backend: NCCL on
the GPUs, gloo in tests).  Three exchanges per step, in the order a training step produces their inputs
(``csrc/dp_exchange.cuh``):

* ``counts()`` — the two token counts summed over ranks through 8-byte peer stores (no collective call);
* ``exchange()`` — the LM-head part: rank r sums vocabulary slice r of all ranks' fp32 ``dW_acc`` over NVLink P2P
  loads in rank order, rounds ONCE to bf16, and every rank then pulls the finished slices; it co-resides with the
  tensor-core CTAs of the last dX contraction (no shared memory), so it hides behind math instead of following it;
* ``embedding_exchange()`` — the embedding part, which only exists after the decoder backward: every rank publishes
  its dH0 rows summed per distinct id (compact list) and adds all ranks' rows, in rank order, to its gradient rows.

The result is bit-identical on every rank and run to run.

There is no reference counterpart (the reference trains on one GPU); semantics are those of torchtune's
distributed recipe (tt/recipes/full_finetune_distributed.py:844-851,867): gradients of the global mean.
"""
from __future__ import annotations

import ctypes
import os

import torch
import torch.distributed as dist

from . import _lib, ops

MAX_WORLD = 8


class _DevicePtr:
    """Minimal ``__cuda_array_interface__`` carrier so torch can view library-owned device memory."""

    def __init__(self, ptr: int, shape: tuple[int, ...], typestr: str):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (ptr, False), "version": 3,
                                         "strides": None}


class TiedGradExchange:
    """Owns this rank's exchange buffer and the mappings of its peers' buffers.

    Attributes:
        dw_acc: fp32 ``[V, D]`` tensor view of the peer-visible accumulator; hand it to the single-pass forward.
        max_rows: most token rows one ``embedding_exchange`` call may carry (sizes the compact-row area).
    """

    def __init__(self, V: int, D: int, group=None, device: torch.device | None = None, max_rows: int = 0,
                 nvls: bool | None = None):
        """``nvls``: allocate the exchange buffers through ``torch.distributed._symmetric_memory`` (plumbing only:
        allocation, handle exchange, multicast binding) and let the dense reduce use the in-switch reduction
        (``multimem.ld_reduce`` on the multicast mapping).  None = the environment variable ``SSI_B200_NVLS`` (default
        off); True raises if the box has no multicast support; the CUDA-IPC path is the fallback otherwise."""
        if not dist.is_initialized():
            raise RuntimeError("ssi_b200.peer: torch.distributed is not initialised")
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        if self.world > MAX_WORLD:
            raise RuntimeError(f"ssi_b200.peer: world size {self.world} > {MAX_WORLD} (one NVSwitch node)")
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.V, self.D = int(V), int(D)
        self.max_rows = int(max_rows)
        self.epoch_dense = self.epoch_sparse = self.epoch_counts = 0
        lib = _lib.load()
        # options that change which flags a rank waits on must be the same everywhere: found here, by all ranks
        # together, instead of as a device-side time-out in the first exchange
        mine = {k: _lib.get_option(k) for k in ("sparse_two_phase",)}
        every: list = [None] * self.world
        dist.all_gather_object(every, mine, group=group)
        if any(e != every[0] for e in every):
            raise RuntimeError(f"ssi_b200.peer: the ranks disagree on an exchange option (set_option / SSI_B200_*): "
                               f"{every}")
        nbytes, off = ctypes.c_size_t(0), ctypes.c_size_t(0)
        _lib.check(lib.ssi_b200_dp_buffer_bytes(self.V, self.D, self.world, self.max_rows, ctypes.byref(nbytes),
                                                ctypes.byref(off)), "dp_buffer_bytes")
        self.nbytes, self._acc_off = nbytes.value, off.value
        own = ctypes.c_void_p(0)
        handle = (ctypes.c_ubyte * 64)()
        self._own = None
        self._peer_ptrs: list[int] = []
        self._opened: list[int] = []
        self._symm = self._symm_tensor = None
        self.multicast_ptr = 0
        want_nvls = (os.environ.get("SSI_B200_NVLS") == "1") if nvls is None else bool(nvls)
        if want_nvls and self._setup_symmetric():
            return
        if nvls:
            raise RuntimeError("ssi_b200.peer: nvls=True, but symmetric memory with multicast is not available here")
        # Set-up is collective: a failure on ANY rank (no IPC in this container, no P2P between two GPUs, out of
        # memory) is agreed on by all ranks before anyone raises, so that callers can fall back together.
        err = None
        try:
            with torch.cuda.device(self.device):
                _lib.check(lib.ssi_b200_dp_alloc(self.nbytes, ctypes.byref(own),
                                                 ctypes.cast(handle, ctypes.c_void_p)), "dp_alloc")
            self._own = own.value
        except RuntimeError as e:
            err = str(e)
        gathered: list = [None] * self.world
        dist.all_gather_object(gathered, (bytes(handle), err), group=group)
        if all(e is None for _, e in gathered):
            try:
                with torch.cuda.device(self.device):
                    for r in range(self.world):
                        if r == self.rank:
                            self._peer_ptrs.append(self._own)
                            continue
                        buf = (ctypes.c_ubyte * 64).from_buffer_copy(gathered[r][0])
                        p = ctypes.c_void_p(0)
                        _lib.check(lib.ssi_b200_dp_open(ctypes.cast(buf, ctypes.c_void_p), ctypes.byref(p)),
                                   "dp_open")
                        self._peer_ptrs.append(p.value)
                        self._opened.append(p.value)
            except RuntimeError as e:
                err = str(e)
        errs: list = [None] * self.world
        dist.all_gather_object(errs, err, group=group)
        if any(e is not None for _, e in gathered) or any(e is not None for e in errs):
            if self._opened or self._own is not None:
                with torch.cuda.device(self.device):
                    for p in self._opened:
                        lib.ssi_b200_dp_close(p)
                    if self._own is not None:
                        lib.ssi_b200_dp_free(self._own)
            self._opened, self._own = [], None
            why = next(e for e in [g[1] for g in gathered] + errs if e is not None)
            raise RuntimeError(f"ssi_b200.peer: peer-memory set-up failed on at least one rank: {why}")
        self._finish_setup()

    def _finish_setup(self) -> None:
        self._ptr_array = (ctypes.c_void_p * self.world)(*self._peer_ptrs)
        self.dw_acc = torch.as_tensor(_DevicePtr(self._own + self._acc_off, (self.V, self.D), "<f4"),
                                      device=self.device)
        # pinned host word a timed-out wait sets to 1: polled by check() without any device synchronisation
        self._host_flag = torch.zeros(1, dtype=torch.int32).pin_memory()
        # the wait limit (option dp_timeout_s / SSI_B200_DP_TIMEOUT_S) goes into the ctrl block: dp_alloc stored it
        # for the IPC buffers, torch's symmetric allocation gets it here
        with torch.cuda.device(self.device):
            _lib.check(_lib.load().ssi_b200_dp_reset(self._own), "dp_reset")
        # every rank must have mapped (and zero-initialised) its buffers before the first exchange touches them
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)

    def _setup_symmetric(self) -> bool:
        """Exchange buffers from torch's symmetric-memory allocator (cuMem + multicast binding); False -- agreed on
        by all ranks -- when any rank cannot (old driver, no NVSwitch multicast, world size 1)."""
        err, t, hdl = None, None, None
        try:
            import torch.distributed._symmetric_memory as symm

            t = symm.empty(self.nbytes, dtype=torch.uint8, device=self.device)
            t.zero_()
            hdl = symm.rendezvous(t, self.group if self.group is not None else dist.group.WORLD)
            if int(hdl.multicast_ptr) == 0:
                err = "no multicast mapping (multicast_ptr == 0)"
        except Exception as e:  # noqa: BLE001 - any failure means "use the IPC path"
            err = f"{type(e).__name__}: {e}"
        errs: list = [None] * self.world
        dist.all_gather_object(errs, err, group=self.group)
        if any(e is not None for e in errs):
            return False
        self._symm, self._symm_tensor = hdl, t
        self._peer_ptrs = [int(p) for p in hdl.buffer_ptrs]
        self._own = self._peer_ptrs[self.rank]
        self.multicast_ptr = int(hdl.multicast_ptr)
        self._finish_setup()
        return True

    # ------------------------------------------------------------------------------------------------ exchanges
    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    def check(self) -> None:
        """Raise if an earlier exchange timed out waiting for a peer (reads a pinned host word: no device sync)."""
        if self._host_flag is not None and int(self._host_flag[0]) != 0:
            raise RuntimeError("ssi_b200.peer: an exchange timed out waiting for a peer (on this rank or another); "
                               "the gradients since were replaced by NaN on every rank -- reset() on all ranks "
                               "recovers; option dp_timeout_s sets the limit")

    def reset(self) -> None:
        """Collective recovery after a time-out (every rank of the group must call it, e.g. from the ``except`` of a
        failed :meth:`check`): waits until no rank has an exchange in flight, clears every rank's time-out mark,
        re-reads the ``dp_timeout_s`` option and realigns the step counters, so that the next step exchanges
        normally.  The gradients of the step that timed out are lost (NaN) -- redo or skip that step."""
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)  # nobody is spinning on, or still poisoning, anybody's ctrl block
        with torch.cuda.device(self.device):
            _lib.check(_lib.load().ssi_b200_dp_reset(self._own), "dp_reset")
        self._host_flag.zero_()
        mine = (self.epoch_counts, self.epoch_dense, self.epoch_sparse)
        every: list = [None] * self.world
        dist.all_gather_object(every, mine, group=self.group)  # doubles as the barrier after the clear
        self.epoch_counts, self.epoch_dense, self.epoch_sparse = (max(e[i] for e in every) for i in range(3))

    def counts(self, local_counts: torch.Tensor):
        """Enqueue on the current stream: returns ``(global_counts int64 [2], hint fp32 [1])`` = the int64 ``[2]``
        ``local_counts`` summed over ranks, and ``float(global_counts[0])``."""
        if local_counts.dtype != torch.int64 or local_counts.numel() != 2 or local_counts.device != self.device:
            raise ValueError("ssi_b200.peer: local_counts must be an int64 [2] tensor on the exchange device")
        self.check()
        self.epoch_counts += 1
        out = torch.empty(2, dtype=torch.int64, device=self.device)
        hint = torch.empty(1, dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            rc = _lib.load().ssi_b200_dp_counts(self._ptr_array, self.rank, self.world,
                                                self.epoch_counts & 0xFFFFFFFF, local_counts.data_ptr(), out.data_ptr(),
                                                hint.data_ptr(), self._host_flag.data_ptr(), self._stream())
        _lib.check(rc, "dp_counts")
        return out, hint

    def exchange(self, dW: torch.Tensor, grad_out: torch.Tensor, grad_hint: torch.Tensor,
                 accumulate: bool = False, post_scale: torch.Tensor | None = None,
                 fold_ws: torch.Tensor | None = None) -> torch.Tensor:
        """Enqueue on the current stream: ``dW = bf16(post_scale * ((accumulate ? dW : 0) + grad_out / grad_hint *
        sum_ranks dw_acc))`` on every rank.  Every rank must call it the same number of times.  ``post_scale`` /
        ``fold_ws``: the scale_grads / grad-norm fold (``ssi_b200_dp_exchange_fold``)."""
        self._check_dw(dW)
        self.epoch_dense += 1
        ps, fw, fwb = ops._fold_args(post_scale, fold_ws, self.device)
        with torch.cuda.device(self.device):
            rc = _lib.load().ssi_b200_dp_exchange_nvls(self._ptr_array, self.multicast_ptr or None, self.rank,
                                                       self.world, self.V, self.D, self.max_rows,
                                                       self.epoch_dense & 0xFFFFFFFF, grad_out.data_ptr(),
                                                       grad_hint.data_ptr(), dW.data_ptr(), 1 if accumulate else 0,
                                                       self._host_flag.data_ptr(), ps, fw, fwb, self._stream())
        _lib.check(rc, "dp_exchange")
        return dW

    def embedding_exchange(self, dH: torch.Tensor, sort_ws: torch.Tensor, dW: torch.Tensor,
                           post_scale: torch.Tensor | None = None,
                           fold_ws: torch.Tensor | None = None) -> torch.Tensor:
        """Enqueue on the current stream: ``dW[id] += sum over ranks (rank order) of that rank's dH rows with this
        id`` on every rank; ``sort_ws`` from :func:`ssi_b200.ops.embedding_grad_sort` of this rank's ids.  Must be
        stream-ordered after this step's :meth:`exchange`."""
        self._check_dw(dW)
        if dH.dtype != torch.bfloat16 or dH.dim() != 2 or dH.size(1) != self.D or not dH.is_contiguous():
            raise ValueError("ssi_b200.peer: dH must be a contiguous bf16 [N, D] tensor")
        if dH.size(0) > self.max_rows:
            raise RuntimeError(f"ssi_b200.peer: {dH.size(0)} token rows exceed max_rows={self.max_rows} of the "
                               "exchange buffer; pass a larger max_rows (install(..., max_tokens_per_rank=...))")
        self.epoch_sparse += 1
        ps, fw, fwb = ops._fold_args(post_scale, fold_ws, self.device)
        with torch.cuda.device(self.device):
            rc = _lib.load().ssi_b200_dp_embedding_exchange_fold(
                dH.data_ptr(), dH.size(0), self.V, self.D, sort_ws.data_ptr(), sort_ws.numel(), self._ptr_array,
                self.rank, self.world, self.max_rows, self.epoch_sparse & 0xFFFFFFFF, dW.data_ptr(),
                self._host_flag.data_ptr(), ps, fw, fwb, self._stream())
        _lib.check(rc, "dp_embedding_exchange")
        return dW

    def _check_dw(self, dW: torch.Tensor) -> None:
        if tuple(dW.shape) != (self.V, self.D) or dW.dtype != torch.bfloat16 or not dW.is_contiguous():
            raise ValueError("ssi_b200.peer: dW must be a contiguous bf16 [V, D] tensor")
        if dW.device != self.device:
            raise ValueError("ssi_b200.peer: dW is on another device than the exchange buffer")

    def timed_out(self) -> bool:
        """True if a wait inside an exchange gave up because a peer never arrived (synchronises)."""
        flag = ctypes.c_int(0)
        with torch.cuda.device(self.device):
            _lib.check(_lib.load().ssi_b200_dp_status(self._own, ctypes.byref(flag)), "dp_status")
        return bool(flag.value)

    def close(self) -> None:
        if self._own is None:
            return
        torch.cuda.synchronize(self.device)
        if self.timed_out():
            raise RuntimeError("ssi_b200.peer: an exchange timed out waiting for a peer; gradients were invalid")
        if dist.is_initialized():
            dist.barrier(group=self.group)  # nobody may unmap while a peer can still read
        lib = _lib.load()
        self.dw_acc = None
        if self._symm is not None:  # torch owns the symmetric allocation: drop the references
            self._symm = self._symm_tensor = None
            self.multicast_ptr = 0
        else:
            with torch.cuda.device(self.device):
                for p in self._opened:
                    lib.ssi_b200_dp_close(p)
                lib.ssi_b200_dp_free(self._own)
        self._opened, self._own = [], None
