"""ssi_b200_rmsnorm_bwd_static: the RMSNorm backward for hosts that cannot hand the library any scratch memory.

Who needs it: a program that binds the C ABI directly and has given the whole GPU to its own arena at start-up
(cudaMalloc has nothing left for it, and there is no torch caching allocator to ask instead), or a training step
that keeps a second stream busy and must not call cudaMalloc / cudaFree mid-step, because either one stalls every
stream of the device.  For them the per-warp partial sums of dscale live in a fixed 18.5 MiB array inside the
library image (rows up to SSI_B200_RMSNORM_BWD_STATIC_MAX_D = 2048 columns wide), shared by all calls on a GPU.

What is checked, through ctypes on the raw entry point (the tensors below are the CALLER's inputs and outputs; the
entry point itself is handed no workspace):
  * two calls in a row -- the second narrower and shorter than the first, so it runs on top of what the first
    left in the shared array -- each give exactly the dx and dscale of ssi_b200_rmsnorm_bwd with a caller workspace
    (same kernels, same grid, fixed summation order: bit-identical, the bar the determinism asserts in
    test_gpu_next_rows.py use);
  * a repeat of the second call leaves the device's free memory where it was;
  * rows wider than the limit are refused on the host with a message that names the limit.

This file is named to be collected last: under `pytest -x` a failure here must not hide the parity tests.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_D = 2048  # SSI_B200_RMSNORM_BWD_STATIC_MAX_D in include/ssi_b200.h


def _inputs(N, D, seed):
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(N, D, generator=g) * 3).to(torch.bfloat16).cuda()
    scale = (1 + 0.2 * torch.randn(D, generator=g)).to(torch.bfloat16).cuda()
    dy = (torch.randn(N, D, generator=g) * 1e-2).to(torch.bfloat16).cuda()
    return x, scale, dy


def _static_bwd(lib, dy, x, scale, rstd, dx, dscale):
    """The raw C-ABI call: inputs, outputs, stream -- and nothing else."""
    N, D = x.shape
    stream = torch.cuda.current_stream().cuda_stream
    return lib.ssi_b200_rmsnorm_bwd_static(dy.data_ptr(), x.data_ptr(), scale.data_ptr(), rstd.data_ptr(), N, D,
                                           dx.data_ptr(), dscale.data_ptr(), ctypes.c_void_p(stream))


def test_static_scratch_rmsnorm_bwd_matches_the_workspace_entry_point_twice_in_a_row():
    from ssi_b200 import _lib, ops

    lib = _lib.load()
    torch.cuda.set_device(0)
    # first call wide and tall, second narrow and short: the second one reuses a dirty array
    cases = [(300, 2048, 11), (77, 520, 12)]
    prepared = []
    for N, D, seed in cases:
        x, scale, dy = _inputs(N, D, seed)
        _, rstd = ops.rmsnorm_fwd(x, scale, 1e-6)
        want_dx, want_dscale = ops.rmsnorm_bwd(dy, x, scale, rstd)  # existing entry point, caller workspace
        dx = torch.full_like(x, float("nan"))
        dscale = torch.full((D,), float("nan"), dtype=torch.float32, device="cuda")
        prepared.append((x, scale, dy, rstd, want_dx, want_dscale, dx, dscale))
    torch.cuda.synchronize()

    for x, scale, dy, rstd, _, _, dx, dscale in prepared:  # back to back on one stream, no sync in between
        _lib.check(_static_bwd(lib, dy, x, scale, rstd, dx, dscale), "rmsnorm_bwd_static")
    torch.cuda.synchronize()
    for (N, D, _), (_, _, _, _, want_dx, want_dscale, dx, dscale) in zip(cases, prepared):
        assert torch.equal(dx, want_dx), f"dx differs from ssi_b200_rmsnorm_bwd at N={N}, D={D}"
        assert torch.equal(dscale, want_dscale), f"dscale differs from ssi_b200_rmsnorm_bwd at N={N}, D={D}"

    # a repeat of the last call (kernels already loaded, address already looked up) takes no device memory
    x, scale, dy, rstd, want_dx, want_dscale, dx, dscale = prepared[-1]
    dx.fill_(float("nan"))
    dscale.fill_(float("nan"))
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    _lib.check(_static_bwd(lib, dy, x, scale, rstd, dx, dscale), "rmsnorm_bwd_static")
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after == free_before, "the static-scratch entry point changed the device's free memory"
    assert torch.equal(dx, want_dx) and torch.equal(dscale, want_dscale)


def test_static_scratch_rmsnorm_bwd_refuses_rows_wider_than_its_array():
    from ssi_b200 import _lib

    lib = _lib.load()
    torch.cuda.set_device(0)
    N, D = 4, MAX_D + 64
    x, scale, dy = _inputs(N, D, 13)
    rstd = torch.ones(N, dtype=torch.float32, device="cuda")
    dx = torch.zeros_like(x)
    dscale = torch.zeros(D, dtype=torch.float32, device="cuda")
    rc = _static_bwd(lib, dy, x, scale, rstd, dx, dscale)
    assert rc < 0
    msg = lib.ssi_b200_last_error().decode()
    assert str(MAX_D) in msg and "ssi_b200_rmsnorm_bwd" in msg, msg
    torch.cuda.synchronize()
    assert not dx.any() and not dscale.any()  # nothing was launched
