"""The shared tcgen05/TMA mainloop against torch fp32 matmul, for every operand-major combination the product
kernels use: (K,K) logits = X W^T; (MN,MN) dW = G^T X; (K,MN) dX = G W.  Inputs are bf16-exact, so the only
difference from an fp32 reference is accumulation order: tolerance 1e-3 relative to the result scale."""
import pytest
import torch

pytestmark = pytest.mark.gpu

# (4864, 2048, 320): 19 m-pairs x 4 wide n-tiles = 76 items on 74 CTA pairs -> 74 wide items + 2 x 2 half items
# (tail splitting of the 256x512 schedule); (1000, 2048, 520): 16 items, all of them split
SHAPES = [(128, 256, 64), (128, 256, 256), (256, 512, 192), (200, 264, 136), (8, 8, 8), (1000, 2048, 520),
          (384, 1024, 2048), (4864, 2048, 320), (4904, 1800, 136)]


def _mk(shape, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return torch.randn(*shape, generator=g).to(torch.bfloat16).cuda()


@pytest.mark.parametrize("M,N,K", SHAPES)
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (True, True), (False, True)])
@pytest.mark.parametrize("wide", [2, 1])
def test_debug_gemm(M, N, K, a_mn, b_mn, wide):
    """wide = 2: the long-K contractions run 256x512 tiles per CTA pair (both TMEM accumulators on one tile);
    wide = 1: 256x256 tiles with a double-buffered accumulator.  (K,K) always uses 256x256."""
    from ssi_b200 import _lib, ops

    if wide == 1 and not a_mn and not b_mn:
        pytest.skip("(K,K) has no wide variant")
    _lib.load()
    _lib.set_option("wide_dw", wide)
    _lib.set_option("wide_dx", wide)
    _lib.set_option("shape_gate", 0)  # force the tile shape under test (the gate would pick 256x256 for small shapes)
    try:
        _run_debug_gemm(ops, M, N, K, a_mn, b_mn)
    finally:
        _lib.set_option("wide_dw", 2)
        _lib.set_option("wide_dx", 2)
        _lib.set_option("shape_gate", 1)


@pytest.mark.parametrize("M,N,K", [(1000, 2048, 520), (4864, 2048, 3200), (9728, 2048, 128)])
@pytest.mark.parametrize("a_mn,b_mn", [(True, True), (False, True)])
def test_debug_gemm_shape_gate(M, N, K, a_mn, b_mn):
    """Default configuration: both tile shapes are enqueued and the device-side predicate picks one (few items or a
    short K loop -> 256x256, else 256x512); exactly one of them must write the result."""
    from ssi_b200 import ops

    _run_debug_gemm(ops, M, N, K, a_mn, b_mn)


def _run_debug_gemm(ops, M, N, K, a_mn, b_mn):

    A = _mk((K, M) if a_mn else (M, K), 1)
    B = _mk((K, N) if b_mn else (N, K), 2)
    Af = A.float().t() if a_mn else A.float()
    Bf = B.float().t() if b_mn else B.float()
    ref = Af @ Bf.t()
    C = ops.debug_gemm(A, B, a_mn, b_mn)
    torch.cuda.synchronize()
    scale = ref.abs().max().item() + 1e-6
    err = (C - ref).abs().max().item()
    assert err <= 1e-3 * scale, f"max err {err} vs scale {scale}"
    # accumulate: TMA reduce-add on top of the previous result
    C2 = ops.debug_gemm(A, B, a_mn, b_mn, C=C.clone())
    torch.cuda.synchronize()
    err2 = (C2 - 2 * ref).abs().max().item()
    assert err2 <= 2e-3 * scale
