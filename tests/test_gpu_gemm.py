"""The tcgen05 GEMM alone (csrc/gemm_tc.cuh) against a float64 matmul of the same operand
values.  bf16 operands are exact in fp32, so only the accumulation order differs from the
reference: tolerance 1e-5 of sum|a||b| per element.  tf32 reads fp32 operands with a 10-bit
mantissa: tolerance 2^-10 of sum|a||b| (two operands, each off by < 2^-11 relative... the
bound used is 1.1e-3 of sum|a||b|)."""
import numpy as np
import pytest

from k8s_b200 import lib as L

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

SHAPES = [
    # M, N, K, block_n, splits
    (128, 64, 64, 64, 1), (128, 128, 128, 128, 1), (256, 128, 512, 128, 1),
    (200, 72, 800, 64, 1), (200, 72, 800, 128, 1),          # ragged M, N, K (K tail of half a block)
    (1, 10, 8, 64, 1), (8, 1024, 3136, 64, 1),               # fewer rows than one tile
    (1568, 64, 800, 64, 1),                                  # conv2 fwd at B = 8
    (800, 64, 1568, 64, 5), (800, 64, 25088, 64, 21),        # conv2 wgrad: split-K
    (3136, 1024, 128, 128, 1), (128, 3136, 1024, 64, 1),     # fc1 wgrad / dgrad at B = 128
    (130, 200, 72, 128, 3),
]


@pytest.mark.parametrize("math", ["bf16", "tf32"])
@pytest.mark.parametrize("M,N,K,block_n,splits", SHAPES)
def test_gemm_tn_matches_float64(lib, math, M, N, K, block_n, splits):
    g = torch.Generator(device="cpu").manual_seed(M * 7 + N * 3 + K)
    a = torch.randn(M, K, generator=g)
    b = torch.randn(N, K, generator=g)
    if math == "bf16":
        a, b = a.bfloat16(), b.bfloat16()
    da, db = a.cuda().contiguous(), b.cuda().contiguous()
    c = torch.full((M, N), float("nan"), device="cuda")
    partial = torch.empty(max(splits, 1) * M * N, device="cuda")
    bias = torch.randn(N, generator=g).cuda() if splits == 1 else None
    rc = lib.gemm_tn(L.MATH_MODES[math], da.data_ptr(), db.data_ptr(), c.data_ptr(), M, N, K, K, K, block_n,
                     splits, partial.data_ptr(), bias.data_ptr() if bias is not None else None, 0, 0)
    lib.check(rc)
    torch.cuda.synchronize()
    ref = a.double() @ b.double().T
    if bias is not None:
        ref = ref + bias.cpu().double()
    bound = a.double().abs() @ b.double().abs().T
    tol = (1e-5 if math == "bf16" else 1.1e-3) * bound + 1e-6
    err = (c.cpu().double() - ref).abs()
    assert not torch.isnan(c).any()
    assert bool((err <= tol).all()), "max err %.3e at bound %.3e" % (float(err.max()), float(bound.max()))


def test_gemm_tn_relu_and_argument_checks(lib):
    a = torch.randn(64, 64).bfloat16().cuda()
    b = torch.randn(32, 64).bfloat16().cuda()
    c = torch.empty(64, 32, device="cuda")
    lib.check(lib.gemm_tn(L.MATH_BF16, a.data_ptr(), b.data_ptr(), c.data_ptr(), 64, 32, 64, 64, 64, 64, 1, None,
                          None, 1, 0))
    torch.cuda.synchronize()
    ref = torch.relu(a.float() @ b.float().T)
    assert torch.allclose(c, ref, rtol=1e-4, atol=1e-4)
    assert lib.gemm_tn(L.MATH_FP32, a.data_ptr(), b.data_ptr(), c.data_ptr(), 64, 32, 64, 64, 64, 64, 1, None, None, 0, 0) == 2
    assert lib.gemm_tn(L.MATH_BF16, a.data_ptr(), b.data_ptr(), c.data_ptr(), 64, 32, 64, 64, 64, 96, 1, None, None, 0, 0) == 2
    assert lib.gemm_tn(L.MATH_BF16, a.data_ptr(), b.data_ptr(), c.data_ptr(), 64, 32, 60, 60, 60, 64, 1, None, None, 0, 0) == 2  # 120-B pitch
    a2 = torch.randn(64, 128).bfloat16().cuda()
    b2 = torch.randn(32, 128).bfloat16().cuda()
    assert lib.gemm_tn(L.MATH_BF16, a2.data_ptr(), b2.data_ptr(), c.data_ptr(), 64, 32, 128, 128, 128, 64, 2, None, None, 0, 0) == 2  # split-K w/o partial
