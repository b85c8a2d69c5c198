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


@pytest.mark.parametrize("majors", ["KK", "KM", "MM", "MK"])     # storage of A and B: K-major or MN-major
@pytest.mark.parametrize("math", ["bf16", "tf32"])
@pytest.mark.parametrize("M,N,K,block_n,splits", SHAPES)
def test_gemm_tc_matches_float64(lib, math, majors, M, N, K, block_n, splits):
    g = torch.Generator(device="cpu").manual_seed(M * 7 + N * 3 + K)
    a = torch.randn(M, K, generator=g)
    b = torch.randn(N, K, generator=g)
    if math == "bf16":
        a, b = a.bfloat16(), b.bfloat16()
    a_mn, b_mn = majors[0] == "M", majors[1] == "M"
    esize = 2 if math == "bf16" else 4
    if (a_mn and (M * esize) % 16) or (b_mn and (N * esize) % 16):
        pytest.skip("row pitch of the MN-major operand is not a multiple of 16 bytes")
    da = (a.t() if a_mn else a).contiguous().cuda()               # MN-major = stored [K, rows]
    db = (b.t() if b_mn else b).contiguous().cuda()
    lda, ldb = (M if a_mn else K), (N if b_mn else K)
    c = torch.full((M, N), float("nan"), device="cuda")
    partial = torch.empty(max(splits, 1) * M * N, device="cuda")
    bias = torch.randn(N, generator=g).cuda()                    # with split-K the reduce pass adds it
    rc = lib.gemm_tc(L.MATH_MODES[math], da.data_ptr(), int(a_mn), db.data_ptr(), int(b_mn), c.data_ptr(), M, N, K,
                     lda, ldb, block_n, splits, partial.data_ptr(), bias.data_ptr() if bias is not None else None, 0, 0)
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


def test_gemm_tc_relu_and_argument_checks(lib):
    a = torch.randn(64, 64).bfloat16().cuda()
    b = torch.randn(32, 64).bfloat16().cuda()
    c = torch.empty(64, 32, device="cuda")

    def call(math, K, ld, block_n, splits, aa=a, bb=b):
        return lib.gemm_tc(math, aa.data_ptr(), 0, bb.data_ptr(), 0, c.data_ptr(), 64, 32, K, ld, ld, block_n, splits,
                           None, None, 1, 0)

    lib.check(call(L.MATH_BF16, 64, 64, 64, 1))
    torch.cuda.synchronize()
    ref = torch.relu(a.float() @ b.float().T)
    assert torch.allclose(c, ref, rtol=1e-4, atol=1e-4)
    assert call(L.MATH_FP32, 64, 64, 64, 1) == 2
    assert call(L.MATH_BF16, 64, 64, 96, 1) == 2
    assert call(L.MATH_BF16, 60, 60, 64, 1) == 2                     # 120-byte pitch
    a2 = torch.randn(64, 128).bfloat16().cuda()
    b2 = torch.randn(32, 128).bfloat16().cuda()
    assert call(L.MATH_BF16, 128, 128, 64, 2, a2, b2) == 2           # split-K without a partial buffer
