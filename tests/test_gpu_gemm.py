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
    (128, 1024, 3136, 128, 17), (128, 3136, 1024, 128, 5),   # fc1 fwd / dgrad in 128-wide tiles, deep split-K, ragged last N tile
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


def _x3_case(lib, majors, M, N, K, block_n, splits, chunk_kb, same_sign):
    g = torch.Generator(device="cpu").manual_seed(M * 11 + N * 5 + K)
    a = torch.randn(M, K, generator=g)
    b = torch.randn(N, K, generator=g)
    if same_sign:
        a, b = a.abs(), b.abs()
    a_mn, b_mn = majors[0] == "M", majors[1] == "M"
    if (a_mn and (M * 4) % 16) or (b_mn and (N * 4) % 16):
        pytest.skip("row pitch of the MN-major operand is not a multiple of 16 bytes")
    da = (a.t() if a_mn else a).contiguous().cuda()
    db = (b.t() if b_mn else b).contiguous().cuda()
    da_lo, db_lo = torch.empty_like(da), torch.empty_like(db)
    lib.check(lib.tf32_split_lo(da.data_ptr(), da_lo.data_ptr(), da.numel(), 0))
    lib.check(lib.tf32_split_lo(db.data_ptr(), db_lo.data_ptr(), db.numel(), 0))
    # the split itself: hi = the top 19 bits, lo = the remainder rounded to tf32, |x - hi - lo| <= 2^-21 |x|
    hi = (da.view(torch.int32) & -8192).view(torch.float32)
    assert bool(((da - hi - da_lo).abs() <= da.abs() * 2.0 ** -21).all())
    assert bool(((da_lo.view(torch.int32) & 8191) == 0).all())
    lda, ldb = (M if a_mn else K), (N if b_mn else K)
    c = torch.full((M, N), float("nan"), device="cuda")
    partial = torch.empty(max(splits, 1) * M * N, device="cuda")
    bias = torch.randn(N, generator=g).cuda()
    lib.check(lib.gemm_tc_x3(da.data_ptr(), da_lo.data_ptr(), int(a_mn), db.data_ptr(), db_lo.data_ptr(), int(b_mn),
                             c.data_ptr(), M, N, K, lda, ldb, block_n, splits, chunk_kb, partial.data_ptr(),
                             bias.data_ptr(), 0, 0))
    torch.cuda.synchronize()
    assert not torch.isnan(c).any()
    ref = a.double() @ b.double().T + bias.cpu().double()
    err = (c.cpu().double() - ref).abs()
    err32 = ((a @ b.T + bias.cpu()).double() - ref).abs()          # what plain fp32 on the CPU makes of it
    return err, err32, ref, a.double().abs() @ b.double().abs().T


@pytest.mark.parametrize("same_sign", [False, True])
@pytest.mark.parametrize("majors", ["KK", "KM", "MM", "MK"])
@pytest.mark.parametrize("M,N,K,block_n,splits", SHAPES)
def test_gemm_tc_x3_is_fp32_equivalent(lib, majors, M, N, K, block_n, splits, same_sign):
    """The 3xTF32 split product with chunked accumulation (gemm_x3.cuh) against float64: the error
    must be of the order of fp32 summation noise, not of the tf32 operand error (1e-3) nor of the
    tensor core's truncating accumulator (6e-5 of the result at K = 3136).  Bound: 8e-7 of
    sum|a||b| per element (the dropped lo*lo term is < 2^-20 and the truncation inside a chunk a few
    ulp; same-sign inputs make both add up coherently and make sum|a||b| the result itself) -- what
    a sequential fp32 sum of a few thousand terms achieves."""
    err, err32, ref, bound = _x3_case(lib, majors, M, N, K, block_n, splits, 0, same_sign)
    assert bool((err <= 8e-7 * bound + 1e-6).all()), "max err/bound %.3e" % float((err / (bound + 1e-30)).max())


def test_gemm_tc_x3_chunking_is_what_removes_the_accumulator_bias(lib):
    """K = 3136 with same-sign operands: one TMEM accumulator for the whole K loop (chunk_kb = 98)
    shows the tensor core's truncating adds as a bias that shrinks the result; chunks of 2 K
    blocks summed in registers do not."""
    big = _x3_case(lib, "KK", 128, 64, 3136, 64, 1, 98, True)
    small = _x3_case(lib, "KK", 128, 64, 3136, 64, 1, 2, True)
    rel_big = float((big[0] / big[2].abs()).mean())
    rel_small = float((small[0] / small[2].abs()).mean())
    print("mean relative error: one accumulator %.3e, chunks of 2 K blocks %.3e, fp32 CPU matmul %.3e"
          % (rel_big, rel_small, float((small[1] / small[2].abs()).mean())))
    assert rel_small <= 5e-7
    assert rel_big >= 20 * rel_small


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
