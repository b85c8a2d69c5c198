"""Which fp32 -> tf32 operand conversion does tcgen05 kind::tf32 apply?  (run on the GPU box)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from k8s_b200 import lib as L
lib = L.load()
for name, val in [("1+0.75ulp", 1 + 2**-11 + 2**-12), ("1+0.5ulp", 1 + 2**-11), ("1+0.25ulp", 1 + 2**-12),
                  ("1+1.5ulp", 1 + 2**-10 + 2**-11), ("-(1+0.75ulp)", -(1 + 2**-11 + 2**-12))]:
    a = torch.full((128, 8), val, dtype=torch.float32, device="cuda")
    b = torch.ones((64, 8), dtype=torch.float32, device="cuda")
    c = torch.zeros((128, 64), device="cuda")
    lib.check(lib.gemm_tn(L.MATH_TF32, a.data_ptr(), b.data_ptr(), c.data_ptr(), 128, 64, 8, 8, 8, 64, 1, None, None, 0, 0))
    torch.cuda.synchronize()
    print("%-14s a=%.10f  C/8=%.10f  (trunc -> %.10f, rn -> %.10f)" % (
        name, val, c[0, 0].item() / 8,
        (torch.tensor([val]).view(torch.int32) & -8192).view(torch.float32).item(),
        ((torch.tensor([val]).view(torch.int32) + 4096) & -8192).view(torch.float32).item()))
