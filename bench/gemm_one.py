"""Runs one GEMM shape a few times (ncu target).  usage: gemm_one.py layout M N K [iters]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_model_parallel_b200 import _native as nat
lib = nat.kernels()
layout, M, N, K = [int(v) for v in sys.argv[1:5]]
iters = int(sys.argv[5]) if len(sys.argv) > 5 else 5
sp = torch.cuda.current_stream().cuda_stream
a = torch.randn(M, K, device="cuda").bfloat16(); b = torch.randn(N, K, device="cuda").bfloat16()
if layout == 0: A, lda, B, ldb = a, K, b, K
elif layout == 1: A, lda, B, ldb = a, K, b.t().contiguous(), N
else: A, lda, B, ldb = a.t().contiguous(), M, b.t().contiguous(), N
D = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16 if layout != 2 else torch.float32)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(iters):
    flush.zero_()
    nat.check(lib.dmp_gemm_bf16(layout, A.data_ptr(), lda, B.data_ptr(), ldb, D.data_ptr(), N, M, N, K, None, 0 if layout != 2 else 2, 0 if layout == 2 else 1, sp))
torch.cuda.synchronize()
print("ok", float(D.float().abs().mean()))
