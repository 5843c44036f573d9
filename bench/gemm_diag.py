"""GEMM bring-up diagnostic: per layout/shape max error of the tcgen05 kernel vs torch fp32, with a
small dump of where mismatches sit (row/col pattern) to localise descriptor/swizzle mistakes."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_model_parallel_b200 import _native as nat

lib = nat.kernels()
sp = torch.cuda.current_stream().cuda_stream
for layout in (0, 1, 2):
    for (M, N, K) in [(128, 128, 64), (128, 64, 16), (256, 128, 128), (128, 128, 256), (384, 256, 192)]:
        torch.manual_seed(1)
        a = torch.randn(M, K, device="cuda").bfloat16()
        b = torch.randn(N, K, device="cuda").bfloat16()
        ref = a.float() @ b.float().t()
        if layout == 0: A, lda, B, ldb = a, K, b, K
        elif layout == 1: A, lda, B, ldb = a, K, b.t().contiguous(), N
        else: A, lda, B, ldb = a.t().contiguous(), M, b.t().contiguous(), N
        D = torch.zeros(M, N, device="cuda")
        rc = lib.dmp_gemm_bf16(layout, A.data_ptr(), lda, B.data_ptr(), ldb, D.data_ptr(), N, M, N, K, None, 1, 1, sp)
        torch.cuda.synchronize()
        err = (D - ref).abs()
        bad = err > 0.05 + 0.01 * ref.abs()
        print("layout", layout, (M, N, K), "rc", rc, "maxerr %.4g" % err.max().item(), "bad", int(bad.sum()),
              "badrows", int(bad.any(1).sum()), "badcols", int(bad.any(0).sum()))
        if bad.any():
            idx = bad.nonzero()[:6].tolist()
            print("   first bad:", [(i, j, round(D[i, j].item(), 3), round(ref[i, j].item(), 3)) for i, j in idx])
