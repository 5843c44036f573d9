"""Tiny-shape pass over the CUDA-core and tcgen05 kernels, meant to be run under compute-sanitizer:
    compute-sanitizer --tool memcheck  python bench/sanitizer_smoke.py
    compute-sanitizer --tool racecheck python bench/sanitizer_smoke.py
Every section prints as soon as it has synchronised, so a partial log still says how far it got.
(compute-sanitizer is closed on the GPU pool this round was measured on -- the runner refuses it -- so the kernels' memory
safety rests on the numerics tests at odd shapes (tails in M, N, K, H, W) and the host-side ASan/UBSan pass,
tools/sanitize_runtime.sh; the script also runs plain, as a tiny-shape smoke of every kernel family.)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_model_parallel_b200 import _native as nat  # noqa: E402

lib = nat.kernels()
dev = "cuda"
sp = torch.cuda.current_stream().cuda_stream


def done(name):
    torch.cuda.synchronize()
    print("ok", name, flush=True)


# BatchNorm: fused forward / backward, split path
M, C = 256, 32
x = torch.randn(M, C, device=dev).bfloat16(); y = torch.empty_like(x); dx = torch.empty_like(x)
g, b = torch.ones(C, device=dev), torch.zeros(C, device=dev)
rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
nbt = torch.zeros((), dtype=torch.int64, device=dev)
mean, invstd = torch.empty(C, device=dev), torch.empty(C, device=dev)
dg, db = torch.empty(C, device=dev), torch.empty(C, device=dev)
ws = torch.zeros(int(lib.dmp_bn_fused_ws_bytes(C)) // 8, dtype=torch.float64, device=dev)
ws2 = torch.zeros_like(ws)
nat.check(lib.dmp_bn_fwd_fused(x.data_ptr(), 1, None, y.data_ptr(), M, C, g.data_ptr(), b.data_ptr(), rm.data_ptr(), rv.data_ptr(), nbt.data_ptr(),
                               0.1, 1e-5, mean.data_ptr(), invstd.data_ptr(), ws.data_ptr(), 1, sp), "bn fwd")
nat.check(lib.dmp_bn_bwd_fused(y.data_ptr(), None, x.data_ptr(), 1, mean.data_ptr(), invstd.data_ptr(), g.data_ptr(), b.data_ptr(), dx.data_ptr(),
                               None, dg.data_ptr(), db.data_ptr(), 0, M, C, 1, ws2.data_ptr(), sp), "bn bwd")
done("bn fused fwd/bwd")
acc = torch.zeros(2 * C, dtype=torch.float64, device=dev)
nat.check(lib.dmp_bn_stats(x.data_ptr(), 1, M, C, acc.data_ptr(), sp), "bn stats")
done("bn stats")

# depthwise 3x3: forward (+statistics), data gradient (both strides), weight gradient
for stride in (1, 2):
    N, H, Cd = 2, 8, 16
    Ho = (H + 2 - 3) // stride + 1
    xd = torch.randn(N, H, H, Cd, device=dev).bfloat16(); w = torch.randn(Cd, 1, 3, 3, device=dev).bfloat16()
    yd = torch.empty(N, Ho, Ho, Cd, device=dev, dtype=torch.bfloat16); dyd = torch.randn_like(yd); dxd = torch.empty_like(xd)
    dw = torch.empty(Cd, 1, 3, 3, device=dev)
    sums = torch.zeros(int(lib.dmp_bn_sums_bytes(Cd)) // 8, dtype=torch.float64, device=dev)
    nat.check(lib.dmp_dwconv3x3_fwd(xd.data_ptr(), w.data_ptr(), 1, yd.data_ptr(), N, H, H, Cd, stride, sp), "dw fwd")
    nat.check(lib.dmp_dwconv3x3_fwd_colstats(xd.data_ptr(), w.data_ptr(), 1, yd.data_ptr(), N, H, H, Cd, stride, sums.data_ptr(), sp), "dw fwd stats")
    nat.check(lib.dmp_dwconv3x3_dgrad(dyd.data_ptr(), w.data_ptr(), 1, dxd.data_ptr(), N, H, H, Cd, stride, sp), "dw dgrad")
    nat.check(lib.dmp_dwconv3x3_wgrad(dyd.data_ptr(), xd.data_ptr(), dw.data_ptr(), N, H, H, Cd, stride, 0, sp), "dw wgrad")
    done("depthwise stride %d" % stride)

# stem convolution
N, H, Co = 2, 8, 32
xs = torch.randn(N, 3, H, H, device=dev); wsd = torch.randn(Co, 3, 3, 3, device=dev)
ys = torch.empty(N, H, H, Co, device=dev, dtype=torch.bfloat16); dws = torch.empty_like(wsd)
nat.check(lib.dmp_stem_conv_fwd(xs.data_ptr(), 0, wsd.data_ptr(), 0, ys.data_ptr(), N, 3, H, H, Co, 3, 3, 1, 1, sp), "stem fwd")
nat.check(lib.dmp_stem_conv_wgrad(ys.data_ptr(), xs.data_ptr(), 0, dws.data_ptr(), N, 3, H, H, Co, 3, 3, 1, 1, 0, sp), "stem wgrad")
done("stem conv")

# tcgen05 GEMM, three operand layouts (+ statistics epilogue)
Mg, Ng, Kg = 256, 64, 64
a = torch.randn(Mg, Kg, device=dev).bfloat16(); bw = torch.randn(Ng, Kg, device=dev).bfloat16()
d = torch.empty(Mg, Ng, device=dev, dtype=torch.bfloat16)
nat.check(lib.dmp_gemm_bf16(0, a.data_ptr(), Kg, bw.data_ptr(), Kg, d.data_ptr(), Ng, Mg, Ng, Kg, None, 0, 1, sp), "gemm NT")
done("gemm NT")
bt = bw.t().contiguous()
nat.check(lib.dmp_gemm_bf16(1, a.data_ptr(), Kg, bt.data_ptr(), Ng, d.data_ptr(), Ng, Mg, Ng, Kg, None, 0, 1, sp), "gemm NN")
done("gemm NN")
at = a.t().contiguous(); df = torch.zeros(Mg, Ng, device=dev)
nat.check(lib.dmp_gemm_bf16(2, at.data_ptr(), Mg, bt.data_ptr(), Ng, df.data_ptr(), Ng, Mg, Ng, Kg, None, 2, 0, sp), "gemm TN")
done("gemm TN (split-K atomics)")
sums = torch.zeros(int(lib.dmp_bn_sums_bytes(Ng)) // 8, dtype=torch.float64, device=dev)
nat.check(lib.dmp_gemm_bf16_colstats(a.data_ptr(), Kg, bw.data_ptr(), Kg, d.data_ptr(), Ng, Mg, Ng, Kg, sums.data_ptr(), sp), "gemm stats")
done("gemm + column statistics")
print("all sections ran", flush=True)
