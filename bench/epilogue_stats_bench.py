"""Producer-side cost of BatchNorm statistics in the convolution epilogue vs the BatchNorm-side saving.
Every variant is captured 20x in one CUDA graph over rotating buffers and timed with CUDA events (see bn_phase_bench.py)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_model_parallel_b200 import _native  # noqa: E402
from bn_phase_bench import timed_graph, REPS  # noqa: E402

lib = _native.kernels()
dev = torch.device("cuda")


def sp():
    return torch.cuda.current_stream().cuda_stream


GEMMS = [(131072, 144, 24), (131072, 24, 144), (32768, 192, 32), (8192, 384, 64), (8192, 64, 384), (2048, 960, 160), (2048, 160, 960)]
for M, N, K in GEMMS:
    nbuf = max(2, min(REPS, int(300e6 // (M * (N + K) * 2)) + 1))
    a = [torch.randn(M, K, device=dev).bfloat16() for _ in range(nbuf)]
    y = [torch.empty(M, N, device=dev, dtype=torch.bfloat16) for _ in range(nbuf)]
    o = [torch.empty(M, N, device=dev, dtype=torch.bfloat16) for _ in range(nbuf)]
    b = torch.randn(N, K, device=dev).bfloat16()
    sums = torch.zeros(int(lib.dmp_bn_sums_bytes(N)) // 8, dtype=torch.float64, device=dev)
    gamma, beta = torch.ones(N, device=dev), torch.zeros(N, device=dev)
    rm, rv = torch.zeros(N, device=dev), torch.ones(N, device=dev)
    nbt = torch.zeros((), dtype=torch.int64, device=dev)
    mean, invstd = torch.empty(N, device=dev), torch.empty(N, device=dev)
    ws = torch.zeros(int(lib.dmp_bn_fused_ws_bytes(N)) // 8, dtype=torch.float64, device=dev)

    def gemm(r):
        _native.check(lib.dmp_gemm_bf16(0, a[r % nbuf].data_ptr(), K, b.data_ptr(), K, y[r % nbuf].data_ptr(), N, M, N, K, None, 0, 1, sp()), "gemm")

    def gemm_stats(r):
        _native.check(lib.dmp_gemm_bf16_colstats(a[r % nbuf].data_ptr(), K, b.data_ptr(), K, y[r % nbuf].data_ptr(), N, M, N, K, sums.data_ptr(), sp()), "gemm+stats")

    def apply_sums(r):
        _native.check(lib.dmp_bn_apply_sums(y[r % nbuf].data_ptr(), 1, None, o[r % nbuf].data_ptr(), M, N, gamma.data_ptr(), beta.data_ptr(), rm.data_ptr(),
                                            rv.data_ptr(), nbt.data_ptr(), 0.1, 1e-5, mean.data_ptr(), invstd.data_ptr(), sums.data_ptr(), 1, sp()), "apply")

    def fused(r):
        _native.check(lib.dmp_bn_fwd_fused(y[r % nbuf].data_ptr(), 1, None, o[r % nbuf].data_ptr(), M, N, gamma.data_ptr(), beta.data_ptr(), rm.data_ptr(),
                                           rv.data_ptr(), nbt.data_ptr(), 0.1, 1e-5, mean.data_ptr(), invstd.data_ptr(), ws.data_ptr(), 1, sp()), "fused")

    def pair_old(r):
        gemm(r); fused(r)

    def pair_new(r):
        gemm_stats(r); apply_sums(r)

    rec = {"op": "pw", "M": M, "N": N, "K": K}
    for name, fn in (("gemm", gemm), ("gemm_stats", gemm_stats), ("bn_fused", fused), ("bn_apply_sums", apply_sums), ("pair_old", pair_old), ("pair_new", pair_new)):
        rec[name + "_us"] = round(timed_graph(fn), 2)
    print(json.dumps(rec), flush=True)

DWS = [(128, 32, 144, 1), (128, 32, 144, 2), (128, 16, 192, 1), (128, 8, 384, 1), (128, 4, 960, 1)]
for Nb, H, C, s in DWS:
    Ho = (H + 2 - 3) // s + 1
    nbuf = max(2, min(REPS, int(300e6 // (Nb * H * H * C * 2)) + 1))
    x = [torch.randn(Nb, H, H, C, device=dev).bfloat16() for _ in range(nbuf)]
    y = [torch.empty(Nb, Ho, Ho, C, device=dev, dtype=torch.bfloat16) for _ in range(nbuf)]
    w = torch.randn(C, 1, 3, 3, device=dev).bfloat16()
    sums = torch.zeros(int(lib.dmp_bn_sums_bytes(C)) // 8, dtype=torch.float64, device=dev)

    def dw(r):
        _native.check(lib.dmp_dwconv3x3_fwd(x[r % nbuf].data_ptr(), w.data_ptr(), 1, y[r % nbuf].data_ptr(), Nb, H, H, C, s, sp()), "dw")

    def dw_stats(r):
        _native.check(lib.dmp_dwconv3x3_fwd_colstats(x[r % nbuf].data_ptr(), w.data_ptr(), 1, y[r % nbuf].data_ptr(), Nb, H, H, C, s, sums.data_ptr(), sp()), "dw+stats")

    rec = {"op": "dw", "N": Nb, "H": H, "C": C, "stride": s}
    for name, fn in (("dw", dw), ("dw_stats", dw_stats)):
        rec[name + "_us"] = round(timed_graph(fn), 2)
    print(json.dumps(rec), flush=True)
