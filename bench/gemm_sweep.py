"""GEMM forward (NT, bf16 out) time vs shape, CUDA-graph timed: which of rows / K bytes / output bytes / tile count sets the time."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_model_parallel_b200 import _native  # noqa: E402
from bn_phase_bench import timed_graph, REPS  # noqa: E402

lib = _native.kernels()
dev = torch.device("cuda")
shapes = [tuple(int(v) for v in a.split("x")) for a in sys.argv[1:]] or [(131072, 144, 24), (131072, 144, 64), (131072, 128, 24),
                                                                        (131072, 128, 64), (131072, 64, 64), (131072, 256, 64),
                                                                        (131072, 128, 128), (65536, 128, 64), (32768, 128, 64)]
for M, N, K in shapes:
    nbuf = max(2, min(REPS, int(300e6 // (M * (N + K) * 2)) + 1))
    a = [torch.randn(M, K, device=dev).bfloat16() for _ in range(nbuf)]
    y = [torch.empty(M, N, device=dev, dtype=torch.bfloat16) for _ in range(nbuf)]
    b = torch.randn(N, K, device=dev).bfloat16()

    def gemm(r):
        _native.check(lib.dmp_gemm_bf16(0, a[r % nbuf].data_ptr(), K, b.data_ptr(), K, y[r % nbuf].data_ptr(), N, M, N, K, None, 0, 1,
                                        torch.cuda.current_stream().cuda_stream), "gemm")

    us = timed_graph(gemm)
    mb = (M * K + M * N) * 2 / 1e6
    print(json.dumps({"M": M, "N": N, "K": K, "us": round(us, 2), "MB": round(mb, 1), "TBps": round(mb / us, 2),
                      "tiles": -(-M // 128) * -(-N // (128 if N > 64 else 64 if N > 32 else 32))}), flush=True)
