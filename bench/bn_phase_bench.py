"""Where does a fused BatchNorm launch spend its time?  Each variant is captured 20x back to back in one CUDA graph (the
way the training step runs it) over rotating buffers larger than L2, and timed with CUDA events:
  fused   dmp_bn_fwd_fused (stats -> grid.sync -> apply, one cooperative launch)
  apply   dmp_bn_apply alone (what is left if the producer's epilogue delivered the statistics)
  stats   dmp_bn_stats alone (memset + reduction)
  bwd     dmp_bn_bwd_fused
  copy    a plain y = x * 1 elementwise torch kernel of the same bytes (floor)
Usage: python bench/bn_phase_bench.py
"""
import ctypes as C
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_model_parallel_b200 import _native  # noqa: E402

lib = _native.kernels()
dev = torch.device("cuda")
SHAPES = [(131072, 144), (131072, 32), (32768, 192), (8192, 384), (8192, 64), (2048, 960), (2048, 160)]
REPS = 20


def timed_graph(fn, iters=5):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn(0)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for r in range(REPS):
                fn(r)
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(iters):
            g.replay()
        e1.record(s)
        torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (iters * REPS)


out = []
for M, Cc in SHAPES:
    nbuf = max(2, min(REPS, int(300e6 // (M * Cc * 2)) + 1))
    xs = [torch.randn(M, Cc, device=dev).bfloat16() for _ in range(nbuf)]
    ys = [torch.empty_like(x) for x in xs]
    dxs = [torch.empty_like(x) for x in xs]
    gamma, beta = torch.ones(Cc, device=dev), torch.zeros(Cc, device=dev)
    rm, rv = torch.zeros(Cc, device=dev), torch.ones(Cc, device=dev)
    nbt = torch.zeros((), dtype=torch.int64, device=dev)
    mean, invstd = torch.empty(Cc, device=dev), torch.empty(Cc, device=dev)
    scale, shift = torch.ones(Cc, device=dev), torch.zeros(Cc, device=dev)
    dg, db = torch.empty(Cc, device=dev), torch.empty(Cc, device=dev)
    ws = torch.zeros(int(lib.dmp_bn_fused_ws_bytes(Cc)) // 8, dtype=torch.float64, device=dev)
    wsb = torch.zeros_like(ws)
    accum = torch.zeros(2 * Cc, dtype=torch.float64, device=dev)

    def sp():
        return torch.cuda.current_stream().cuda_stream

    def fused(r):
        _native.check(lib.dmp_bn_fwd_fused(xs[r % nbuf].data_ptr(), 1, None, ys[r % nbuf].data_ptr(), M, Cc, gamma.data_ptr(), beta.data_ptr(),
                                           rm.data_ptr(), rv.data_ptr(), nbt.data_ptr(), 0.1, 1e-5, mean.data_ptr(), invstd.data_ptr(),
                                           ws.data_ptr(), 1, sp()), "fused")

    def apply(r):
        _native.check(lib.dmp_bn_apply(xs[r % nbuf].data_ptr(), 1, scale.data_ptr(), shift.data_ptr(), None, ys[r % nbuf].data_ptr(), M, Cc, 1,
                                       sp()), "apply")

    def stats(r):
        _native.check(lib.dmp_bn_stats(xs[r % nbuf].data_ptr(), 1, M, Cc, accum.data_ptr(), sp()), "stats")

    def bwd(r):
        _native.check(lib.dmp_bn_bwd_fused(ys[r % nbuf].data_ptr(), None, xs[r % nbuf].data_ptr(), 1, mean.data_ptr(), invstd.data_ptr(),
                                           gamma.data_ptr(), beta.data_ptr(), dxs[r % nbuf].data_ptr(), None, dg.data_ptr(), db.data_ptr(), 0,
                                           M, Cc, 1, wsb.data_ptr(), sp()), "bwd")

    def copy(r):
        torch.mul(xs[r % nbuf], 1.0, out=ys[r % nbuf])

    fused(0)
    rec = {"M": M, "C": Cc, "MB": round(M * Cc * 2 / 1e6, 2)}
    for name, fn in (("fused", fused), ("apply", apply), ("stats", stats), ("bwd", bwd), ("copy", copy)):
        rec[name + "_us"] = round(timed_graph(fn), 2)
    out.append(rec)
    print(json.dumps(rec), flush=True)
