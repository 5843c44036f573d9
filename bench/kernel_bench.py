"""Per-kernel roofline microbenchmark on the MobileNetV2 (B=128) layer shapes.
Times each kernel with CUDA events (L2 flushed between iterations by writing a 256 MB buffer), reports achieved GB/s of
algorithmic bytes and the fraction of the measured HBM copy bandwidth (MEASURED_PEAKS.json, fallback 6650 GB/s)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_model_parallel_b200 import _native as nat

lib = nat.kernels()
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]; SRC = "measured"
except Exception:
    PEAK, SRC = 6650.0, "fallback"
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
sp = torch.cuda.current_stream().cuda_stream


def timeit(fn, iters=10):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        lib.dmp_l2_flush(flush.data_ptr(), flush.numel(), sp)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


rows = []
def report(name, us, bytes_):
    gbs = bytes_ / us / 1e3
    rows.append((name, us, gbs, gbs / PEAK))
    print("%-44s %9.1f us %9.0f GB/s  %5.2f of %s HBM copy peak" % (name, us, gbs, gbs / PEAK, SRC))


B = 128
# (H, Cin, planes, Cout, stride) of the 17 blocks
cfg, blocks, cin, H = [(1, 16, 1, 1), (6, 24, 2, 1), (6, 32, 3, 2), (6, 64, 4, 2), (6, 96, 3, 1), (6, 160, 3, 2), (6, 320, 1, 1)], [], 32, 32
for e, co, n, s in cfg:
    for st in [s] + [1] * (n - 1):
        blocks.append((H, cin, e * cin, co, st)); cin = co; H = H // st
sel = [blocks[i] for i in (1, 3, 6, 10, 13, 16)]
tot = {"gemm_fwd": 0.0, "gemm_dgrad": 0.0, "gemm_wgrad": 0.0, "dw_fwd": 0.0, "dw_dgrad": 0.0, "dw_wgrad": 0.0, "bn_stats": 0.0,
       "bn_apply": 0.0, "bn_bwd_reduce": 0.0, "bn_bwd_dx": 0.0}
for (Hh, ci, pl, co, st) in blocks:
    M = B * Hh * Hh
    Ho = Hh // st
    Mo = B * Ho * Ho
    for (m, k, n, tag) in ((M, ci, pl, "expand"), (Mo, pl, co, "project")):
        a = torch.randn(m, k, device="cuda").bfloat16(); w = torch.randn(n, k, device="cuda").bfloat16()
        y = torch.empty(m, n, device="cuda", dtype=torch.bfloat16); dw = torch.zeros(n, k, device="cuda")
        by = 2 * (m * k + n * k + m * n)
        t = timeit(lambda: lib.dmp_gemm_bf16(0, a.data_ptr(), k, w.data_ptr(), k, y.data_ptr(), n, m, n, k, None, 0, 1, sp)); tot["gemm_fwd"] += t
        if (Hh, ci, pl, co, st) in sel: report("gemm fwd   %s M=%d K=%d N=%d" % (tag, m, k, n), t, by)
        dx = torch.empty(m, k, device="cuda", dtype=torch.bfloat16)
        t = timeit(lambda: lib.dmp_gemm_bf16(1, y.data_ptr(), n, w.data_ptr(), k, dx.data_ptr(), k, m, k, n, None, 0, 1, sp)); tot["gemm_dgrad"] += t
        if (Hh, ci, pl, co, st) in sel: report("gemm dgrad %s M=%d K=%d N=%d" % (tag, m, n, k), t, by)
        t = timeit(lambda: lib.dmp_gemm_bf16(2, y.data_ptr(), n, a.data_ptr(), k, dw.data_ptr(), k, n, k, m, None, 2, 0, sp)); tot["gemm_wgrad"] += t
        if (Hh, ci, pl, co, st) in sel: report("gemm wgrad %s M=%d N=%d K=%d" % (tag, n, k, m), t, 2 * (m * k + m * n) + 4 * n * k)
    x = torch.randn(B, Hh, Hh, pl, device="cuda").bfloat16(); wd = torch.randn(pl, 9, device="cuda").bfloat16()
    yo = torch.empty(B, Ho, Ho, pl, device="cuda", dtype=torch.bfloat16); dxx = torch.empty_like(x); dwd = torch.empty(pl, 9, device="cuda")
    t = timeit(lambda: lib.dmp_dwconv3x3_fwd(x.data_ptr(), wd.data_ptr(), 1, yo.data_ptr(), B, Hh, Hh, pl, st, sp)); tot["dw_fwd"] += t
    if (Hh, ci, pl, co, st) in sel: report("dw fwd     C=%d H=%d s=%d" % (pl, Hh, st), t, 2 * (x.numel() + yo.numel()))
    t = timeit(lambda: lib.dmp_dwconv3x3_dgrad(yo.data_ptr(), wd.data_ptr(), 1, dxx.data_ptr(), B, Hh, Hh, pl, st, sp)); tot["dw_dgrad"] += t
    if (Hh, ci, pl, co, st) in sel: report("dw dgrad   C=%d H=%d s=%d" % (pl, Hh, st), t, 2 * (x.numel() + yo.numel()))
    t = timeit(lambda: lib.dmp_dwconv3x3_wgrad(yo.data_ptr(), x.data_ptr(), dwd.data_ptr(), B, Hh, Hh, pl, st, 0, sp)); tot["dw_wgrad"] += t
    if (Hh, ci, pl, co, st) in sel: report("dw wgrad   C=%d H=%d s=%d" % (pl, Hh, st), t, 2 * (x.numel() + yo.numel()))
    for (m, c) in ((M, pl), (Mo, pl), (Mo, co)):
        xb = torch.randn(m, c, device="cuda").bfloat16(); yb = torch.empty_like(xb); dyb = torch.randn(m, c, device="cuda").bfloat16()
        acc = torch.empty(2 * c, device="cuda", dtype=torch.float64); mean = torch.zeros(c, device="cuda"); inv = torch.ones(c, device="cuda")
        t = timeit(lambda: lib.dmp_bn_stats(xb.data_ptr(), 1, m, c, acc.data_ptr(), sp)); tot["bn_stats"] += t
        if (Hh, ci, pl, co, st) in sel: report("bn stats      M=%d C=%d" % (m, c), t, 2 * m * c)
        t = timeit(lambda: lib.dmp_bn_apply(xb.data_ptr(), 1, inv.data_ptr(), mean.data_ptr(), None, yb.data_ptr(), m, c, 1, sp)); tot["bn_apply"] += t
        if (Hh, ci, pl, co, st) in sel: report("bn apply      M=%d C=%d" % (m, c), t, 4 * m * c)
        t = timeit(lambda: lib.dmp_bn_bwd_reduce(dyb.data_ptr(), yb.data_ptr(), xb.data_ptr(), 1, mean.data_ptr(), inv.data_ptr(), m, c, 1, acc.data_ptr(), sp)); tot["bn_bwd_reduce"] += t
        if (Hh, ci, pl, co, st) in sel: report("bn bwd reduce M=%d C=%d" % (m, c), t, 6 * m * c)
        t = timeit(lambda: lib.dmp_bn_bwd_dx(dyb.data_ptr(), yb.data_ptr(), xb.data_ptr(), 1, mean.data_ptr(), inv.data_ptr(), inv.data_ptr(), acc.data_ptr(), float(m), yb.data_ptr(), None, m, c, 1, sp)); tot["bn_bwd_dx"] += t
        if (Hh, ci, pl, co, st) in sel: report("bn bwd dx     M=%d C=%d" % (m, c), t, 8 * m * c)
print("---- per-step totals over the 17 blocks (cold L2), us:")
for k, v in tot.items():
    print("  %-14s %9.1f" % (k, v))
print("  %-14s %9.1f" % ("sum", sum(tot.values())))
