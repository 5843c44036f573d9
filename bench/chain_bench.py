"""In-graph per-kernel time of the step's kernels on the MobileNetV2 layer shapes: REPS back-to-back launches of one kernel are
captured in a CUDA graph and replayed, so the number is what the kernel costs inside the captured training step (launch
latency hidden by the graph, operands warm in the 126 MB L2 when they fit) -- unlike bench/kernel_bench.py, whose event
pair around a single launch has a ~8 us floor.  Prints us per launch and the effective bandwidth of the algorithmic bytes.

    python bench/chain_bench.py [--batch 128] [--only bn,gemm,dw,misc]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_model_parallel_b200 import _native as nat

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=128)
ap.add_argument("--reps", type=int, default=20)
ap.add_argument("--only", default="bn,gemm,dw,misc")
args = ap.parse_args()
lib = nat.kernels()
B = args.batch
REPS = args.reps
only = set(args.only.split(","))
dev = torch.device("cuda", 0)


def chain(fn, reps=REPS):
    """fn(sp) launches the kernel once on raw stream sp.  Returns us per launch inside a graph replay."""
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(2):
            fn(s.cuda_stream)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        for _ in range(reps):
            fn(s.cuda_stream)
    for _ in range(2):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / 5 / reps


tot = {}


def report(group, name, us, nbytes):
    tot[group] = tot.get(group, 0.0) + us
    print("%-14s %-40s %8.2f us %8.0f GB/s" % (group, name, us, nbytes / us / 1e3), flush=True)


def ck(rc):
    if rc != 0:
        raise RuntimeError(lib.dmp_last_error().decode())


cfg, blocks, cin, H = [(1, 16, 1, 1), (6, 24, 2, 1), (6, 32, 3, 2), (6, 64, 4, 2), (6, 96, 3, 1), (6, 160, 3, 2), (6, 320, 1, 1)], [], 32, 32
for e, co, n, s in cfg:
    for st in [s] + [1] * (n - 1):
        blocks.append((H, cin, e * cin, co, st)); cin = co; H = H // st

bf = torch.bfloat16
for (Hh, ci, pl, co, st) in blocks:
    M = B * Hh * Hh
    Ho = Hh // st
    Mo = B * Ho * Ho
    tag = "H%d %d->%d->%d s%d" % (Hh, ci, pl, co, st)
    if "gemm" in only:
        for (m, k, n, nm) in ((M, ci, pl, "expand"), (Mo, pl, co, "project")):
            if nm == "expand" and pl == ci:
                continue
            a = torch.randn(m, k, device=dev).to(bf); w = torch.randn(n, k, device=dev).to(bf)
            y = torch.empty(m, n, device=dev, dtype=bf); dw = torch.zeros(n, k, device=dev); dx = torch.empty(m, k, device=dev, dtype=bf)
            sums = torch.zeros(int(lib.dmp_bn_sums_bytes(n)) // 8, dtype=torch.float64, device=dev)
            by = 2 * (m * k + n * k + m * n)
            report("gemm_fwd", "%s %s M=%d K=%d N=%d" % (tag, nm, m, k, n),
                   chain(lambda sp: ck(lib.dmp_gemm_bf16_colstats(a.data_ptr(), k, w.data_ptr(), k, y.data_ptr(), n, m, n, k, sums.data_ptr(), sp))), by)
            report("gemm_dgrad", "%s %s" % (tag, nm),
                   chain(lambda sp: ck(lib.dmp_gemm_bf16(1, y.data_ptr(), n, w.data_ptr(), k, dx.data_ptr(), k, m, k, n, None, 0, 1, sp))), by)
            report("gemm_wgrad", "%s %s" % (tag, nm),
                   chain(lambda sp: ck(lib.dmp_gemm_bf16(2, y.data_ptr(), n, a.data_ptr(), k, dw.data_ptr(), k, n, k, m, None, 2, 0, sp))), 2 * (m * k + m * n) + 4 * n * k)
    if "dw" in only:
        x = torch.randn(B, Hh, Hh, pl, device=dev).to(bf); wd = torch.randn(pl, 9, device=dev).to(bf)
        yo = torch.empty(B, Ho, Ho, pl, device=dev, dtype=bf); dxx = torch.empty_like(x); dwd = torch.zeros(pl, 9, device=dev)
        sums = torch.zeros(int(lib.dmp_bn_sums_bytes(pl)) // 8, dtype=torch.float64, device=dev)
        by = 2 * (x.numel() + yo.numel())
        report("dw_fwd", tag, chain(lambda sp: ck(lib.dmp_dwconv3x3_fwd_colstats(x.data_ptr(), wd.data_ptr(), 1, yo.data_ptr(), B, Hh, Hh, pl, st, sums.data_ptr(), sp))), by)
        report("dw_dgrad", tag, chain(lambda sp: ck(lib.dmp_dwconv3x3_dgrad(yo.data_ptr(), wd.data_ptr(), 1, dxx.data_ptr(), B, Hh, Hh, pl, st, sp))), by)
        report("dw_wgrad", tag, chain(lambda sp: ck(lib.dmp_dwconv3x3_wgrad(yo.data_ptr(), x.data_ptr(), dwd.data_ptr(), B, Hh, Hh, pl, st, 1, sp))), by)
    if "bn" in only:
        for (m, c, relu) in ((M, pl, 1), (Mo, pl, 1), (Mo, co, 0)):
            if c == ci and m == M and pl == ci:
                continue
            xb = torch.randn(m, c, device=dev).to(bf); yb = torch.empty_like(xb); dyb = torch.randn(m, c, device=dev).to(bf); dxb = torch.empty_like(xb)
            g = torch.ones(c, device=dev); b_ = torch.zeros(c, device=dev); rm = torch.zeros(c, device=dev); rv = torch.ones(c, device=dev)
            nbt = torch.zeros((), dtype=torch.int64, device=dev); mean = torch.zeros(c, device=dev); inv = torch.ones(c, device=dev)
            dg = torch.zeros(c, device=dev); db = torch.zeros(c, device=dev)
            sums = torch.zeros(int(lib.dmp_bn_sums_bytes(c)) // 8, dtype=torch.float64, device=dev)
            ws = torch.zeros(int(lib.dmp_bn_fused_ws_bytes(c)) // 8, dtype=torch.float64, device=dev)
            report("bn_apply", "%s M=%d C=%d" % (tag, m, c),
                   chain(lambda sp: ck(lib.dmp_bn_apply_sums(xb.data_ptr(), 1, None, yb.data_ptr(), m, c, g.data_ptr(), b_.data_ptr(), rm.data_ptr(), rv.data_ptr(),
                                                             nbt.data_ptr(), 0.1, 1e-5, mean.data_ptr(), inv.data_ptr(), sums.data_ptr(), relu, sp))), 4 * m * c)
            report("bn_apply_plain", "%s M=%d C=%d (scale/shift given)" % (tag, m, c),
                   chain(lambda sp: ck(lib.dmp_bn_apply(xb.data_ptr(), 1, inv.data_ptr(), mean.data_ptr(), None, yb.data_ptr(), m, c, relu, sp))), 4 * m * c)
            report("bn_bwd", "%s M=%d C=%d" % (tag, m, c),
                   chain(lambda sp: ck(lib.dmp_bn_bwd_fused(dyb.data_ptr(), None, xb.data_ptr(), 1, mean.data_ptr(), inv.data_ptr(), g.data_ptr(), b_.data_ptr(),
                                                            dxb.data_ptr(), None, dg.data_ptr(), db.data_ptr(), 1, m, c, relu, ws.data_ptr(), sp))), 6 * m * c)
if "misc" in only:
    x = torch.randn(B, 3, 32, 32, device=dev); w = torch.randn(32, 3, 3, 3, device=dev); y = torch.empty(B, 32, 32, 32, device=dev, dtype=bf)
    dy = torch.randn(B, 32, 32, 32, device=dev).to(bf); dw = torch.zeros(32, 27, device=dev)
    report("misc", "stem fwd", chain(lambda sp: ck(lib.dmp_stem_conv_fwd(x.data_ptr(), 0, w.data_ptr(), 0, y.data_ptr(), B, 3, 32, 32, 32, 3, 3, 1, 1, sp))), x.numel() * 4 + y.numel() * 2)
    report("misc", "stem wgrad", chain(lambda sp: ck(lib.dmp_stem_conv_wgrad(dy.data_ptr(), x.data_ptr(), 0, dw.data_ptr(), B, 3, 32, 32, 32, 3, 3, 1, 1, 1, sp))), x.numel() * 4 + y.numel() * 2)
    a = torch.randn(64 << 20, device=dev).view(torch.uint8)
    for mb in (1, 4, 16):
        n = mb << 20
        src, dst = a[:n], a[n:2 * n]
        report("misc", "add_act %d MB (copy-like floor)" % mb, chain(lambda sp: ck(lib.dmp_add_act(src.data_ptr(), None, dst.data_ptr(), n // 2, 0, sp))), 2 * n)
print("---- totals over the 17 blocks, us per step if nothing overlapped:")
for k, v in tot.items():
    print("  %-12s %9.1f" % (k, v))
print("  %-12s %9.1f" % ("sum", sum(tot.values())))
