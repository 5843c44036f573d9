"""Device timeline of the captured training step (CUPTI activity records through torch.profiler): per-kernel durations INSIDE the
graph replay (warm L2, real overlap with the side streams), idle gaps on the main stream, and the critical-path share per kernel.

    python bench/step_trace.py --model mobilenetv2 --batch 128 > profiles/step_trace_b128.txt
"""
import argparse
import os
import re
import sys
from collections import defaultdict

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="mobilenetv2")
ap.add_argument("--batch", type=int, default=128)
ap.add_argument("--replays", type=int, default=3)
ap.add_argument("--list", action="store_true")
args = ap.parse_args()

from distributed_model_parallel_b200.models import build_model
from distributed_model_parallel_b200.optim import FlatFusedSGD
from distributed_model_parallel_b200.train.engine import TrainStep

dev = torch.device("cuda", 0)
torch.manual_seed(0)
shape, nc = ((3, 224, 224), 1000) if args.model == "resnet50" else ((3, 32, 32), 10)
model = build_model(args.model, num_classes=nc).to(dev)
opt = FlatFusedSGD(model.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
x = torch.randn((args.batch,) + shape, device=dev)
y = torch.randint(0, nc, (args.batch,), device=dev)
step = TrainStep(model, opt, use_graph=True)
for _ in range(5):
    step(x, y)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    step(x, y)
e1.record()
torch.cuda.synchronize()
ms_plain = e0.elapsed_time(e1) / 20
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(args.replays):
        step(x, y)
    torch.cuda.synchronize()


def short(name):
    name = re.sub(r"\(.*$", "", name)
    name = re.sub(r"^void ", "", name)
    return re.sub(r"\(anonymous namespace\)::", "", name)


evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and "emcpy" not in e.name and "emset" not in e.name]
evs.sort(key=lambda e: e.time_range.start)
# split into replays by the largest gaps
per = len(evs) // args.replays
last = evs[-per:]
t0 = last[0].time_range.start
t1 = max(e.time_range.end for e in last)
streams = defaultdict(list)
for e in last:
    streams[getattr(e, "stream", None) if hasattr(e, "stream") else 0].append(e)
tot, cnt = defaultdict(float), defaultdict(int)
for e in last:
    tot[short(e.name)] += e.time_range.end - e.time_range.start
    cnt[short(e.name)] += 1
span = t1 - t0
busy = sum(tot.values())
# union of busy intervals (any stream)
iv = sorted((e.time_range.start, e.time_range.end) for e in last)
union, cur_s, cur_e = 0.0, iv[0][0], iv[0][1]
for s, e_ in iv[1:]:
    if s > cur_e:
        union += cur_e - cur_s
        cur_s, cur_e = s, e_
    else:
        cur_e = max(cur_e, e_)
union += cur_e - cur_s
print("# %s batch %d: %.3f ms/step by CUDA events (not under the profiler); traced replay: %d kernels, span %.1f us, "
      "sum of kernel durations %.1f us, time with at least one kernel running %.1f us (idle %.1f us)" % (
          args.model, args.batch, ms_plain, len(last), span, busy, union, span - union))
print("%-78s %6s %10s %7s %9s" % ("kernel", "count", "total_us", "share", "us/launch"))
for k in sorted(tot, key=lambda k: -tot[k]):
    print("%-78s %6d %10.1f %6.1f%% %9.2f" % (k[:78], cnt[k], tot[k], 100.0 * tot[k] / busy, tot[k] / cnt[k]))
if args.list:
    print("\n# kernels of one replay in start order: start_us, dur_us, name")
    for e in last:
        print("%9.1f %8.2f %s" % (e.time_range.start - t0, e.time_range.end - e.time_range.start, short(e.name)[:100]))
