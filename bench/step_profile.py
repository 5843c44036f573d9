"""One eager training step of a model inside a cudaProfilerStart/Stop window, for a per-kernel launch list:

    ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum \
        --clock-control none --csv --log-file gpurun_out/launches.csv python bench/step_profile.py --model mobilenetv2 --batch 128
    python tools/launch_summary.py gpurun_out/launches.csv > profiles/launches_summary.txt

Without ncu it prints the eager and graph-replay step times (CUDA events), which is the check that the program exits 0 first."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="mobilenetv2")
ap.add_argument("--batch", type=int, default=128)
ap.add_argument("--steps", type=int, default=1)
ap.add_argument("--time", action="store_true", help="also time eager and graph replay (not under ncu)")
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
step = TrainStep(model, opt, use_graph=False)
for _ in range(3):
    step(x, y)
torch.cuda.synchronize()
torch.cuda.profiler.start()
for _ in range(args.steps):
    step(x, y)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
if args.time:
    def timed(s, n=20):
        for _ in range(3):
            s(x, y)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            s(x, y)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n

    eager = timed(step)
    g = TrainStep(model, opt, use_graph=True)
    print("%s batch %d: eager %.3f ms/step, graph %.3f ms/step, %d launches/step" % (args.model, args.batch, eager, timed(g), g.launches_per_iteration))
