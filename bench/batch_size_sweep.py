"""The reference's batch-size experiments, the part that can be reproduced offline: throughput and short-run trainability of
MobileNetV2 (with BatchNorm) and MobileNetV2_nobn as the batch grows 128 -> 2048 with the learning rate scaled linearly
(lr = 0.05 * batch / 128, the reference's 0.05 / 0.1 / 0.2 / 0.4 at 128 / 256 / 512 / 1024, /root/reference/Readme.md:159-211;
its no-BN large-batch runs, Readme.md:161-176).  The reference reports final CIFAR-10 accuracies after 90-100 epochs; without
the dataset this driver reports, per batch size: images/s of the captured training step (device-timed), ms/step, and the loss
and top-1 accuracy on a learnable synthetic set (data/synthetic.py: the label is a function of the image) after a fixed number
of SAMPLES, so the rows are comparable the way the reference's curves are.

    python bench/batch_size_sweep.py [--models mobilenetv2,mobilenetv2_nobn] [--batches 128,256,512,1024,2048] [--samples 65536]

Prints one JSON line per (model, batch) and a table; --out writes the rows as JSON."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

ap = argparse.ArgumentParser()
ap.add_argument("--models", default="mobilenetv2,mobilenetv2_nobn")
ap.add_argument("--batches", default="128,256,512,1024,2048")
ap.add_argument("--samples", type=int, default=65536, help="training samples per configuration (the same for every batch size)")
ap.add_argument("--base-lr", type=float, default=0.05, help="learning rate at batch 128, scaled linearly with the batch")
ap.add_argument("--timing-steps", type=int, default=20)
ap.add_argument("--out", default="")
args = ap.parse_args()

from distributed_model_parallel_b200.data.synthetic import SyntheticImages
from distributed_model_parallel_b200.models import build_model
from distributed_model_parallel_b200.optim import FlatFusedSGD
from distributed_model_parallel_b200.train.engine import TrainStep

dev = torch.device("cuda", 0)
rows = []
for name in args.models.split(","):
    for B in [int(b) for b in args.batches.split(",")]:
        torch.manual_seed(0)
        ds = SyntheticImages(length=max(args.samples, B), seed=0)
        pool, labels = ds.pool.to(dev), ds.labels.to(dev)
        model = build_model(name, num_classes=10).to(dev)
        # without BatchNorm the reference needed a much smaller step to train at all (Readme.md:161-176: 0.2 at 256, 0.8 at 2048)
        lr = args.base_lr * B / 128 * (0.25 if name.endswith("nobn") else 1.0)
        opt = FlatFusedSGD(model.parameters(), lr=lr, momentum=0.9, weight_decay=1e-4)
        step = TrainStep(model, opt, use_graph=True)
        n_steps = max(1, args.samples // B)
        idx = torch.arange(args.samples, device=dev) % pool.shape[0]
        stats = None
        for s in range(n_steps):
            sel = idx[s * B:(s + 1) * B]
            stats = step(pool[sel], labels[s * B:(s + 1) * B])
        torch.cuda.synchronize()
        loss, acc = float(stats[0]), float(stats[1]) / B if stats.numel() > 1 else float("nan")
        x, y = pool[idx[:B]], labels[:B]
        for _ in range(3):
            step(x, y)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.timing_steps):
            step(x, y)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.timing_steps
        row = {"model": name, "batch": B, "lr": lr, "steps_trained": n_steps, "samples_trained": n_steps * B, "final_loss": loss,
               "final_batch_top1": acc, "ms_per_step": ms, "images_per_s": B / ms * 1e3, "finite": loss == loss}
        rows.append(row)
        print(json.dumps(row), flush=True)
        del model, opt, step
        torch.cuda.empty_cache()
print("\n%-18s %6s %7s %10s %12s %11s %10s" % ("model", "batch", "lr", "ms/step", "images/s", "final loss", "top-1"))
for r in rows:
    print("%-18s %6d %7.3f %10.3f %12.0f %11.4f %10.3f" % (r["model"], r["batch"], r["lr"], r["ms_per_step"], r["images_per_s"], r["final_loss"],
                                                            r["final_batch_top1"]))
if args.out:
    json.dump(rows, open(args.out, "w"), indent=1)
