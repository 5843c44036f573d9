#!/usr/bin/env python
"""Headline benchmark: training throughput (images/s) of MobileNetV2 on CIFAR-10-shaped synthetic data,
per-GPU batch 128 (global 128*N: the reference's published 256 @ 2 GPUs / 512 @ 4 GPUs configs, weak scaling).

    python bench.py --gpus N --steps K --warmup W                   # this framework
    python bench.py --impl reference --gpus N --steps K --warmup W  # unmodified reference model + torch DP/DDP

For N > 1 launch with ``python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1
--master-port P bench.py --gpus N ...`` (one rank per GPU).  Rank 0 prints ONE JSON line.

ours:       DistributedDataParallel (peer-memory all-reduce fused with SGD) around the sm_100a kernel model,
            whole step replayed from a CUDA graph; bf16 compute, fp32 master weights.
reference:  the reference's own ``model.MobileNetV2`` (from baseline/_ref, unmodified) trained exactly like its
            data_parallel.py does: torch.nn.DataParallel over N GPUs driven by rank 0, CrossEntropyLoss,
            SGD(momentum 0.9, wd 1e-4), cudnn.benchmark, fp32.  torch DistributedDataParallel (NCCL) on the same
            model is measured too and reported as ``torch_ddp_value``.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PUBLISHED = {2: 705.0, 4: 1293.0}  # nn.DataParallel img/s derived from /root/reference/Readme.md:283-292


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="mobilenetv2")
    ap.add_argument("--batch-per-gpu", type=int, default=128)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--comm", default="peer", choices=["peer", "torch"])
    ap.add_argument("--ref-dtype", default="fp32", choices=["fp32", "bf16"])
    ap.add_argument("--lr", type=float, default=0.1)
    return ap.parse_args()


class ClockSampler:
    """Samples SM clocks / throttle reasons with nvidia-smi while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); power.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


def setup_dist(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return world, rank, local


def barrier_sync(world):
    import torch
    import torch.distributed as dist

    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(value: float, world: int) -> float:
    import torch
    import torch.distributed as dist

    if world <= 1:
        return value
    t = torch.tensor([value], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def input_shape(model_name):
    return (3, 224, 224) if model_name == "resnet50" else (3, 32, 32)


def num_classes(model_name):
    return 1000 if model_name == "resnet50" else 10


# ---------------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    from distributed_model_parallel_b200 import _native
    from distributed_model_parallel_b200.data import pinned_batches
    from distributed_model_parallel_b200.models import build_model
    from distributed_model_parallel_b200.ops import functional as OF
    from distributed_model_parallel_b200.optim import FlatFusedSGD
    from distributed_model_parallel_b200.parallel import DistributedDataParallel
    from distributed_model_parallel_b200.train.engine import TrainStep

    world, rank, local = setup_dist(args)
    dev = torch.device("cuda", local)
    torch.manual_seed(1234)
    B = args.batch_per_gpu
    model = build_model(args.model, num_classes=num_classes(args.model)).to(dev)
    sgd = dict(lr=args.lr, momentum=0.9, weight_decay=1e-4)
    if world > 1:
        model = DistributedDataParallel(model, device_ids=[local], comm_backend=args.comm,
                                        fused_sgd=sgd if args.comm == "peer" else None)
        opt = model.optimizer if args.comm == "peer" else FlatFusedSGD(model.parameters(), **sgd)
    else:
        opt = FlatFusedSGD(model.parameters(), **sgd)
    step = TrainStep(model, opt, use_graph=not args.no_graph)

    n_batches = 8
    host = pinned_batches(n_batches, B, input_shape(args.model), num_classes(args.model), seed=rank)
    devb = [(x.to(dev), y.to(dev)) for x, y in host]
    _native.kernels().dmp_reset_launch_count()
    step(*devb[0])  # builds static buffers, warm-up iterations, graph capture
    torch.cuda.synchronize()
    launches_per_step = step.launches_per_iteration  # our kernels per iteration (what each graph replay re-issues)

    for i in range(args.warmup):
        step(*devb[i % n_batches])
    # ---- kernel-only: inputs already on the device
    barrier_sync(world)
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step(*devb[i % n_batches])
    e1.record()
    barrier_sync(world)
    ms = max_over_ranks(e0.elapsed_time(e1), world) / args.steps
    clocks = sampler.stop()
    # ---- end to end: pinned host batch -> device every step, loss/accuracy read back every step
    for i in range(2):
        step(*host[i % n_batches]); step.read_stats()
    barrier_sync(world)
    t0 = time.perf_counter()
    e0.record()
    for i in range(args.steps):
        step(*host[i % n_batches])
        stats = step.read_stats()
    e1.record()
    barrier_sync(world)
    ms_e2e = max_over_ranks(e0.elapsed_time(e1), world) / args.steps
    wall_e2e = (time.perf_counter() - t0) / args.steps * 1e3
    if hasattr(model, "check_health"):
        model.check_health()
    x0, y0 = host[0]
    gb = B * world
    res = {
        "metric": "images/sec (device-timed, max over ranks)", "impl": "ours",
        "value": gb / ms * 1e3, "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": (gb / ms * 1e3) / PUBLISHED[world] if world in PUBLISHED else None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"model": args.model, "global_batch": gb, "per_gpu_batch": B, "image": list(input_shape(args.model)),
                   "parallelism": "ddp%d" % world if world > 1 else "single", "optimizer": "sgd(m=0.9,wd=1e-4) fused",
                   "comm": args.comm if world > 1 else "none", "cuda_graph": step.graph is not None,
                   "l2": "per-step activation working set (>1 GB) exceeds the 126 MB L2; inputs rotate over %d batches" % n_batches},
        "clocks": clocks,
        "e2e": {"value": gb / ms_e2e * 1e3, "unit": "images/s", "ms_per_step": ms_e2e, "wall_ms_per_step": wall_e2e,
                "h2d_bytes_per_step": x0.numel() * x0.element_size() + y0.numel() * y0.element_size(),
                "d2h_bytes_per_step": 12},
        "gpu_launches": int(launches_per_step) * args.steps,
        "launches_per_step": int(launches_per_step),
        "op_paths": dict(OF.path_counts),
        "final_loss": stats[0],
    }
    if rank == 0:
        print(json.dumps(res))
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------------------
def run_reference(args):
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isdir(os.path.join(ref_dir, "model")):
        try:
            subprocess.run(["bash", os.path.join(ROOT, "baseline", "install_reference.sh")], check=True, capture_output=True,
                           timeout=300)
        except Exception as e:  # noqa: BLE001
            print(json.dumps({"impl": "reference", "unavailable": "reference install failed: %s" % str(e)[:120]}))
            return
    sys.path.insert(0, ref_dir)
    import torch
    import torch.nn as nn
    import torch.backends.cudnn as cudnn
    try:
        from model import MobileNetV2 as RefMobileNetV2  # the reference's own, unmodified, model file
    except Exception as e:  # noqa: BLE001
        print(json.dumps({"impl": "reference", "unavailable": "cannot import reference model: %s" % str(e)[:120]}))
        return
    if args.model != "mobilenetv2":
        print(json.dumps({"impl": "reference", "unavailable": "the reference only ships MobileNetV2 (no %s)" % args.model}))
        return

    world, rank, local = setup_dist(args)
    import torch.distributed as dist
    B = args.batch_per_gpu
    gb = B * world
    cudnn.benchmark = True  # data_parallel.py:78
    dtype_ctx = (lambda: torch.autocast("cuda", dtype=torch.bfloat16)) if args.ref_dtype == "bf16" else None

    def make_batches(n, batch, device, seed):
        g = torch.Generator().manual_seed(seed)
        out = []
        for _ in range(n):
            x = torch.randn(batch, 3, 32, 32, generator=g).pin_memory()
            y = torch.randint(0, 10, (batch,), generator=g).pin_memory()
            out.append((x, y))
        return out

    def train_loop(net, optimizer, criterion, batches, steps, device, e2e):
        """The body of the reference's train() (data_parallel.py:105-115)."""
        total = correct = 0
        for i in range(steps):
            inputs, targets = batches[i % len(batches)]
            inputs, targets = inputs.to(device), targets.to(device)
            optimizer.zero_grad()
            if dtype_ctx is not None:
                with dtype_ctx():
                    outputs = net(inputs)
                    loss = criterion(outputs.float(), targets)
            else:
                outputs = net(inputs)
                loss = criterion(outputs, targets)
            loss.backward()
            optimizer.step()
            if e2e:
                _ = loss.item()
                _, predicted = outputs.max(1)
                total += targets.size(0)
                correct += predicted.eq(targets).sum().item()
        return total

    def timed(net, optimizer, criterion, host_batches, device, everyone):
        dev_batches = [(x.to(device), y.to(device)) for x, y in host_batches]
        train_loop(net, optimizer, criterion, dev_batches, max(args.warmup, 3), device, False)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sampler = ClockSampler(local)
        sampler.start()
        e0.record()
        train_loop(net, optimizer, criterion, dev_batches, args.steps, device, False)
        e1.record()
        torch.cuda.synchronize()
        clocks = sampler.stop()
        ms = e0.elapsed_time(e1) / args.steps
        train_loop(net, optimizer, criterion, host_batches, 2, device, True)
        torch.cuda.synchronize()
        e0.record()
        train_loop(net, optimizer, criterion, host_batches, args.steps, device, True)
        e1.record()
        torch.cuda.synchronize()
        ms_e2e = e0.elapsed_time(e1) / args.steps
        return ms, ms_e2e, clocks

    torch.manual_seed(1234)
    dev0 = torch.device("cuda", local)
    result = {}
    # ---- (a) the reference's literal path: nn.DataParallel, single process (rank 0 drives all N GPUs).
    # The other ranks wait on a CPU (gloo) barrier so that nothing of theirs runs on the GPUs rank 0 is using.
    ctrl = dist.new_group(backend="gloo") if world > 1 else None
    barrier_sync(world)
    if rank == 0:
        host_batches = make_batches(8, gb, "cuda:0", 0)
        variants = [True] if world == 1 else [True, False]
        for bench_flag in variants:
            # stock: cudnn.benchmark = True (data_parallel.py:78).  With N > 1 torch's DataParallel starts fresh replica
            # threads every forward and cuDNN's autotune cache is per thread, so the stock setting re-tunes every
            # convolution every step; the run with the flag off is reported next to it and the better one is `value`.
            cudnn.benchmark = bench_flag
            net = RefMobileNetV2().to("cuda:0")
            net = torch.nn.DataParallel(net, device_ids=list(range(world)))
            criterion = nn.CrossEntropyLoss()
            optimizer = torch.optim.SGD(net.parameters(), lr=args.lr, momentum=0.9, weight_decay=1e-4)
            ms, ms_e2e, clocks = timed(net, optimizer, criterion, host_batches, torch.device("cuda:0"), False)
            result["dp_%s" % ("stock" if bench_flag else "nobench")] = (ms, ms_e2e, clocks)
            del net, optimizer
            torch.cuda.empty_cache()
        cudnn.benchmark = True
    if ctrl is not None:
        torch.cuda.synchronize()
        dist.barrier(group=ctrl)
    barrier_sync(world)
    # ---- (b) torch DistributedDataParallel + NCCL on the same reference model (stronger baseline)
    ddp_ms = ddp_ms_e2e = None
    if world > 1:
        net = RefMobileNetV2().to(dev0)
        net = torch.nn.parallel.DistributedDataParallel(net, device_ids=[local])
        criterion = nn.CrossEntropyLoss()
        optimizer = torch.optim.SGD(net.parameters(), lr=args.lr, momentum=0.9, weight_decay=1e-4)
        host_batches = make_batches(8, B, dev0, rank)
        barrier_sync(world)
        ms, ms_e2e, _ = timed(net, optimizer, criterion, host_batches, dev0, True)
        ddp_ms = max_over_ranks(ms, world)
        ddp_ms_e2e = max_over_ranks(ms_e2e, world)
    if rank == 0:
        best = min(result, key=lambda k: result[k][0])
        ms, ms_e2e, clocks = result[best]
        x_bytes = gb * 3 * 32 * 32 * 4 + gb * 8
        out = {
            "metric": "images/sec (device-timed, max over ranks)", "impl": "reference",
            "value": gb / ms * 1e3, "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": (gb / ms * 1e3) / PUBLISHED[world] if world in PUBLISHED else None,
            "dtype": args.ref_dtype, "data": "synthetic",
            "config": {"model": "mobilenetv2 (reference model/mobilenetv2.py, unmodified)", "global_batch": gb, "per_gpu_batch": B,
                       "image": [3, 32, 32], "parallelism": "torch.nn.DataParallel x%d (reference data_parallel.py path)" % world,
                       "optimizer": "torch.optim.SGD(m=0.9,wd=1e-4)", "cudnn_benchmark": best == "dp_stock",
                       "variant_used_for_value": best},
            "clocks": clocks,
            "e2e": {"value": gb / ms_e2e * 1e3, "unit": "images/s", "ms_per_step": ms_e2e, "h2d_bytes_per_step": x_bytes,
                    "d2h_bytes_per_step": 8},
            "gpu_launches": 0,
            "dataparallel_variants_images_per_s": {k: gb / v[0] * 1e3 for k, v in result.items()},
        }
        if ddp_ms is not None:
            out["torch_ddp_value"] = gb / ddp_ms * 1e3
            out["torch_ddp_ms_per_step"] = ddp_ms
            out["torch_ddp_e2e_value"] = gb / ddp_ms_e2e * 1e3
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
