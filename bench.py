#!/usr/bin/env python
"""Headline benchmark: training throughput (images/s) on synthetic data of the named shape.

    python bench.py --gpus N --steps K --warmup W                   # this framework
    python bench.py --impl reference --gpus N --steps K --warmup W  # unmodified reference (baseline/reference_arm.py)

Default config: MobileNetV2 on CIFAR-10-shaped data (what the reference trains), per-GPU batch 128 (global 128*N: the
reference's published 256 @ 2 GPUs / 512 @ 4 GPUs, weak scaling), DistributedDataParallel.  Other configs of BASELINE.json:

    --model {mobilenetv2,resnet18,resnet50,vgg16}  --batch-per-gpu B  --sync-bn
    --parallelism ddp   multi-process DistributedDataParallel: peer-memory all-reduce fused with SGD, CUDA-graph step (default)
    --parallelism dp    single-process DataParallel (rank 0 drives all N GPUs, thread per replica) -- the comparison point
    --parallelism mp    layer-split pipeline over N stages with micro-batches (strong scaling: --global-batch 512)

For N > 1 launch with ``python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1
--master-port P bench.py --gpus N ...`` (one rank per GPU).  Rank 0 prints ONE JSON line.  With N > 1 the default (ddp)
line also carries ``dp_value`` (our DataParallel on the same N GPUs), ``allreduce_busGBs`` (largest gradient bucket,
device-timed fused all-reduce+SGD kernel, NCCL bus-bandwidth convention) and ``roofline_frac`` (that kernel against the
NVLink floor 2(W-1)/W * bytes / 900 GB/s).
"""
from __future__ import annotations

import argparse
import contextlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time
import types

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PUBLISHED = {2: 705.0, 4: 1293.0}  # nn.DataParallel img/s derived from /root/reference/Readme.md:283-292
PUBLISHED_MP = {2: 332.0, 4: 317.0}
NVLINK_GBS = 900.0                 # per direction per GPU (NVLink 5 through NVSwitch)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="mobilenetv2")
    ap.add_argument("--parallelism", default="ddp", choices=["ddp", "dp", "mp"])
    ap.add_argument("--batch-per-gpu", type=int, default=128)
    ap.add_argument("--global-batch", type=int, default=0, help="mp only (default 512, the reference's)")
    ap.add_argument("--microbatches", type=int, default=4)
    ap.add_argument("--schedule", default="1f1b", choices=["naive", "gpipe", "1f1b"])
    ap.add_argument("--reference-split", action="store_true", help="mp, MobileNetV2 on 4 GPUs: the reference's 3/6/6/2 split")
    ap.add_argument("--sync-bn", action="store_true")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip dp_value / allreduce_busGBs on the ddp line")
    ap.add_argument("--comm", default="peer", choices=["peer", "torch"])
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


TOOLS = types.SimpleNamespace(setup_dist=setup_dist, barrier_sync=barrier_sync, max_over_ranks=max_over_ranks,
                              ClockSampler=ClockSampler)


def input_shape(model_name):
    return (3, 224, 224) if model_name == "resnet50" else (3, 32, 32)


def num_classes(model_name):
    return 1000 if model_name == "resnet50" else 10


L2_NOTE = "per-step activation working set (>1 GB) exceeds the 126 MB L2; inputs rotate over %d batches"


def timed_steps(step, batches, n, world, local, read_stats=False, sample=True):
    """Device-timed loop (CUDA events, barrier + synchronize on both sides, max over ranks).  Returns ms/step, clocks."""
    import torch

    barrier_sync(world)
    sampler = ClockSampler(local) if sample else None
    if sampler:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    stats = None
    for i in range(n):
        step(*batches[i % len(batches)])
        if read_stats:
            stats = step.read_stats()
    e1.record()
    barrier_sync(world)
    wall = (time.perf_counter() - t0) / n * 1e3
    ms = max_over_ranks(e0.elapsed_time(e1), world) / n
    return ms, (sampler.stop() if sampler else None), wall, stats


# ---------------------------------------------------------------------------------------------------------
def measure_largest_bucket(model, world, iters=20):
    """Device-timed launches of the fused all-reduce(+SGD) kernel on the largest gradient bucket, all ranks in lock-step on
    the communication stream.  Returns (bytes, ms per launch max over ranks, bus GB/s, fraction of the NVLink floor)."""
    import torch

    red = getattr(model, "reducer", None)
    if red is None or world < 2 or not hasattr(red.backend, "stream"):
        return None
    bk = max(red.buckets, key=lambda b: b.nbytes())
    be = red.backend
    for _ in range(3):
        be.launch(bk)
    be.end_iteration()
    barrier_sync(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(be.stream):
        e0.record()
    for _ in range(iters):
        be.launch(bk)
    with torch.cuda.stream(be.stream):
        e1.record()
    be.end_iteration()
    barrier_sync(world)
    ms = max_over_ranks(e0.elapsed_time(e1), world) / iters
    nbytes = bk.nbytes()
    bus = nbytes / (ms * 1e-3) * 2 * (world - 1) / world / 1e9
    return {"bucket_bytes": nbytes, "n_buckets": len(red.buckets), "bucket_sizes": red.bucket_sizes(), "ms": ms,
            "busGBs": bus, "roofline_frac": bus / NVLINK_GBS}


def measure_dataparallel(args, world, steps, warmup):
    """Our single-process DataParallel over the same N GPUs (rank 0's process drives all of them), same model and batch."""
    import torch
    from distributed_model_parallel_b200.data import pinned_batches
    from distributed_model_parallel_b200.models import build_model
    from distributed_model_parallel_b200.optim import FlatFusedSGD
    from distributed_model_parallel_b200.parallel import DataParallel
    from distributed_model_parallel_b200.train.engine import TrainStep

    gb = args.batch_per_gpu * world
    torch.manual_seed(1234)
    dev = torch.device("cuda", 0)
    with torch.cuda.device(dev):
        net = build_model(args.model, num_classes=num_classes(args.model)).to(dev)
        dp = DataParallel(net, device_ids=list(range(world))) if world > 1 else net
        opt = FlatFusedSGD(net.parameters(), lr=args.lr, momentum=0.9, weight_decay=1e-4)
        step = TrainStep(dp, opt, use_graph=(world == 1 and not args.no_graph))
        host = pinned_batches(4, gb, input_shape(args.model), num_classes(args.model), seed=0)
        devb = [(x.to(dev), y.to(dev)) for x, y in host]
        for i in range(max(warmup, 3)):
            step(*devb[i % 4])
        ms, clocks, _, _ = timed_steps(step, devb, steps, 1, 0, sample=False)
        for i in range(2):
            step(*host[i % 4]); step.read_stats()
        ms_e2e, _, _, _ = timed_steps(step, host, steps, 1, 0, read_stats=True, sample=False)
        res = {"value": gb / ms * 1e3, "ms_per_step": ms, "e2e_value": gb / ms_e2e * 1e3, "launches_per_step": step.launches_per_iteration,
               "path": "thread per replica, persistent replicas" if getattr(dp, "_replica_set", None) is not None else "thread per replica"}
        if world > 1 and not args.no_graph:
            # the same DataParallel wrapper driven by one thread: every replica's forward + loss + backward is a CUDA graph
            from distributed_model_parallel_b200.parallel.dp_engine import DataParallelStep
            gstep = DataParallelStep(dp, opt)
            for i in range(max(warmup, 3)):
                gstep(*host[i % 4])
            gms, _, _, _ = timed_steps(gstep, host, steps, 1, 0, read_stats=True, sample=False)
            res["graph_per_replica_e2e_value"] = gb / gms * 1e3
            res["graph_per_replica_ms_per_step"] = gms
    return res


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
    B = args.batch_per_gpu
    gb = B * world
    n_batches = 8
    dp_only = args.parallelism == "dp"
    res = None
    if not dp_only:
        torch.manual_seed(1234)
        model = build_model(args.model, num_classes=num_classes(args.model)).to(dev)
        if args.sync_bn and world > 1:
            from distributed_model_parallel_b200.parallel import SyncBatchNorm
            model = SyncBatchNorm.convert_sync_batchnorm(model)
        sgd = dict(lr=args.lr, momentum=0.9, weight_decay=1e-4)
        if world > 1:
            model = DistributedDataParallel(model, device_ids=[local], comm_backend=args.comm,
                                            fused_sgd=sgd if args.comm == "peer" else None)
            opt = model.optimizer if args.comm == "peer" else FlatFusedSGD(model.parameters(), **sgd)
        else:
            opt = FlatFusedSGD(model.parameters(), **sgd)
        step = TrainStep(model, opt, use_graph=not args.no_graph)

        host = pinned_batches(n_batches, B, input_shape(args.model), num_classes(args.model), seed=rank)
        devb = [(x.to(dev), y.to(dev)) for x, y in host]
        _native.kernels().dmp_reset_launch_count()
        step(*devb[0])  # builds static buffers, warm-up iterations, graph capture
        torch.cuda.synchronize()
        launches_per_step = step.launches_per_iteration  # our kernels per iteration (what each graph replay re-issues)
        for i in range(args.warmup):
            step(*devb[i % n_batches])
        # ---- kernel-only: inputs already on the device
        ms, clocks, _, _ = timed_steps(step, devb, args.steps, world, local)
        # ---- end to end: pinned host batch -> device every step, loss/accuracy read back every step
        for i in range(2):
            step(*host[i % n_batches]); step.read_stats()
        ms_e2e, _, wall_e2e, stats = timed_steps(step, host, args.steps, world, local, read_stats=True, sample=False)
        if hasattr(model, "check_health"):
            model.check_health()
        x0, y0 = host[0]
        res = {
            "metric": "images/sec (device-timed, max over ranks)", "impl": "ours",
            "value": gb / ms * 1e3, "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": (gb / ms * 1e3) / PUBLISHED[world] if (world in PUBLISHED and args.model == "mobilenetv2") else None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"model": args.model, "global_batch": gb, "per_gpu_batch": B, "image": list(input_shape(args.model)),
                       "parallelism": "ddp%d" % world if world > 1 else "single", "optimizer": "sgd(m=0.9,wd=1e-4) fused",
                       "sync_bn": bool(args.sync_bn and world > 1),
                       "comm": args.comm if world > 1 else "none", "cuda_graph": step.graph is not None,
                       "l2": L2_NOTE % n_batches},
            "clocks": clocks,
            "e2e": {"value": gb / ms_e2e * 1e3, "unit": "images/s", "ms_per_step": ms_e2e, "wall_ms_per_step": wall_e2e,
                    "h2d_bytes_per_step": x0.numel() * x0.element_size() + y0.numel() * y0.element_size(),
                    "d2h_bytes_per_step": 12},
            "gpu_launches": int(launches_per_step) * args.steps,
            "launches_per_step": int(launches_per_step),
            "op_paths": dict(OF.path_counts),
            "final_loss": stats[0],
        }
        if world > 1 and not args.no_extras and args.comm == "peer":
            try:
                ar = measure_largest_bucket(model, world)
                if ar is not None:
                    res["allreduce_busGBs"] = ar["busGBs"]
                    res["roofline_frac"] = ar["roofline_frac"]
                    res["allreduce"] = ar
            except Exception as e:  # noqa: BLE001
                res["allreduce_error"] = repr(e)[:300]
    # ---- our DataParallel on the same GPUs (single process: rank 0 drives, the others wait on a CPU barrier)
    if dp_only or (world > 1 and not args.no_extras):
        import torch.distributed as dist
        ctrl = dist.new_group(backend="gloo") if world > 1 else None
        barrier_sync(world)
        dpres = None
        if rank == 0:
            try:
                dpres = measure_dataparallel(args, world, args.steps if dp_only else max(5, min(args.steps, 20)), args.warmup)
            except Exception as e:  # noqa: BLE001
                dpres = {"error": repr(e)[:400]}
        if ctrl is not None:
            torch.cuda.synchronize()
            dist.barrier(group=ctrl)
        if rank == 0:
            if dp_only:
                ok = "value" in dpres
                res = {"metric": "images/sec (device-timed, max over ranks)", "impl": "ours",
                       "value": dpres.get("value"), "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                       "ms_per_step": dpres.get("ms_per_step"), "higher_is_better": True, "scaling": "weak",
                       "vs_baseline": dpres["value"] / PUBLISHED[world] if ok and world in PUBLISHED and args.model == "mobilenetv2" else None,
                       "dtype": "bf16", "data": "synthetic",
                       "config": {"model": args.model, "global_batch": gb, "per_gpu_batch": B, "image": list(input_shape(args.model)),
                                  "parallelism": "dp%d (single process, thread per replica)" % world, "l2": L2_NOTE % 4},
                       "e2e": {"value": dpres.get("e2e_value"), "unit": "images/s",
                               "h2d_bytes_per_step": gb * (4 * 3 * input_shape(args.model)[1] * input_shape(args.model)[2] + 8),
                               "d2h_bytes_per_step": 12},
                       "gpu_launches": int(dpres.get("launches_per_step", 0)) * args.steps, "error": dpres.get("error"), "dp": dpres}
            else:
                # DataParallel has two drivers over the same wrapper: a thread per replica (torch's model) and one thread
                # replaying a captured graph per replica; the comparison point is the faster one
                cands = [(dpres.get("graph_per_replica_e2e_value"), "one thread, graph per replica (end to end)"),
                         (dpres.get("value"), "thread per replica")]
                cands = [c for c in cands if c[0]]
                res["dp_value"], res["dp_path"] = max(cands) if cands else (None, None)
                res["dp"] = dpres
    if rank == 0:
        print(json.dumps(res))
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


def bubble_report(kind, stages, microbatches):
    """Idle fraction of the schedule from the native event simulator (equal stages, backward = 2x forward): what the pipeline
    loses to fill and drain even with perfectly balanced stages and free communication -- the floor the measured step sits on."""
    from distributed_model_parallel_b200.parallel.pipeline import simulate_schedule

    try:
        makespan, peak = simulate_schedule(kind, stages, microbatches, 1.0, 2.0)
        busy = microbatches * 3.0
        return {"schedule": kind, "ideal_bubble_fraction": 1.0 - busy / makespan, "peak_in_flight_per_stage": peak}
    except Exception as e:  # noqa: BLE001
        return {"error": repr(e)[:200]}


# ---------------------------------------------------------------------------------------------------------
def run_ours_mp(args):
    """Layer-split pipeline (one stage per GPU) with micro-batches: strong scaling, global batch 512 like the reference's
    model_parallel.py:92.  Every step stage 0 copies the batch from pinned host memory and the loss is read back from the
    stage that owns it, so the device-timed number is already end to end."""
    import torch
    import torch.distributed as dist
    from distributed_model_parallel_b200 import _native
    from distributed_model_parallel_b200.data import pinned_batches
    from distributed_model_parallel_b200.models import build_model
    from distributed_model_parallel_b200.ops import functional as OF
    from distributed_model_parallel_b200.optim import FlatFusedSGD
    from distributed_model_parallel_b200.parallel.pipeline import PipelineStage, measured_boundaries, reference_split, split_sequential

    world, rank, local = setup_dist(args)
    dev = torch.device("cuda", local)
    gb = args.global_batch or 512
    torch.manual_seed(1234)
    full = build_model(args.model, num_classes=num_classes(args.model))
    if args.model == "mobilenetv2" and world == 4 and args.reference_split:
        stages = reference_split(full, stem_relu=True)
        split = "reference 3/6/6/2"
    else:
        seq = full.as_sequential()
        sample = torch.randn((max(1, gb // max(1, args.microbatches)),) + input_shape(args.model)) if rank == 0 else None
        bounds = measured_boundaries(seq, world, sample, dev, rank)
        stages = split_sequential(seq, world, boundaries=bounds)
        split = "balanced by measured layer time %s" % bounds
    module = stages[rank].to(dev)
    params = list(module.parameters())
    opt = FlatFusedSGD(params, lr=args.lr, momentum=0.9, weight_decay=1e-4) if params else None

    def loss_fn(logits, target):
        return OF.cross_entropy(logits if logits.dtype in (torch.float32, torch.bfloat16) else logits.float(), target)

    stage = PipelineStage(module, rank, world, optimizer=opt, loss_fn=loss_fn, schedule=args.schedule, loss_on="last", device=dev)
    host = pinned_batches(8, gb, input_shape(args.model), num_classes(args.model), seed=0) if rank == 0 else None
    loss_host = torch.zeros(1).pin_memory()

    def one(i):
        x, y = host[i % 8] if rank == 0 else (None, None)
        loss = stage.train_step(x, y, args.microbatches)
        if loss is not None:
            loss_host.copy_(loss.detach().reshape(1), non_blocking=True)
            torch.cuda.current_stream().synchronize()

    for i in range(max(args.warmup, 3)):
        one(i)
    c0 = _native.launch_count()
    barrier_sync(world)
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        one(i)
    e1.record()
    barrier_sync(world)
    ms = max_over_ranks(e0.elapsed_time(e1), world) / args.steps
    clocks = sampler.stop()
    launches = _native.launch_count() - c0
    lv = torch.tensor([float(loss_host[0]) if rank == world - 1 else 0.0], device=dev)
    if world > 1:
        dist.all_reduce(lv)
    if rank == 0:
        shape = input_shape(args.model)
        print(json.dumps({
            "metric": "images/sec (device-timed, max over ranks)", "impl": "ours", "value": gb / ms * 1e3, "unit": "images/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": (gb / ms * 1e3) / PUBLISHED_MP[world] if world in PUBLISHED_MP and args.model == "mobilenetv2" else None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"model": args.model, "global_batch": gb, "image": list(shape), "split": split,
                       "parallelism": "mp%d layer-split pipeline, %s x%d micro-batches, NCCL p2p" % (world, args.schedule, args.microbatches),
                       "cuda_graph": bool(stage.use_graphs), "l2": L2_NOTE % 8},
            "clocks": clocks,
            "e2e": {"value": gb / ms * 1e3, "unit": "images/s", "ms_per_step": ms,
                    "h2d_bytes_per_step": gb * (4 * shape[0] * shape[1] * shape[2] + 8), "d2h_bytes_per_step": 4,
                    "note": "the timed loop copies each batch from pinned host memory and reads the loss back: it is the end-to-end number"},
            "gpu_launches": int(launches), "graph_replays_rank0": stage.stats["graph_replays"], "final_loss": float(lv[0]),
            "bubble": bubble_report(args.schedule if args.microbatches > 1 else "naive", world, max(1, args.microbatches)),
        }))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        sys.path.insert(0, os.path.join(ROOT, "baseline"))
        import reference_arm
        real_stdout = sys.stdout

        class _Tee:
            """The reference's own loops print progress lines (utils.py:69-70); anything that is not our JSON line goes to stderr."""
            json_line = False

            def write(self, s):
                if s.lstrip().startswith('{"'):
                    self.json_line = True
                (real_stdout if self.json_line else sys.stderr).write(s)
                if s.endswith("\n"):
                    self.json_line = False

            def flush(self):
                real_stdout.flush(); sys.stderr.flush()

        sys.stdout = _Tee()
        try:
            (reference_arm.run_mp if a.parallelism == "mp" else reference_arm.run_dp)(a, TOOLS)
        finally:
            sys.stdout = real_stdout
    elif a.parallelism == "mp":
        run_ours_mp(a)
    else:
        run_ours(a)
