"""Secondary benchmarks (BASELINE.json configs 2-4), all device-timed with the max over ranks:

  --mode dp        single-process DataParallel (thread per replica) over --gpus devices      [ours vs torch.nn.DataParallel]
  --mode ddp       multi-process DistributedDataParallel (launch with torchrun)              [ours vs torch DDP + NCCL]
  --mode pipeline  layer-split model parallel with micro-batches (launch with torchrun)      [naive / gpipe / 1f1b]

  --model resnet18 | resnet50 | vgg16 | bert_base | mobilenetv2      --sync-bn (ddp)      --impl ours | torch
"""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import torch.nn.functional as F

ap = argparse.ArgumentParser()
ap.add_argument("--mode", default="ddp", choices=["dp", "ddp", "pipeline"])
ap.add_argument("--impl", default="ours", choices=["ours", "torch"])
ap.add_argument("--model", default="resnet18")
ap.add_argument("--gpus", type=int, default=2)
ap.add_argument("--batch-per-gpu", type=int, default=128)
ap.add_argument("--steps", type=int, default=20)
ap.add_argument("--warmup", type=int, default=5)
ap.add_argument("--sync-bn", action="store_true")
ap.add_argument("--schedule", default="1f1b")
ap.add_argument("--microbatches", type=int, default=8)
ap.add_argument("--seq-len", type=int, default=128)
ap.add_argument("--image", type=int, default=0, help="square image side (default 32, 224 for resnet50)")
ap.add_argument("--partition", default="graph_time", choices=["graph_time", "params"])
args = ap.parse_args()

from distributed_model_parallel_b200.models import build_model
from distributed_model_parallel_b200.ops import functional as OF
from distributed_model_parallel_b200.optim import FlatFusedSGD


def shape():
    if args.image:
        return (3, args.image, args.image), (1000 if args.model == "resnet50" else 10)
    if args.model == "resnet50":
        return (3, 224, 224), 1000
    if args.model == "bert_base":
        return (args.seq_len,), 2
    return (3, 32, 32), 10


def batch(n, dev, seed=0):
    g = torch.Generator().manual_seed(seed)
    shp, nc = shape()
    x = torch.randint(0, 30000, (n,) + shp, generator=g) if args.model == "bert_base" else torch.randn((n,) + shp, generator=g)
    return x.to(dev), torch.randint(0, nc, (n,), generator=g).to(dev)


def timed_loop(step, steps, warmup, world=1):
    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    if world > 1:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms


sgd = dict(lr=0.05, momentum=0.9, weight_decay=1e-4)
nc = shape()[1]
if args.mode == "dp":
    devs = list(range(args.gpus))
    net = build_model(args.model, num_classes=nc).cuda(0)
    x, y = batch(args.batch_per_gpu * args.gpus, "cuda:0")
    if args.impl == "ours":
        from distributed_model_parallel_b200.parallel.data_parallel import DataParallel
        dp = DataParallel(net, device_ids=devs)
        opt = FlatFusedSGD(net.parameters(), **sgd)
        def step():
            loss = OF.cross_entropy(dp(x).float(), y); loss.backward(); opt.step()
    else:
        ctx = OF.use_reference_ops(); ctx.__enter__()
        dp = torch.nn.DataParallel(net, device_ids=devs)
        opt = torch.optim.SGD(net.parameters(), **sgd)
        def step():
            opt.zero_grad(); loss = F.cross_entropy(dp(x), y); loss.backward(); opt.step()
    ms = timed_loop(step, args.steps, args.warmup)
    print(json.dumps({"mode": "dp", "impl": args.impl, "model": args.model, "gpus": args.gpus, "global_batch": x.shape[0], "ms_per_step": ms,
                      "images_per_s": x.shape[0] / ms * 1e3}))
    sys.exit(0)

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
torch.manual_seed(0)
if args.mode == "ddp":
    net = build_model(args.model, num_classes=nc).to(dev)
    x, y = batch(args.batch_per_gpu, dev, seed=rank)
    if args.impl == "ours":
        from distributed_model_parallel_b200.parallel import DistributedDataParallel
        from distributed_model_parallel_b200.parallel.sync_batchnorm import convert_sync_batchnorm
        from distributed_model_parallel_b200.train.engine import TrainStep
        if args.sync_bn:
            net = convert_sync_batchnorm(net)
        ddp = DistributedDataParallel(net, device_ids=[local], fused_sgd=sgd)
        ts = TrainStep(ddp, ddp.optimizer, use_graph=True)
        step = lambda: ts(x, y)
    else:
        ctx = OF.use_reference_ops(); ctx.__enter__()
        if args.sync_bn:
            net = torch.nn.SyncBatchNorm.convert_sync_batchnorm(net)
        net = net.to(memory_format=torch.channels_last)
        ddp = torch.nn.parallel.DistributedDataParallel(net, device_ids=[local])
        opt = torch.optim.SGD(ddp.parameters(), **sgd)
        xc = x.contiguous(memory_format=torch.channels_last) if x.dim() == 4 else x
        def step():
            opt.zero_grad()
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out = ddp(xc)
            F.cross_entropy(out.float(), y).backward(); opt.step()
    ms = timed_loop(step, args.steps, args.warmup, world)
    if rank == 0:
        print(json.dumps({"mode": "ddp", "impl": args.impl, "model": args.model, "sync_bn": args.sync_bn, "gpus": world,
                          "global_batch": args.batch_per_gpu * world, "ms_per_step": ms, "images_per_s": args.batch_per_gpu * world / ms * 1e3,
                          "paths": dict(OF.path_counts) if args.impl == "ours" else None}))
else:
    from distributed_model_parallel_b200.parallel.pipeline import PipelineStage, layer_costs, measured_boundaries, split_sequential
    full = build_model(args.model, num_classes=nc)
    seq = full.as_sequential()
    gb = args.batch_per_gpu * world
    mbs = 1 if args.schedule == "naive" else args.microbatches
    if args.partition == "params":
        stages = split_sequential(seq, world, layer_costs(seq, "params"))
    else:
        sample = batch(max(1, gb // mbs), "cpu")[0] if rank == 0 else None
        stages = split_sequential(seq, world, boundaries=measured_boundaries(seq, world, sample, dev, rank))
    module = stages[rank].to(dev)
    params = list(module.parameters())
    opt = FlatFusedSGD(params, **sgd) if params else None
    st = PipelineStage(module, rank, world, optimizer=opt, loss_fn=lambda o, t: OF.cross_entropy(o.float() if o.dtype not in (torch.float32, torch.bfloat16) else o, t),
                       schedule=args.schedule, loss_on="last", device=dev)
    gb = args.batch_per_gpu * world
    x, y = batch(gb, dev)
    mbs = 1 if args.schedule == "naive" else args.microbatches
    step = lambda: st.train_step(x if rank == 0 else None, y if rank == 0 else None, mbs)
    ms = timed_loop(step, args.steps, args.warmup, world)
    if rank == 0:
        print(json.dumps({"mode": "pipeline", "model": args.model, "schedule": args.schedule, "microbatches": mbs, "stages": world, "global_batch": gb,
                          "ms_per_step": ms, "samples_per_s": gb / ms * 1e3, "stage_layers": [len(s) for s in stages]}))
dist.destroy_process_group()
