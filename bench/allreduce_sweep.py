"""Gradient-bucket all-reduce sweep (BASELINE.json config 5): our peer-memory kernels (one-shot / two-shot, plain and
fused with SGD) vs torch.distributed.all_reduce on NCCL, 1 MB - 1 GB, device-timed, max over ranks.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/allreduce_sweep.py [--max-mb 1024]

Bus bandwidth uses the NCCL convention (S/t)*2(W-1)/W.  The roofline column is the fraction of the link-time floor of
the chosen algorithm at the measured 770 GB/s per direction per GPU (B200_PROFILING.md): one-shot moves (W-1)*S into
every GPU, two-shot 2*S*(W-1)/W (for the fused variant the second S is the fp32 master + bf16 shadow written back)."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from distributed_model_parallel_b200 import _native
from distributed_model_parallel_b200.parallel.symm import DTYPE_CODE, ProcessSymmGroup

LINK = 770e9
ap = argparse.ArgumentParser()
ap.add_argument("--max-mb", type=int, default=1024)
ap.add_argument("--iters", type=int, default=10)
ap.add_argument("--out", default="")
args = ap.parse_args()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
lib = _native.kernels()
g = ProcessSymmGroup()
state = g.new_state()
sp = torch.cuda.current_stream().cuda_stream
hp = torch.tensor([0.1, 0.9, 1e-4, 1.0], device="cuda")


def timed(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()) * 1e-3


try:  # NVLS multicast buffer (switch-side reduction); absent on platforms without multicast support
    from distributed_model_parallel_b200.parallel.symm import ProcessMulticast
    mcbuf = ProcessMulticast(args.max_mb * 1024 * 1024)
    mc_master = ProcessMulticast(args.max_mb * 1024 * 1024)      # fused variant: fp32 master and bf16 shadow live in multicast
    mc_shadow = ProcessMulticast(args.max_mb * 1024 * 1024 // 2)  # buffers too (multimem.st writes every rank's copy)
except Exception as e:  # noqa: BLE001
    mcbuf = mc_master = mc_shadow = None
    if rank == 0:
        print("# NVLS multicast unavailable: %s" % str(e)[:200], flush=True)

rows = []
mb = 1
sizes = []
while mb <= args.max_mb:
    sizes.append(mb); mb *= 4
if sizes[-1] != args.max_mb:
    sizes.append(args.max_mb)
for mb in sizes:
    n = mb * 1024 * 1024 // 4  # fp32 gradients
    n = n // (8 * world) * (8 * world)
    S = n * 4
    buf = g.alloc(S); grad = buf.view(torch.float32, n); grad.normal_()
    out = torch.empty_like(grad)
    pbuf = g.alloc(S); master = pbuf.view(torch.float32, n); master.normal_()
    sbuf = g.alloc(n * 2); shadow = sbuf.view(torch.bfloat16, n)
    mom = torch.zeros(n, device="cuda")
    ref = torch.empty(n, device="cuda").normal_()
    iters = max(3, min(args.iters, int(2e9 // S) + 3))
    res = {"MB": mb, "world": world}
    res["nccl_s"] = timed(lambda: dist.all_reduce(ref), iters)
    for algo, name in ((0, "oneshot"), (1, "twoshot")):
        if algo == 0 and S * (world - 1) > (1 << 30) * 2:
            continue
        res[name + "_s"] = timed(lambda: _native.check(lib.dmp_allreduce_fused(algo, 0, buf.ptrs, state.ptrs, rank, world, n, 1.0 / world,
                                 out.data_ptr() if algo == 0 else None, None, None, None, None, 0, None, None, 0, 0, sp)), iters)
        res[name + "_fused_sgd_s"] = timed(lambda: _native.check(lib.dmp_allreduce_fused(algo, 0, buf.ptrs, state.ptrs, rank, world, n, 1.0 / world,
                                           None, master.data_ptr(), mom.data_ptr(), shadow.data_ptr(), hp.data_ptr(), 0, pbuf.ptrs, sbuf.ptrs, 1, 0, sp)), iters)
    if mcbuf is not None:
        mcbuf.view(torch.float32, n).normal_()
        res["nvls_s"] = timed(lambda: _native.check(lib.dmp_allreduce_nvls(mcbuf.mc_ptr, state.ptrs, rank, world, n, 1.0 / world, 0, sp)), iters)
        if n % (8 * world) == 0:
            mc_master.view(torch.float32, n).normal_()
            res["nvls_fused_sgd_s"] = timed(lambda: _native.check(lib.dmp_allreduce_nvls_sgd(
                mcbuf.mc_ptr, mc_master.mc_ptr, mc_shadow.mc_ptr, mc_master.view(torch.float32, n).data_ptr(), mom.data_ptr(), hp.data_ptr(), 0,
                state.ptrs, rank, world, n, 1.0 / world, 0, sp)), iters)
    # NCCL + separate fused SGD kernel = what the fused variants replace
    res["nccl_plus_sgd_s"] = timed(lambda: (dist.all_reduce(ref), lib.dmp_sgd_flat(master.data_ptr(), shadow.data_ptr(), ref.data_ptr(), 0, mom.data_ptr(), n, hp.data_ptr(), 0, 1, sp)), iters)
    f = 2.0 * (world - 1) / world
    for k in [k for k in res if k.endswith("_s")]:
        res[k[:-2] + "_busGBs"] = S / res[k] * f / 1e9
    if "oneshot_s" in res:
        res["oneshot_roofline_frac"] = ((world - 1) * S / LINK) / res["oneshot_s"]
    res["twoshot_roofline_frac"] = (2.0 * S * (world - 1) / world / LINK) / res["twoshot_s"]
    res["twoshot_fused_roofline_frac"] = ((S + S + S / 2) * (world - 1) / world / LINK) / res["twoshot_fused_sgd_s"]
    rows.append(res)
    if rank == 0:
        print(json.dumps(res), flush=True)
    del buf, pbuf, sbuf, grad, master, shadow, out, mom, ref
    torch.cuda.empty_cache()
if rank == 0 and args.out:
    json.dump(rows, open(args.out, "w"), indent=1)
assert lib.dmp_comm_status(state.local.data_ptr(), sp) == 0
dist.destroy_process_group()
