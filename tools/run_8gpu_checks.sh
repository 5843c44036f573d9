#!/bin/bash
# Everything that needs the 8-GPU node, in one call: all-GPU DDP tests, the headline bench (ours + reference arm), the
# all-reduce sweep against NCCL, ResNet-50 + SyncBatchNorm and ResNet-18 DP-vs-DDP.  Logs go to gpurun_out/.
mkdir -p gpurun_out
N=${1:-8}
P=29700
tr() { P=$((P+1)); timeout ${TMO:-300} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P "$@"; }
(timeout 400 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -k "all_gpus or allreduce or nvls" 2>&1 | grep -v "^\[W\|^\[rank" | tail -15) > gpurun_out/pytest_multi_${N}gpu.log 2>&1
tail -3 gpurun_out/pytest_multi_${N}gpu.log
tr bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "ours rc=$?"
[ -n "$SKIP_REF" ] || TMO=500 tr bench.py --impl reference --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_ref_${N}gpu.json 2> gpurun_out/bench_ref_${N}gpu.err; echo "ref rc=$?"
tr bench/allreduce_sweep.py --max-mb 256 --out gpurun_out/allreduce_sweep_${N}gpu.json > gpurun_out/allreduce_sweep_${N}gpu.log 2>&1; echo "sweep rc=$?"
tr bench.py --gpus $N --model resnet50 --sync-bn --batch-per-gpu 64 --steps 10 --warmup 3 --no-extras > gpurun_out/bench_resnet50_syncbn_${N}gpu.json 2> gpurun_out/bench_resnet50_${N}gpu.err; echo "r50 rc=$?"
[ -n "$SKIP_R18" ] || tr bench.py --gpus $N --model resnet18 --steps 20 --warmup 5 > gpurun_out/bench_resnet18_${N}gpu.json 2> gpurun_out/bench_resnet18_${N}gpu.err; echo "r18 rc=$?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/bench_*gpu.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, {k: (round(d[k], 1) if isinstance(d.get(k), float) else d.get(k)) for k in ("n_gpus", "value", "ms_per_step", "dp_value", "allreduce_busGBs", "roofline_frac", "torch_ddp_fp32_value", "torch_ddp_bf16_value", "unavailable")}, "e2e", d.get("e2e", {}).get("value"))
    except Exception as e:
        print(f, "ERR", e)
PY
tail -8 gpurun_out/allreduce_sweep_${N}gpu.log | cut -c1-400
