mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -s 2>&1 | grep -v "^\[W\|^\[rank" | tail -40) > gpurun_out/pytest_multi_2gpu.log 2>&1
tail -4 gpurun_out/pytest_multi_2gpu.log
P=29511
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $P bench.py --gpus 2 --steps 30 --warmup 5 > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; echo "ours rc=$?"
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((P+1)) bench.py --impl reference --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_ref_2gpu.json 2> gpurun_out/bench_ref_2gpu.err; echo "ref rc=$?"
python - <<'P'
import json
for f in ("gpurun_out/bench_2gpu.json","gpurun_out/bench_ref_2gpu.json"):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, {k:d.get(k) for k in ("value","ms_per_step","dp_value","allreduce_busGBs","roofline_frac","torch_ddp_fp32_value","torch_ddp_bf16_value","unavailable")}, d.get("e2e",{}).get("value"))
    except Exception as e:
        print(f, "ERR", e)
P
tail -3 gpurun_out/bench_2gpu.err
