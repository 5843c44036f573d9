#!/bin/bash
# Single-GPU evidence run: the whole GPU test suite, smoke(), the headline bench at batch 128 and 512 (ours + reference arm), the
# other single-GPU configurations, the in-graph kernel times and the batch-size sweep.  Everything lands in gpurun_out/.
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_1gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_1gpu.log)
tail -3 gpurun_out/pytest_1gpu.log
timeout 120 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log | cut -c1-300
timeout 200 python bench.py --steps 100 --warmup 10 > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; echo "bench rc=$?"
timeout 200 python bench.py --steps 30 --warmup 5 --batch-per-gpu 512 > gpurun_out/bench_1gpu_b512.json 2>> gpurun_out/bench_1gpu.err
[ -n "$SKIP_REF" ] || { timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref_1gpu.json 2> gpurun_out/bench_ref_1gpu.err; echo "ref rc=$?"; }
timeout 200 python bench.py --model resnet18 --steps 30 --warmup 5 > gpurun_out/bench_resnet18_1gpu.json 2>> gpurun_out/bench_1gpu.err
timeout 300 python bench.py --model resnet50 --batch-per-gpu 64 --steps 10 --warmup 3 > gpurun_out/bench_resnet50_1gpu.json 2>> gpurun_out/bench_1gpu.err
[ -n "$WITH_SWEEPS" ] && { timeout 300 python bench/chain_bench.py > gpurun_out/chain_b128_final.txt 2>&1; timeout 400 python bench/batch_size_sweep.py --samples 32768 --out gpurun_out/batch_size_sweep.json > gpurun_out/batch_size_sweep.txt 2>&1; echo "sweep rc=$?"; }
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/bench_*1gpu*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d.get("config", {}).get("model"), d.get("config", {}).get("per_gpu_batch"), round(d["value"], 1), round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), "launches", d.get("launches_per_step"), d.get("torch_ddp_bf16_value"))
    except Exception as e:
        print(f, "ERR", e)
PY
if [ -n "$WITH_SWEEPS" ]; then tail -14 gpurun_out/batch_size_sweep.txt; fi
