#!/bin/bash
# Pipeline (model-parallel) measurements on 4 GPUs of one node: our schedules against no micro-batching and against the
# reference's own 4-rank loop.  Writes gpurun_out/pipeline_4gpu.log (one JSON line per run).
mkdir -p gpurun_out
OUT=gpurun_out/pipeline_4gpu.log
: > $OUT
P=29600
run() { P=$((P+1)); timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port $P "$@" > gpurun_out/pp_out.log 2> gpurun_out/pp_err.log; if grep -q '^{' gpurun_out/pp_out.log; then grep '^{' gpurun_out/pp_out.log | tail -1 >> $OUT; else echo "FAILED: $*" >> $OUT; grep -v "^\[W\|OMP_NUM\|^\*\*\*" gpurun_out/pp_err.log | tail -6 >> $OUT; fi; }
# the reference's configuration: MobileNetV2, CIFAR shape, global batch 512, 4 stages
run bench.py --gpus 4 --parallelism mp --schedule naive --microbatches 1 --steps 10 --warmup 3
run bench.py --gpus 4 --parallelism mp --schedule 1f1b --microbatches 4 --steps 10 --warmup 3
run bench.py --gpus 4 --parallelism mp --schedule 1f1b --microbatches 8 --steps 10 --warmup 3
run bench.py --gpus 4 --parallelism mp --schedule 1f1b --microbatches 4 --reference-split --steps 10 --warmup 3
run bench.py --gpus 4 --parallelism mp --impl reference --steps 10 --warmup 3
# bigger per-stage work: VGG-16 at 224x224 (global batch 128) and BERT-base at sequence length 512 (global batch 64)
for cfg in "naive 1" "1f1b 4" "1f1b 8" "gpipe 8"; do set -- $cfg
  run bench/model_bench.py --mode pipeline --model vgg16 --image 224 --batch-per-gpu 32 --schedule $1 --microbatches $2 --steps 8 --warmup 3
done
for cfg in "naive 1" "1f1b 4" "1f1b 8"; do set -- $cfg
  run bench/model_bench.py --mode pipeline --model bert_base --seq-len 512 --batch-per-gpu 16 --schedule $1 --microbatches $2 --steps 8 --warmup 3
done
cat $OUT | cut -c1-600
