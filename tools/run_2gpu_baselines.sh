#!/bin/bash
# Stock-torch baselines of BASELINE.md section 2 on 2 GPUs (same box, same synthetic shapes): torch.nn.DataParallel and torch DDP
# (bf16 autocast, channels_last, NCCL) on ResNet-18, torch DDP + nn.SyncBatchNorm on ResNet-50.  One JSON line each.
mkdir -p gpurun_out
OUT=gpurun_out/torch_baselines_2gpu.log
: > $OUT
timeout 200 python bench/model_bench.py --mode dp --impl torch --model resnet18 --gpus 2 --steps 20 --warmup 5 2>/dev/null | grep '^{' | tail -1 >> $OUT
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29801 bench/model_bench.py --mode ddp --impl torch --model resnet18 --steps 20 --warmup 5 2>/dev/null | grep '^{' | tail -1 >> $OUT
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29802 bench/model_bench.py --mode ddp --impl torch --model resnet50 --sync-bn --batch-per-gpu 64 --steps 10 --warmup 3 2>/dev/null | grep '^{' | tail -1 >> $OUT
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29803 bench/model_bench.py --mode ddp --impl torch --model mobilenetv2 --steps 20 --warmup 5 2>/dev/null | grep '^{' | tail -1 >> $OUT
cat $OUT
