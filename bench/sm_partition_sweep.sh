#!/bin/bash
# Sweep of the SM split between the data-gradient chain and the side-stream weight-gradient kernels (full captured step).
#   bash bench/sm_partition_sweep.sh [batch] > gpurun_out/sm_partition_sweep.txt
# columns: wgrad GEMM CTAs, depthwise wgrad blocks, cooperative BatchNorm blocks, SMs the chain's GEMMs size for, side streams
B=${1:-128}
for cfg in "0 0 0 0 2" "48 48 200 0 2" "48 48 0 0 2" "48 48 200 100 2" "40 40 216 0 2" "56 56 184 0 2" "48 48 200 0 1" "48 48 200 0 3" "32 32 200 0 3" "48 32 200 0 2" "32 48 200 0 2" "48 48 148 100 2"; do
  set -- $cfg
  r=$(DMP_WGRAD_MAX_CTAS=$1 DMP_DW_WGRAD_MAX_BLOCKS=$2 DMP_BN_COOP_MAX_BLOCKS=$3 DMP_GEMM_CHAIN_MAX_SMS=$4 DMP_WGRAD_STREAMS=$5 python bench.py --steps 30 --warmup 5 --batch-per-gpu $B 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('%.3f ms  %.0f img/s' % (d['ms_per_step'], d['value']))")
  echo "wgrad_ctas=$1 dw_wgrad_blocks=$2 bn_coop_blocks=$3 chain_gemm_sms=$4 side_streams=$5 : $r"
done
