for cfg in "0 0 0 0 2" "48 48 200 0 2" "74 74 148 0 2" "96 96 0 0 2" "64 64 168 0 2"; do
  set -- $cfg
  r=$(DMP_WGRAD_MAX_CTAS=$1 DMP_DW_WGRAD_MAX_BLOCKS=$2 DMP_BN_COOP_MAX_BLOCKS=$3 DMP_GEMM_CHAIN_MAX_SMS=$4 DMP_WGRAD_STREAMS=$5 python bench.py --steps 20 --warmup 5 --batch-per-gpu 512 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('%.3f ms  %.0f img/s' % (d['ms_per_step'], d['value']))")
  echo "b512 wgrad_ctas=$1 dw_wgrad_blocks=$2 bn_coop_blocks=$3 chain_gemm_sms=$4 side_streams=$5 : $r"
done
