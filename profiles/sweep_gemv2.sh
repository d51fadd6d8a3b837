#!/bin/bash
# second sweep: two CTAs per SM (profiles only; not a bench number)
for cfg in "16 1 76 2" "8 2 76 2" "8 1 40 4" "16 2 200 1"; do
  set -- $cfg
  echo "== warps=$1 rpw=$2 ring_kb=$3 ctas_per_sm=$4"
  MLLMCU_GEMV_WARPS=$1 MLLMCU_GEMV_RPW=$2 MLLMCU_GEMV_RING_KB=$3 MLLMCU_GEMV_CPS=$4 python profiles/bench_gemv.py fused 2>&1 | grep "pdl=1"
done
