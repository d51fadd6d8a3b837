#!/bin/bash
# sweep of the persistent GEMV launch shape (profiles only; not a bench number)
for cfg in "8 2 96" "16 1 96" "16 1 144" "16 2 160" "32 1 160" "16 1 208" "32 1 208" "8 1 96"; do
  set -- $cfg
  echo "== warps=$1 rpw=$2 ring_kb=$3"
  MLLMCU_GEMV_WARPS=$1 MLLMCU_GEMV_RPW=$2 MLLMCU_GEMV_RING_KB=$3 python profiles/bench_gemv.py fused 2>&1 | grep "pdl=1"
done
