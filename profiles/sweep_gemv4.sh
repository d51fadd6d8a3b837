#!/bin/bash
# fourth sweep: more warps per SM (the stream phase runs at ~5.2 TB/s-equivalent with 16 warps at IPC 1.4: issue/latency bound, below HBM)
for cfg in "32 1 200" "32 2 200" "16 1 200" "16 2 200"; do
  set -- $cfg
  echo "== warps=$1 rpw=$2 ring_kb=$3"
  MLLMCU_GEMV_WARPS=$1 MLLMCU_GEMV_RPW=$2 MLLMCU_GEMV_RING_KB=$3 python profiles/bench_gemv.py fused 2>&1 | grep "pdl=1"
done
