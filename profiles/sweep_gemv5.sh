#!/bin/bash
# fifth sweep: warp counts that change the number of dealing rounds (qkv 6144 pair-groups, gate|up 11008 groups over 148 x W warps)
for w in 16 18 19 20 21 22 24; do
  echo "== warps=$w"
  MLLMCU_GEMV_WARPS=$w python profiles/bench_gemv.py fused 2>&1 | grep "pdl=1"
done
