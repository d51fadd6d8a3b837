#!/bin/bash
# third sweep: rings small enough that CONSECUTIVE launches co-reside on an SM (one CTA per SM per launch): the next launch fills its ring
# and sits in griddepcontrol.wait while the current one streams (profiles only; not a bench number)
for cfg in "0 0 200" "16 1 76" "16 1 100" "8 2 100" "8 1 100" "12 1 100"; do
  set -- $cfg
  echo "== warps=$1 rpw=$2 ring_kb=$3"
  MLLMCU_GEMV_WARPS=$1 MLLMCU_GEMV_RPW=$2 MLLMCU_GEMV_RING_KB=$3 python profiles/bench_gemv.py fused 2>&1 | grep "pdl=1"
done
