#!/bin/bash
# Full-size Qwen2-VL-2B Q4_0 request (1036 x 1036 image -> 1369 image tokens) on the in-tree CUDA backend: a timed run, then the ncu launch list.
# usage: profiles/c4_launches.sh <out.csv>
OUT=${1:-gpurun_out/r02_launches_c4.csv}
R=integration/_build/bin/run_vlm_cuda
F=/tmp/vlm_qwen2vl.mllm
$R --model qwen2vl --write $F 2>&1 | tail -1
$R --model qwen2vl --device 4 --file $F --text 32 --decode 32 --limit 1600 2>&1 | tail -1
$R --model qwen2vl --device 4 --file $F --text 32 --decode 32 --limit 1600 2>&1 | tail -1
ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file $OUT $R --model qwen2vl --device 4 --file $F --text 32 --decode 4 --limit 1600 > gpurun_out/c4_ncu.log 2>&1
tail -1 gpurun_out/c4_ncu.log | cut -c1-300
wc -l $OUT
rm -f $F
