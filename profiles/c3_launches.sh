#!/bin/bash
# Full-size LLaVA-1.5-7B Q4_K request on the in-tree CUDA backend (prefill 608 rows + decode): a timed run, then the ncu launch list
# (gpu__time_duration only) of prefill + 4 decode steps.
# usage: profiles/c3_launches.sh <out.csv>
OUT=${1:-gpurun_out/r02_launches_c3.csv}
R=integration/_build/bin/run_vlm_cuda
F=/tmp/vlm_llava.mllm
$R --model llava --write $F 2>&1 | tail -1
$R --model llava --device 4 --file $F --text 32 --decode 32 --limit 700 2>&1 | tail -1
MLLM_CUDA_Q4K_CHAIN_OFF=1 $R --model llava --device 4 --file $F --text 32 --decode 32 --limit 700 2>&1 | tail -1 | sed 's/^{/{"q4k_chain": "off", /'
ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file $OUT $R --model llava --device 4 --file $F --text 32 --decode 4 --limit 700 > gpurun_out/c3_ncu.log 2>&1
tail -1 gpurun_out/c3_ncu.log
wc -l $OUT
rm -f $F
