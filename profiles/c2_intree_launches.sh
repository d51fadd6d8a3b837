#!/bin/bash
# LLaMA-2-7B Q4_0 (BASELINE configs[1]) through the UNCHANGED models/llama/modeling_llama.hpp on the in-tree CUDA backend: a timed run, then
# the ncu launch list of the prefill (512 tokens) + 3 decode steps.   usage: profiles/c2_intree_launches.sh <out.csv>
OUT=${1:-gpurun_out/r02_launches_c2_intree.csv}
F=/tmp/mllm_test_llama2-7b_q4_0.mllm
PYTHONPATH=. python - <<'PY'
from mllm_b200 import mllm_file as mf
cfg = mf.config("llama2-7b", cache_limit=1024)
mf.write_random_model_torch(cfg, "/tmp/mllm_test_llama2-7b_q4_0.mllm", device="cuda", seed=1234)
PY
R=integration/_build/bin/run_llama_cuda
$R --device 4 --model llama2-7b --file $F --prompt 512 --decode 16 --limit 1024 2>&1 | tail -1 | cut -c1-300
$R --device 4 --model llama2-7b --file $F --prompt 512 --decode 16 --limit 1024 2>&1 | tail -1 | cut -c1-300
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file $OUT $R --device 4 --model llama2-7b --file $F --prompt 512 --decode 3 --limit 1024 > gpurun_out/c2_intree_ncu.log 2>&1
wc -l $OUT
