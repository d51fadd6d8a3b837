#!/bin/bash
# LLaMA-2-7B Q4_0 through the unchanged modeling_llama.hpp on the in-tree CUDA backend: the same request four times in one process
F=/tmp/mllm_test_llama2-7b_q4_0.mllm
PYTHONPATH=. python - <<'PY'
from mllm_b200 import mllm_file as mf
cfg = mf.config("llama2-7b", cache_limit=1024)
mf.write_random_model_torch(cfg, "/tmp/mllm_test_llama2-7b_q4_0.mllm", device="cuda", seed=1234)
PY
integration/_build/bin/run_llama_cuda --device 4 --model llama2-7b --file $F --prompt 512 --decode 64 --limit 1024 --repeat 4 2>&1 | tail -1 | cut -c1-420
