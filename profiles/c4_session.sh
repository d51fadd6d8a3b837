timeout 600 python -m pytest tests/test_vision_kq_gpu.py -q -x 2>&1 | tail -4
PYTHONPATH=. timeout 300 python profiles/bench_vit_attn.py 2>&1 | tail -4 | tee gpurun_out/r02_vit_attn_micro.jsonl
timeout 600 python -m pytest tests/test_integration_vlm_gpu.py -q -x 2>&1 | tail -3
R=integration/_build/bin/run_vlm_cuda; F=/tmp/vlm_qwen2vl.mllm
$R --model qwen2vl --write $F 2>&1 | tail -1
for i in 1 2; do timeout 300 $R --model qwen2vl --device 4 --file $F --text 32 --decode 8 --limit 1600 2>&1 | tail -1 | cut -c1-420; done
