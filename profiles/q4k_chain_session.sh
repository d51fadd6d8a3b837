timeout 600 python -m pytest tests/test_vision_kq_gpu.py -q -x -k "q4k or q8_K or k_quants" 2>&1 | tail -3
PYTHONPATH=. timeout 300 python profiles/bench_q4k_gemv.py 2>&1 | tail -10 | cut -c1-260 | tee gpurun_out/r02_q4k_gemv_micro.jsonl
timeout 600 python -m pytest tests/test_integration_vlm_gpu.py tests/test_integration_gpu.py -q -x 2>&1 | tail -3
R=integration/_build/bin/run_vlm_cuda; F=/tmp/vlm_llava.mllm
$R --model llava --write $F 2>&1 | tail -1
timeout 300 $R --model llava --device 4 --file $F --text 32 --decode 32 --limit 700 2>&1 | tail -1 | cut -c1-330
timeout 300 $R --model llava --device 4 --file $F --text 32 --decode 32 --limit 700 2>&1 | tail -1 | cut -c1-330
