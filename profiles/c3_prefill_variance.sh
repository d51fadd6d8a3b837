R=integration/_build/bin/run_vlm_cuda; F=/tmp/vlm_llava.mllm
$R --model llava --write $F 2>&1 | tail -1
for i in 1 2 3 4; do timeout 300 $R --model llava --device 4 --file $F --text 32 --decode 8 --limit 700 2>&1 | tail -1 | cut -c1-420; done
MLLM_CUDA_POOL_MB=0 timeout 300 $R --model llava --device 4 --file $F --text 32 --decode 8 --limit 700 2>&1 | tail -1 | cut -c1-420
