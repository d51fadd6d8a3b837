set -e
python -m pytest tests/test_decode_gpu.py -x -q 2>&1 | tail -3
for L in 0 16 32 48 64; do
  echo "== lead $L MB"; MLLM_CUDA_L2_LEAD_MB=$L timeout 300 python profiles/prof_decode.py 512 64 2>&1 | grep -o "decode ms [0-9.]*"
done
for A in 1.0 2.0 4.0; do
  echo "== lead 32 alpha $A"; MLLMCU_GEMV_PF_ALPHA=$A MLLM_CUDA_L2_LEAD_MB=32 timeout 300 python profiles/prof_decode.py 512 64 2>&1 | grep -o "decode ms [0-9.]*"
done
