for cfg in "4 0" "2 0" "2 1" "4 1" "1 0"; do set -- $cfg; MLLMCU_ATTN_SPLIT_MULT=$1 MLLM_CUDA_NO_ROPE_FUSION=$2 python bench.py --steps 2 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('mult $1 nofuse $2', d['decode_step']['ms'], d['value'])"; done
