#!/bin/bash
# One multi-GPU box session: tensor-parallel 7B request (peer-memory exchange and NCCL) and the C5 batched decode, data-parallel.
# usage: profiles/multi_gpu_session.sh N out.jsonl
N=$1; OUT=${2:-gpurun_out/r02_multi_gpu_n$1.jsonl}
: > $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29660"
$TR bench.py --gpus $N --config tp --steps 3 --warmup 1 --no-cpu 2>&1 | grep '^{' | tail -1 >> $OUT
MLLM_TP_NCCL=1 $TR bench.py --gpus $N --config tp --steps 3 --warmup 1 --no-cpu 2>&1 | grep '^{' | tail -1 >> $OUT
$TR bench.py --gpus $N --config c5 --batch 64 --steps 2 --warmup 1 --no-cpu 2>&1 | grep '^{' | tail -1 >> $OUT
$TR bench.py --gpus $N --config c5 --batch 8 --steps 2 --warmup 1 --no-cpu 2>&1 | grep '^{' | tail -1 >> $OUT
$TR bench.py --gpus $N --steps 2 --warmup 1 --no-cpu 2>&1 | grep '^{' | tail -1 >> $OUT
cut -c1-420 $OUT
