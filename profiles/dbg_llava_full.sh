R=integration/_build_dbg/bin/run_vlm_cuda
$R --model llava --write /tmp/l.mllm | tail -1
mkdir -p /tmp/dcpu /tmp/dgpu
MLLM_DUMP_MAXROWS=4 MLLM_DUMP_OPS=/tmp/dcpu $R --model llava --device 0 --file /tmp/l.mllm --text 32 --decode 1 --limit 700 --thread 8 | tail -1
MLLM_DUMP_MAXROWS=4 MLLM_DUMP_OPS=/tmp/dgpu $R --model llava --device 4 --file /tmp/l.mllm --text 32 --decode 1 --limit 700 | tail -1
python profiles/compare_op_dumps.py /tmp/dcpu /tmp/dgpu > gpurun_out/vlm_llava_full_ops.txt 2>&1
wc -l gpurun_out/vlm_llava_full_ops.txt
