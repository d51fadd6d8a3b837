#!/bin/bash
# BASELINE configs[2] and configs[3] at FULL size through the unchanged reference graphs (integration/_build/bin/run_vlm_cuda):
#   LLaVA-1.5-7B Q4_K (CLIP ViT-L/14-336: 576 image tokens + text into the LLaMA-2-7B body), Qwen2-VL-2B Q4_0 (1036 x 1036 image =
#   74 x 74 patches -> 1369 image tokens, M-RoPE LLM), random-init weights written by the binary itself, synthetic pixels and ids.
# CPU backend (reference) and CUDA backend on the same file; the CPU run decodes CPU_DEC steps, the CUDA run GPU_DEC steps.
# usage: profiles/vlm_full.sh <out.jsonl> [llava|qwen2vl|both]
OUT=${1:-gpurun_out/r02_vlm_full.jsonl}; WHICH=${2:-both}
R=integration/_build/bin/run_vlm_cuda
T=$(python - <<'PY'
import os
print(max(1, len(os.sched_getaffinity(0)) // 2))
PY
)
CPU_DEC=${CPU_DEC:-4}; GPU_DEC=${GPU_DEC:-32}
: > $OUT
run_case() { # model limit text
  local m=$1 lim=$2 txt=$3 f=/tmp/vlm_$1.mllm
  local t0=$(date +%s)
  $R --model $m --write $f 2>&1 | tail -1 >> $OUT
  echo "{\"case\": \"$m\", \"write_s\": $(( $(date +%s) - t0 ))}" >> $OUT
  $R --model $m --device 4 --file $f --text $txt --decode $GPU_DEC --limit $lim --dump /tmp/vlm_$m.gpu.bin 2>&1 | tail -3 | grep -E '^\{|error|Error' >> $OUT
  $R --model $m --device 0 --file $f --text $txt --decode $CPU_DEC --limit $lim --thread $T --dump /tmp/vlm_$m.cpu.bin 2>&1 | grep '^{' | sed "s/^{/{\"threads\": $T, /" >> $OUT
  # noise floor: the SAME reference backend with every input pixel moved by one ulp
  $R --model $m --device 0 --file $f --text $txt --decode $CPU_DEC --limit $lim --thread $T --perturb 1 --dump /tmp/vlm_$m.cpu1.bin 2>&1 | grep '^{' | sed "s/^{/{\"perturbed_pixels_1ulp\": 1, /" >> $OUT
  python - $m $CPU_DEC $GPU_DEC >> $OUT <<'PY'
import sys, json
import numpy as np
m, nc, ng = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
a = np.fromfile(f"/tmp/vlm_{m}.cpu.bin", dtype=np.float32).reshape(nc, -1).astype(np.float64)
b = np.fromfile(f"/tmp/vlm_{m}.gpu.bin", dtype=np.float32).reshape(ng, -1).astype(np.float64)[:nc]
rows = []
for k in range(nc):
    err = np.abs(a[k] - b[k]).max() / np.abs(a[k]).max()
    cos = (a[k] @ b[k]) / np.sqrt((a[k] @ a[k]) * (b[k] @ b[k]))
    top = np.sort(a[k])[-2:]
    rows.append({"step": k, "rel_max_err": float(err), "cos": float(cos), "cpu_tok": int(a[k].argmax()), "gpu_tok": int(b[k].argmax()), "cpu_margin_over_max": float((top[1] - top[0]) / np.abs(a[k]).max())})
    if rows[-1]["cpu_tok"] != rows[-1]["gpu_tok"]:
        break
print(json.dumps({"case": m, "logits_cpu_vs_cuda": rows}))
c = np.fromfile(f"/tmp/vlm_{m}.cpu1.bin", dtype=np.float32).reshape(nc, -1).astype(np.float64)
rows = []
for k in range(nc):
    err = np.abs(a[k] - c[k]).max() / np.abs(a[k]).max()
    cos = (a[k] @ c[k]) / np.sqrt((a[k] @ a[k]) * (c[k] @ c[k]))
    rows.append({"step": k, "rel_max_err": float(err), "cos": float(cos), "cpu_tok": int(a[k].argmax()), "cpu_1ulp_tok": int(c[k].argmax())})
    if rows[-1]["cpu_tok"] != rows[-1]["cpu_1ulp_tok"]:
        break
print(json.dumps({"case": m, "floor_cpu_vs_cpu_with_pixels_moved_1ulp": rows}))
PY
  rm -f $f /tmp/vlm_$m.*.bin
}
[ $WHICH = both -o $WHICH = llava ] && run_case llava 700 32
[ $WHICH = both -o $WHICH = qwen2vl ] && run_case qwen2vl 1600 32
cat $OUT
