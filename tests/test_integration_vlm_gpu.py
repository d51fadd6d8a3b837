"""BASELINE configs[2] / configs[3] graphs: the reference's UNCHANGED models/llava/modeling_llava.hpp (CLIP tower + projector + LLaMA
body, Q4_K Linears) and models/qwen2_vl/modeling_qwen2_vl.hpp (Conv3D patch embed, 2-D rotary ViT, patch merger, M-RoPE LLM, Q4_0
Linears) run by integration/_build/bin/run_vlm_cuda on the reference's CPU backend (--device 0) and on the in-tree CUDA backend
(--device 4), the way examples/demo_llava.cpp / demo_qwen2_vl.cpp drive them (prefill with the image, greedy steps with the image
tensors emptied).  Small dimensions, same graph: every op type the full-size models use is on the path (Convolution2D/3D, Parameter,
transpose, flatten, cat, range, F32 Embedding, LayerNorm, QuickGELU, GELU, fp32-KV flash attention, split, VisionRoPE,
apply_rotary_pos_emb_vision, where, index_put, MultimodalRoPE, clip, Q4_K / Q4_0 Linear with bias).

Two kinds of cases, compared on the last-position logits of every step and on the greedy tokens:
  * "-f32": the same graphs with F32 Linear weights.  Nothing is re-quantised between ops, so the two backends must agree to fp32
    summation-order noise plus the two fp16 roundings both backends share (the GELU / QuickGELU lookup tables indexed by the fp16 of
    the input, the fp16 KV cache) through the WHOLE graph: max-abs <= 1e-3 * max|logit| and cosine >= 1 - 1e-6 over 2 + 2 (LLaVA) and
    32 + 2 (Qwen2-VL) blocks (measured: 1.1e-4 / 3.1e-4, 1 - 4e-9 / 1 - 4e-8).  This is the check of every op and of the graph wiring (a wrong transpose, cat offset, splice position, rotary table or mask cannot hide under it).
  * quantised (Q4_K LLaVA, Q4_0 Qwen2-VL, the shipped formats): both backends re-quantise every Linear input to Q8_K / Q8_0, so a
    1e-7 difference in one activation can move one int8 step, and each later Linear multiplies the number of moved steps: op by op
    (profiles/r02_vlm_ops_llava.txt, profiles/r02_vlm_ops_qwen2vl.txt: every op output of both backends side by side) the two
    agree to <= 1e-6 up to the first such step and drift to 1e-2 over the depth of the towers.  Bars: max-abs <= 1e-1 * max|logit|,
    cosine >= 0.995, and the same token wherever the CPU backend's top-2 margin exceeds twice the measured difference."""
import json
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "integration", "_build", "bin", "run_vlm_cuda")

# llava: vocab,hidden,heads,ffn,layers,vhidden,vheads,vffn,vlayers,patch,img  (vocab > 32000: the graph looks for the image token 32000)
# qwen2vl: vocab,hidden,heads,kvheads,ffn,layers,vembed,grid_h,grid_w,mrope sections  (vocab > 151655; the tower always has 32 blocks of 16 heads)
CASES = {
    "llava": ("32064,256,4,512,2,256,4,256,2,14,56", 32064, "default"),
    "llava-d128": ("32064,512,4,512,2,512,8,512,3,14,84", 32064, "default"),
    "llava-f32": ("32064,256,4,512,2,256,4,256,2,14,56", 32064, "f32"),
    "qwen2vl": ("151936,256,4,2,512,2,512,4,6,8,12,12", 151936, "default"),
    "qwen2vl-d128": ("151936,512,4,2,512,2,512,8,8,16,24,24", 151936, "default"),
    "qwen2vl-f32": ("151936,256,4,2,512,2,512,4,6,8,12,12", 151936, "f32"),
}


def run(model, dims, *extra):
    cmd = [BIN, "--model", model, "--dims", dims, "--thread", str(max(1, min(8, len(os.sched_getaffinity(0)) // 2)))] + [str(e) for e in extra]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, (r.stderr[-3000:], r.stdout[-500:])
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert lines, (r.stderr[-3000:], r.stdout[-1000:])  # the reference's load pass reports errors and exits 0 (Module.hpp:128-137)
    return json.loads(lines[-1])


@pytest.mark.parametrize("case", sorted(CASES))
def test_unchanged_vlm_graph_cpu_and_cuda_backends_agree(tmp_path, case):
    assert os.path.exists(BIN), "integration/_build/bin/run_vlm_cuda missing: run `make -C integration` where /root/reference exists"
    dims, vocab, linear = CASES[case]
    model = case.split("-")[0]
    path = str(tmp_path / "m.mllm")
    assert run(model, dims, "--linear", linear, "--write", path)["tensors"] > 50
    n_dec = 8
    cpu = run(model, dims, "--device", 0, "--file", path, "--text", 10, "--decode", n_dec, "--limit", 256, "--dump", tmp_path / "cpu.bin")
    gpu = run(model, dims, "--device", 4, "--file", path, "--text", 10, "--decode", n_dec, "--limit", 256, "--dump", tmp_path / "gpu.bin")
    assert cpu["prompt_tokens"] == gpu["prompt_tokens"] and cpu["prompt_tokens"] > 15
    if linear != "f32":  # the decode plan recognises both LLM bodies (LLaVA: Q4_K op-level kernels + the trailing clip; Qwen2-VL: M-RoPE as a
        assert gpu["fused_steps"] == n_dec - 2, gpu  # shifted rotate-half table): step 0 is the prefill, step 1 is recorded, the rest replay the graph
    lc = np.fromfile(tmp_path / "cpu.bin", dtype=np.float32).reshape(n_dec, vocab)
    lg = np.fromfile(tmp_path / "gpu.bin", dtype=np.float32).reshape(n_dec, vocab)
    assert np.isfinite(lg).all()
    for k in range(n_dec):
        a, b = lc[k].astype(np.float64), lg[k].astype(np.float64)
        err = np.abs(a - b).max()
        cos = (a @ b) / np.sqrt((a @ a) * (b @ b))
        if linear == "f32":
            assert err <= 1e-3 * np.abs(a).max() and cos >= 1 - 1e-6, (case, k, err / np.abs(a).max(), cos)
            assert gpu["tokens"][k] == cpu["tokens"][k] or np.sort(a)[-1] - np.sort(a)[-2] <= 2 * err, (case, k)
            if gpu["tokens"][k] != cpu["tokens"][k]:
                break
            continue
        assert err <= 1e-1 * np.abs(a).max() and cos >= 0.995, (case, k, err / np.abs(a).max(), cos)
        top2 = np.sort(a)[-2:]
        if top2[1] - top2[0] > 2 * err:
            assert gpu["tokens"][k] == cpu["tokens"][k], (case, k)
        if gpu["tokens"][k] != cpu["tokens"][k]:
            break  # from here on the two backends decode different sequences
    os.remove(path)
    assert gpu["tokens"][0] == cpu["tokens"][0] or (np.sort(lc[0])[-1] - np.sort(lc[0])[-2]) <= 2 * np.abs(lc[0] - lg[0]).max()
