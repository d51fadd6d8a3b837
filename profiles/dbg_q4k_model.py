import sys; sys.path.insert(0,'.')
import numpy as np
from mllm_b200 import mllm_file as mf, backend as be
from oracle import pyoracle as po
be.init_backend(0)
P=po.port()
name,wtype,B,S="toy-llama",mf.Q4_K,1,8
cfg=mf.config(name,cache_limit=48); cfg["wtype"]=wtype
t=mf.random_tensors(cfg,seed=S+wtype)
path="/tmp/x.mllm"; mf.write_model(path,cfg,t)
rng=np.random.default_rng(B*100+S)
ids=rng.integers(0,cfg["vocab"],(B,S)).astype(np.float32)
cpu=po.ModelRunner(P,cfg,t,mf.Names(cfg["rope_type"]),max_batch=B)
gpu=be.LlamaModel(cfg).load(path)
lc,dc=cpu.forward(ids,dump=True); lg,dg=gpu.forward(ids,dump=True)
cos=lambda a,b: float((a.astype(np.float64)*b).sum()/(np.linalg.norm(a.astype(np.float64))*np.linalg.norm(b.astype(np.float64))))
for L in range(cfg["layers"]+1):
    print(L, np.abs(dg[L]-dc[L]).max(), np.abs(dc[L]).max(), cos(dg[L],dc[L]))
print("logits", cos(lg,lc), np.abs(lg-lc).max(), np.abs(lc).max(), lg.argmax(1), lc.argmax(1))
for step in range(6):
    nxt=lc.argmax(1).astype(np.float32)[:,None]
    lc,lg=cpu.forward(nxt),gpu.forward(nxt)
    print(step, cos(lg,lc), lg.argmax(1), lc.argmax(1), np.sort(lc[0])[-3:])
