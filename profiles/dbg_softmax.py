import sys; sys.path.insert(0,'.')
import numpy as np, torch
from mllm_b200 import capi as cu
from oracle import pyoracle as po
P=po.port()
for (outer,S,n,causal,old) in [(8,16,16,True,0),(16,577,577,False,0)]:
    rng=np.random.default_rng(outer+S+n)
    x=(rng.standard_normal((outer,S,n))*4).astype(np.float32)
    got=cu.softmax(torch.from_numpy(x).cuda(),causal,old).cpu().numpy()
    want=P.softmax(x,causal,old)
    d=np.argwhere(got.view(np.uint32)!=want.view(np.uint32))
    print(len(d), d[:12].tolist())
    for o,s,j in d[:6]:
        print(o,s,j,got[o,s,j],want[o,s,j], got[o,s,j].view(np.uint32)-want[o,s,j].view(np.uint32))
