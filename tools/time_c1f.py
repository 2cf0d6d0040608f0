import torch, statistics, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from libskylark_b200 import lib, sketch
n,d,s=100000,500,int(os.environ.get("C1_S","2000"))
A=torch.randn(d,n,dtype=torch.float64,device="cuda").t()
flush=torch.empty(256<<20,dtype=torch.uint8,device="cuda")
S=sketch.CWT(n,s,lib.Context(123834))
SA=S.apply(A,None,0); torch.cuda.synchronize()
ts=[]
for _ in range(9):
    flush.zero_()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record(); S.apply(A,SA,0); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
ms=statistics.median(ts); print("C1 col-major:", round(ms,4), "ms", round(408e6/ms/1e6), "GB/s", flush=True)
