import sys, os
sys.path.insert(0, os.getcwd())
import numpy as np, torch
from mlcomp_b200 import ops
dev='cuda:0'
def timeit(fn, iters=10, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    ev=[(torch.cuda.Event(True), torch.cuda.Event(True)) for _ in range(iters)]
    for a,b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return float(np.median([a.elapsed_time(b) for a,b in ev]))
for n,c,dt in [(42000,10,torch.float32),(4_000_000,10,torch.float32),(1_000_000,100,torch.float32),(1_000_000,1000,torch.float32),(1_000_000,1000,torch.float16),(4_000_000,64,torch.float16),(2_000_000,65,torch.float32)]:
    x=torch.randn((n,c),device=dev,dtype=dt); y=torch.randint(0,c,(n,),device=dev)
    t=timeit(lambda: ops.confusion(x,y))
    gb=(x.numel()*x.element_size()+n*8)/1e9
    print(n,c,dt, round(t,3),'ms', round(gb/t*1e3,1),'GB/s', round(gb/t*1e3/6550.4,3))
    del x,y
