import sys; sys.path.insert(0,"sq-learn_b200"); sys.path.insert(0,"tests"); sys.path.insert(0,".")
import torch, time, numpy as np
from sqlearn import _cabi, _device as dv
import bench
from sqlearn.QuantumUtility import Utility as U
B=8_400_000
for n,p in [(2.5e7,0.3),(1000.0,0.3),(50.0,0.05),(5.0,0.3)]:
    nn=torch.full((B,),n,dtype=torch.float64,device="cuda"); a=torch.full((B,),p,dtype=torch.float64,device="cuda"); b=1-a
    out=torch.empty_like(nn)
    ms=bench._timed(torch, lambda: _cabi.check(_cabi.lib.sq_binomial_sample(nn.data_ptr(),a.data_ptr(),b.data_ptr(),B,1,0,out.data_ptr(),dv.stream_ptr()),"b"),3)
    print("binomial", n, p, ms, "ms for", B, "draws ->", B/ms/1e6, "Gdraws/s")
g=torch.Generator(device="cuda"); g.manual_seed(1)
for B,d,delta in [(4096,1000,0.1),(592,1000,0.1),(148,1000,0.1),(4096,128,0.1),(4096,4096,0.1)]:
    V=torch.rand((B,d),generator=g,device="cuda",dtype=torch.float64)*2-1
    ms=bench._timed(torch, lambda: U._tomography_device(V,delta,None,"L2",True,False,0),3)
    print("tomo single-shot",B,d,ms,"ms")
