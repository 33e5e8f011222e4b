import torch, sys, os
sys.path.insert(0, "/root/repo")
from nerf_b200 import _lib
L=_lib.lib(); P=_lib.ptr
b,s=65536,63
out=torch.randn(b*s,54,device="cuda"); pts=torch.randn(b,s,3,device="cuda").cumsum(1).contiguous()
img=torch.empty(b,3,device="cuda"); seg=torch.empty(b,50,device="cuda")
for flag in (0,1):
    for i in range(3):
        e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
        e0.record(); L.nerf_head_composite(P(out),54,50,P(pts),None,0.0,b,s,P(img),P(seg),None,flag,_lib.stream()); e1.record(); torch.cuda.synchronize()
    print("flag",flag,"ms",e0.elapsed_time(e1), "GB/s", b*s*(54*4+12)/e0.elapsed_time(e1)/1e6)
