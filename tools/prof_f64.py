import sys,os
sys.path.insert(0,"tests"); sys.path.insert(0,".")
import numpy as np, torch, szref, sz_b200, bench
sz_b200.set_device(0); sz_b200.init(dict(errorBoundMode="REL",relBoundRatio=1e-3))
a=bench.make_field((400,400,400)).astype(np.float64); x=torch.from_numpy(a).cuda(); buf=torch.empty(sz_b200.compress_bound(np.float64,a.shape),dtype=torch.uint8,device="cuda"); out=torch.empty_like(x)
for _ in range(3):
    n=sz_b200.compress_device(x,buf,sz_b200.REL,1e-4,1e-3); st=sz_b200.last_stats(); c=(round(st.ms_total,3), {k: round(v,3) for k,v in st.as_dict()["ms_stage"].items() if v>0.004})
    sz_b200.decompress_device(buf,n,out); st=sz_b200.last_stats(); d=(round(st.ms_total,3), {k: round(v,3) for k,v in st.as_dict()["ms_stage"].items() if v>0.004})
print(a.shape, a.dtype, n, a.nbytes/n); print(" C", c, round(a.nbytes/c[0]/1e6,1),"GB/s"); print(" D", d, round(a.nbytes/d[0]/1e6,1),"GB/s")
