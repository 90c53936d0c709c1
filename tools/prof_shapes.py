import sys,os
sys.path.insert(0,"tests"); sys.path.insert(0,".")
import numpy as np, torch, szref, sz_b200
sz_b200.set_device(0); sz_b200.init(dict(errorBoundMode="REL",relBoundRatio=1e-4))
for shape,mode,ab,rl in (((1800,3600),sz_b200.REL,1e-4,1e-4),((100,500,500),sz_b200.ABS,1e-2,1e-3)):
    a=szref.smooth_field(shape,np.float32,span=8*np.pi,field_id=5); x=torch.from_numpy(a).cuda(); buf=torch.empty(sz_b200.compress_bound(np.float32,a.shape),dtype=torch.uint8,device="cuda"); out=torch.empty_like(x)
    for _ in range(3):
        n=sz_b200.compress_device(x,buf,mode,ab,rl); st=sz_b200.last_stats(); c=(round(st.ms_total,3), {k: round(v,3) for k,v in st.as_dict()["ms_stage"].items() if v>0.004})
        sz_b200.decompress_device(buf,n,out); st=sz_b200.last_stats(); d=(round(st.ms_total,3), {k: round(v,3) for k,v in st.as_dict()["ms_stage"].items() if v>0.004})
    print(shape, n, "reg", st.reg_count, "/", st.num_blocks); print(" C", c); print(" D", d)
