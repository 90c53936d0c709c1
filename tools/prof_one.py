"""Small driver for ncu captures: one device-resident compress + decompress of an n^3 fp32 smooth field (REL 1e-3).
    python tools/prof_one.py [n] [repeats]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402
import sz_b200  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
rep = int(sys.argv[2]) if len(sys.argv) > 2 else 1
sz_b200.set_device(0)
sz_b200.init(dict(errorBoundMode="REL", relBoundRatio=1e-3))
a = bench.make_field((n, n, n))
x = torch.from_numpy(a).cuda()
buf = torch.empty(sz_b200.compress_bound(np.float32, a.shape), dtype=torch.uint8, device="cuda")
out = torch.empty_like(x)
for _ in range(rep):
    nb = sz_b200.compress_device(x, buf, sz_b200.REL, 1e-4, 1e-3)
    st = sz_b200.last_stats()
    print("compress %d -> %d bytes, %.3f ms, launches %d, reg %d/%d" % (a.nbytes, nb, st.ms_total, st.launches, st.reg_count, st.num_blocks), {k: round(v, 3) for k, v in st.as_dict()["ms_stage"].items() if v > 0})
    sz_b200.decompress_device(buf, nb, out)
    st = sz_b200.last_stats()
    print("decompress %.3f ms" % st.ms_total, {k: round(v, 3) for k, v in st.as_dict()["ms_stage"].items() if v > 0})
torch.cuda.synchronize()
