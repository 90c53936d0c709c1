"""Where the end-to-end time of the host-pointer call goes: SZB200_HOST_PROFILE=1 python tools/e2e_probe.py [n]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench, sz_b200
n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
sz_b200.set_device(0)
sz_b200.init(dict(errorBoundMode="REL", relBoundRatio=1e-3))
a = bench.make_field((n, n, n))
host = torch.from_numpy(a).pin_memory()
harr = host.numpy()
x = torch.empty_like(host, device="cuda")
for _ in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter(); x.copy_(host, non_blocking=True); torch.cuda.synchronize()
    print("torch pinned H2D of %d MiB: %.2f ms" % (a.nbytes >> 20, (time.perf_counter() - t0) * 1e3))
for _ in range(4):
    t0 = time.perf_counter(); s = sz_b200.compress(harr, sz_b200.REL, 1e-4, 1e-3); t1 = time.perf_counter()
    st = sz_b200.last_stats()
    print("compress() wall %.2f ms, %d bytes" % ((t1 - t0) * 1e3, len(s)), {k: round(v, 2) for k, v in st.as_dict()["ms_stage"].items() if v > 0})
for _ in range(3):
    t0 = time.perf_counter(); s2 = sz_b200.compress_view(a, sz_b200.REL, 1e-4, 1e-3); t1 = time.perf_counter()
    print("compress_view() from PAGEABLE input: wall %.2f ms, equal=%s" % ((t1 - t0) * 1e3, bytes(s2) == bytes(s)))
for _ in range(3):
    t0 = time.perf_counter(); b = sz_b200.decompress(s, np.float32, a.shape); t1 = time.perf_counter()
    print("decompress() wall %.2f ms" % ((t1 - t0) * 1e3))
