"""BASELINE config 5 at full size (S3D-shape: 11 x 500 x 500 x 500 fp64, ABS_AND_REL) as ONE device-resident call:
round trip, error bound, determinism.  The field is generated on the device.   python tools/try_cfg5.py [r4]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import sz_b200
r4 = int(sys.argv[1]) if len(sys.argv) > 1 else 11
shape = (r4, 500, 500, 500)
sz_b200.set_device(0)
sz_b200.init(dict(errorBoundMode="ABS_AND_REL", absErrBound=1e-3, relBoundRatio=1e-3))
dev = "cuda"
x = torch.empty(shape, dtype=torch.float64, device=dev)
g = [torch.linspace(0, 8 * np.pi, n, dtype=torch.float64, device=dev) for n in shape[1:]]
for s in range(r4):      # slab by slab to bound temporary memory
    z, y, xx = g[0][:, None, None], g[1][None, :, None], g[2][None, None, :]
    x[s] = torch.sin(xx) * torch.cos(1.3 * y) * torch.sin(0.7 * z + 0.3 + 0.37 * s) + 0.25 * torch.sin(3.1 * xx + 1.7 * y) * torch.cos(2.3 * z)
    x[s] += 1e-3 * torch.randn(shape[1:], dtype=torch.float64, device=dev, generator=torch.Generator(device=dev).manual_seed(1234 + s))
torch.cuda.synchronize()
nbytes = x.numel() * 8
cap = sz_b200.compress_bound(np.float64, shape)
buf = torch.empty(cap, dtype=torch.uint8, device=dev)
for rep in range(2):
    t0 = time.perf_counter(); n = sz_b200.compress_device(x, buf, sz_b200.ABS_AND_REL, 1e-3, 1e-3); torch.cuda.synchronize(); dt = time.perf_counter() - t0
    st = sz_b200.last_stats()
    print("compress %.1f GB -> %d bytes (ratio %.2f) in %.1f ms = %.1f GB/s" % (nbytes / 1e9, n, nbytes / n, dt * 1e3, nbytes / dt / 1e9),
          {k: round(v, 1) for k, v in st.as_dict()["ms_stage"].items() if v > 0.05}, flush=True)
first = buf[:n].clone()
out = torch.empty_like(x)
t0 = time.perf_counter(); sz_b200.decompress_device(buf, n, out); torch.cuda.synchronize(); dt = time.perf_counter() - t0
print("decompress in %.1f ms = %.1f GB/s" % (dt * 1e3, nbytes / dt / 1e9))
rng = float(x.max() - x.min()); eb = min(1e-3, 1e-3 * rng)
err = 0.0
for s in range(r4):
    err = max(err, float((out[s] - x[s]).abs().max()))
print("max abs err %.3e  eb %.3e  within bound: %s" % (err, eb, err <= eb))
n2 = sz_b200.compress_device(x, buf, sz_b200.ABS_AND_REL, 1e-3, 1e-3)
print("deterministic:", n2 == n and bool(torch.equal(buf[:n], first)))
print("peak device memory (torch + library not included): %.1f GB allocated by torch" % (torch.cuda.max_memory_allocated() / 1e9))
free, total = torch.cuda.mem_get_info(); print("device memory in use: %.1f GB of %.1f" % ((total - free) / 1e9, total / 1e9))
