"""Aggregate throughput of batches of independent fields (BASELINE configs 3 / 4 and a batch of config-2 fields) through
szb200_compress_batch_device / szb200_decompress_batch_device for 1, 2, 4, 8 worker lanes.
    python tools/bench_batch.py [cfg2|cfg3|cfg4] [n_fields]"""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import bench, sz_b200

which = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
CFG = {"cfg2": ((512, 512, 512), sz_b200.REL, 1e-4, 1e-3, 6), "cfg3": ((100, 500, 500), sz_b200.ABS, 1e-2, 1e-3, 13),
       "cfg4": ((1800, 3600), sz_b200.REL, 1e-4, 1e-4, 77)}
shape, mode, ab, rl, nf = CFG[which]
if len(sys.argv) > 2:
    nf = int(sys.argv[2])
sz_b200.set_device(0)
sz_b200.init(dict(errorBoundMode={sz_b200.ABS: "ABS", sz_b200.REL: "REL"}[mode], absErrBound=ab, relBoundRatio=rl))
if len(shape) == 3:
    base = torch.from_numpy(bench.make_field(shape)).cuda()
else:
    import szref
    base = torch.from_numpy(szref.smooth_field(shape, np.float32, span=8 * np.pi, field_id=5)).cuda()
# distinct fields of the same statistics: the base field with a small per-field offset and scale
xs = [(base * (1.0 + 0.01 * k) + 0.003 * k).contiguous() for k in range(nf)]
cap = sz_b200.compress_bound(np.float32, shape)
outs = [torch.empty(cap, dtype=torch.uint8, device="cuda") for _ in range(nf)]
recs = [torch.empty_like(xs[0]) for _ in range(nf)]
nbytes = nf * xs[0].numel() * 4
res = {"config": which, "shape": shape, "fields": nf, "bytes": nbytes, "lanes": {}}
for lanes in (1, 2, 4, 8):
    sizes = sz_b200.compress_batch_device(xs, outs, mode, ab, rl, lanes=lanes)      # warm-up (allocates the lanes' workspaces)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    sizes = sz_b200.compress_batch_device(xs, outs, mode, ab, rl, lanes=lanes)
    torch.cuda.synchronize(); tc = time.perf_counter() - t0
    sz_b200.decompress_batch_device(outs, sizes, recs, lanes=lanes)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    sz_b200.decompress_batch_device(outs, sizes, recs, lanes=lanes)
    torch.cuda.synchronize(); td = time.perf_counter() - t0
    err = max(float((r.double() - x.double()).abs().max()) for r, x in zip(recs[:3], xs[:3]))
    res["lanes"][lanes] = {"compress_ms": tc * 1e3, "compress_GBps": nbytes / tc / 1e9, "decompress_ms": td * 1e3, "decompress_GBps": nbytes / td / 1e9,
                           "ratio": nbytes / float(sum(sizes)), "max_abs_err_first3": err}
    print(lanes, res["lanes"][lanes], flush=True)
print(json.dumps(res))
