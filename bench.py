#!/usr/bin/env python
"""bench.py -- headline benchmark of the SZ2 prediction/quantisation/entropy hot path on B200.

    python bench.py --gpus N --steps K --warmup W            (N>1: launched by torchrun, one rank per GPU)
    python bench.py --impl reference --gpus N --steps K --warmup W

Workload (BASELINE.json configs[1]): NYX-shape 512x512x512 fp32 synthetic smooth field, REL 1e-3, blocked regression +
Lorenzo, szMode=SZ_BEST_SPEED.  One "step" = one SZ_compress_args of one field per GPU (weak scaling: every rank owns one
512^3 slab of a (N*512)x512x512 array; for N>1 the step ends with the NCCL all-gather of stream sizes + streams).
`value` = uncompressed GB/s of the whole job with the input already resident in HBM (CUDA events on the library's stream,
max over ranks); `e2e` = the same through the host-pointer C API (pinned host input, H2D + kernels + D2H inside the timer).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
_JSON_OUT = None          # duplicate of the real stdout when file descriptor 1 had to be redirected (multi-rank runs)
sys.path.insert(0, ROOT)

SHAPE = (512, 512, 512)
REL_BOUND = 1e-3
METRIC = "compress_GBps"
UNIT = "GB/s"


def make_field(shape, field_id=0, dtype=np.float32):
    """SURVEY.md 8d synthetic smooth field: sin(x)cos(1.3y)sin(0.7z+0.3) + 0.25 sin(3.1x+1.7y)cos(2.3z) + 1e-3 N(0,1),
    coordinates linspace(0, 8pi, n), numpy default_rng(1234 + field_id), phase shift 0.37*field_id."""
    rng = np.random.default_rng(1234 + field_id)
    z, y, x = np.meshgrid(*[np.linspace(0, 8 * np.pi, n, dtype=np.float64) for n in shape], indexing="ij", sparse=True)
    ph = 0.37 * field_id
    f = (np.sin(x + ph) * np.cos(1.3 * y)) * np.sin(0.7 * z + 0.3)
    f = f + (0.25 * np.sin(3.1 * x + 1.7 * y)) * np.cos(2.3 * z)
    out = f.astype(dtype)
    del f
    out += (1e-3 * rng.standard_normal(shape, dtype=np.float32)).astype(dtype)
    return np.ascontiguousarray(out)


class ClockSampler:
    """samples nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:  # noqa: BLE001
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = sorted(int(float(r[0])) for r in self.rows if r and r[0].replace(".", "").isdigit())
        mx = [int(float(r[1])) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 7 for n, v in zip(names, r[3:7]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


def peak_hbm():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ----------------------------------------------------------------------------------------------------------------------
def ref_lib():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import szref
    if not szref.have_ref():
        return None, szref
    return szref.SZLib(szref.REF_SO), szref


def _omp_variant(a, ncores):
    """the reference's OpenMP compressor (sz/src/sz_omp.c:63, float 3D, ABS only, its own stream format) on the same field, in a
    child process; a reported baseline beside the slab-parallel one.  None when the library is absent."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import szref
    if not os.path.exists(szref.REF_OMP_SO):
        return None
    threads = 1
    while threads * 2 <= ncores:
        threads *= 2
    code = (
        "import sys, os, time, ctypes as C, numpy as np\n"
        "sys.path.insert(0, %r)\n"
        "import szref\n"
        "a = np.load(%r)\n"
        "L = C.CDLL(szref.REF_OMP_SO)\n"
        "cfg = os.path.join(os.path.dirname(%r), 'omp.config'); szref.write_config(cfg, dict(errorBoundMode='ABS', absErrBound=1e-3))\n"
        "L.SZ_Init.argtypes = [C.c_char_p]; assert L.SZ_Init(cfg.encode()) == 0\n"
        "f = L.SZ_compress_float_3D_MDQ_openmp; f.restype = C.c_void_p\n"
        "f.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_float, C.POINTER(C.c_size_t)]\n"
        "eb = %r * float(a.max() - a.min()); n = C.c_size_t(0); best = 1e9\n"
        "for _ in range(3):\n"
        "    t0 = time.perf_counter(); p = f(a.ctypes.data, a.shape[0], a.shape[1], a.shape[2], eb, C.byref(n)); best = min(best, time.perf_counter() - t0)\n"
        "print('OMP', best, n.value)\n"
    )
    import tempfile
    tmp = tempfile.NamedTemporaryFile(suffix=".npy", delete=False)
    np.save(tmp.name, a)
    env = dict(os.environ, OMP_NUM_THREADS=str(threads))
    try:
        out = subprocess.run([sys.executable, "-c", code % (os.path.join(ROOT, "tests"), tmp.name, tmp.name, REL_BOUND)], env=env, capture_output=True, text=True, timeout=300)
        for ln in out.stdout.splitlines():
            if ln.startswith("OMP "):
                _, t, nbytes = ln.split()
                return {"value": a.nbytes / float(t) / 1e9, "unit": UNIT, "threads": threads, "stream_bytes": int(nbytes),
                        "note": "reference sz_omp.c float 3D, ABS bound = rel x range; a different stream format, timing baseline only"}
    except Exception:
        pass
    finally:
        os.unlink(tmp.name)
    return None


def _ref_worker(args):
    """one slab compressed by the UNMODIFIED reference in its own process (its global state is not thread-safe)."""
    path, lo, hi, shape, rel = args
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import szref
    a = np.load(path, mmap_mode="r")[lo:hi]
    a = np.ascontiguousarray(a)
    lib = szref.SZLib(szref.REF_SO)
    lib.init(dict(errorBoundMode="REL", relBoundRatio=rel, absErrBound=1e-4))
    t0 = time.perf_counter()
    s = lib.compress(a, szref.REL, 1e-4, rel)
    dt = time.perf_counter() - t0
    lib.finalize()
    return dt, len(s)


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path (oracle/_ref/libSZ_ref.so, compiled from the
    reference's sources), all host cores via its slab-parallel pattern (example/test_threadsafe.c:40-58), same config."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    lib, szref = ref_lib()
    ncores = os.cpu_count() or 1
    nworkers = max(1, min(ncores, 64))
    # bounded sample: one 512^3 field cut into `nworkers` slabs along the slowest dimension
    import multiprocessing as mp
    import tempfile
    a = make_field(SHAPE)
    tmp = tempfile.NamedTemporaryFile(suffix=".npy", delete=False)
    np.save(tmp.name, a)
    from sz_b200.slabs import slab_bounds
    bounds = [b for b in slab_bounds(SHAPE[0], nworkers) if b[1] - b[0] >= 12]
    if lib is None:
        # no compiled reference on this box: the oracle port stands in (kind "port")
        kind = "port"
    else:
        kind = "reference"
    times = []
    sizes = None
    ctx = mp.get_context("spawn")
    with ctx.Pool(len(bounds)) as pool:
        jobs = [(tmp.name, lo, hi, SHAPE, REL_BOUND) for lo, hi in bounds]
        for it in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            res = pool.map(_ref_worker, jobs)
            dt = time.perf_counter() - t0
            if it >= args.warmup:
                times.append(dt)
            sizes = [r[1] for r in res]
    os.unlink(tmp.name)
    ms = 1e3 * sum(times) / len(times)
    gbps = a.nbytes / (ms * 1e-3) / 1e9
    omp = _omp_variant(a, ncores) if lib is not None else None
    line = {"impl": "reference", "metric": METRIC, "value": gbps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "NYX-shape 512x512x512 fp32 synthetic smooth field, REL 1e-3, SZ_BEST_SPEED (BASELINE.json configs[1])",
                       "parallelism": "%d slab processes on %d host cores" % (len(bounds), ncores)},
            "cpu_baseline": {"value": gbps, "unit": UNIT, "cores": len(bounds), "kind": kind,
                             "sample": "one 512^3 field per step, cut into %d independent slabs (reference's slab-parallel pattern)" % len(bounds)},
            "e2e": {"value": gbps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "compressed_bytes": int(sum(sizes)), "ratio": a.nbytes / float(sum(sizes))}
    if omp is not None:
        line["openmp_variant"] = omp
    print(json.dumps(line))


# ----------------------------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import sz_b200
    from sz_b200 import api, slabs
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback exists)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        # NCCL prints its version banner on the C stdout; stdout must carry the one JSON line only, so file descriptor 1 is
        # pointed at stderr for the run and the JSON line goes to a duplicate of the real stdout
        global _JSON_OUT
        sys.stdout.flush()
        _JSON_OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sz_b200.set_device(local)
    sz_b200.init(dict(errorBoundMode="REL", relBoundRatio=REL_BOUND, absErrBound=1e-4))
    L = sz_b200.lib()
    lib_stream = torch.cuda.ExternalStream(L.szb200_stream())

    host = torch.from_numpy(make_field(SHAPE, field_id=rank)).pin_memory()
    x = host.cuda(non_blocking=False)
    nbytes = x.numel() * 4
    cap = sz_b200.compress_bound(np.float32, SHAPE)
    stream_buf = torch.empty(cap, dtype=torch.uint8, device="cuda")
    recon = torch.empty_like(x)

    def step():
        n = sz_b200.compress_device(x, stream_buf, sz_b200.REL, 1e-4, REL_BOUND)
        if world > 1:
            slabs.gather_streams(stream_buf, n)
        return n

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        n = step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stage_ms, launches = {}, 0
    wave_ms = []
    ev0.record(lib_stream)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        n = step()
        st = sz_b200.last_stats()
        launches += st.launches
        for k, v in st.as_dict()["ms_stage"].items():
            stage_ms[k] = stage_ms.get(k, 0.0) + v
        wave_ms.append(st.ms_stage[4])
    ev1.record(lib_stream)
    barrier()
    wall = time.perf_counter() - t0
    ms_dev = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([ms_dev, wall * 1e3], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t[0]) / args.steps
    value = world * nbytes / (ms_step * 1e-3) / 1e9

    # ---- decompress (device resident), same timing rules ----
    for _ in range(max(1, args.warmup // 2)):
        sz_b200.decompress_device(stream_buf, n, recon)
    barrier()
    d0, d1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    d0.record(lib_stream)
    dsteps = max(1, args.steps)
    dstage = {}
    for _ in range(dsteps):
        sz_b200.decompress_device(stream_buf, n, recon)
        st = sz_b200.last_stats()
        for k, v in st.as_dict()["ms_stage"].items():
            dstage[k] = dstage.get(k, 0.0) + v
    d1.record(lib_stream)
    barrier()
    td = torch.tensor([d0.elapsed_time(d1)], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(td, op=dist.ReduceOp.MAX)
    dms = float(td[0]) / dsteps
    eb = REL_BOUND * float((x.max() - x.min()).item())
    max_err = float((recon.double() - x.double()).abs().max().item())

    # ---- e2e through the host-pointer C API (pinned host input; H2D + kernels + D2H inside the timer) ----
    harr = host.numpy()
    # compress_view = SZ_compress_args on the host array; the returned numpy array wraps the malloc'ed stream of the C call
    for _ in range(2):
        s_host = sz_b200.compress_view(harr, sz_b200.REL, 1e-4, REL_BOUND)
    barrier()
    e_steps = max(2, min(args.steps, 5))
    t0 = time.perf_counter()
    for _ in range(e_steps):
        s_host = sz_b200.compress_view(harr, sz_b200.REL, 1e-4, REL_BOUND)
    torch.cuda.synchronize()
    e_ms = (time.perf_counter() - t0) * 1e3 / e_steps
    te = torch.tensor([e_ms], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e_ms = float(te[0])
    e2e = world * nbytes / (e_ms * 1e-3) / 1e9

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    peak, peak_src = peak_hbm()
    algo_bytes = nbytes + n                                  # N*sizeof(T) read + outSize written
    wave = sum(wave_ms) / len(wave_ms)
    # DRAM bytes of the same stage from the committed `ncu --set full` capture (profiles/), per launch; None if not captured
    traffic, traffic_src = None, None
    tj = os.path.join(ROOT, "profiles", "r01_wavefront_traffic.json")
    if os.path.exists(tj):
        try:
            with open(tj) as f:
                tinfo = json.load(f)
            traffic, traffic_src = tinfo.get("dram_bytes_per_launch"), "profiles/r01_wavefront_traffic.json"
        except Exception:
            pass
    side = stage_ms.get("wait_coeff_trees_side_stream", 0.0) / args.steps
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "NYX-shape 512x512x512 fp32 synthetic smooth field, REL 1e-3, SZ_BEST_SPEED (BASELINE.json configs[1])",
                       "per_gpu_bytes": nbytes, "l2": "input (512 MiB) larger than L2, no flush needed",
                       "parallelism": "1 slab per GPU%s" % ("; NCCL all-gather of sizes + streams inside the step" if world > 1 else "")},
            "compressed_bytes": int(n), "ratio": nbytes / float(n),
            "decompress": {"value": world * nbytes / (dms * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": dms, "max_abs_err": max_err, "eb": eb,
                           "within_bound": bool(max_err <= eb), "stage_ms": {k: v / dsteps for k, v in dstage.items() if v > 0}},
            "stage_ms": {k: v / args.steps for k, v in stage_ms.items() if v > 0},
            "roofline": {"bound": "hbm", "kernel": "prepass_compress_kernel + wavefront_kernel<float,3> (predict + quantise: the stage that touches every byte)",
                         "achieved": algo_bytes / (wave * 1e-3) / 1e9, "peak": peak,
                         "unit": "GB/s", "frac": algo_bytes / (wave * 1e-3) / 1e9 / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                         "algorithmic_bytes": algo_bytes, "kernel_ms": wave, "job_frac": algo_bytes / (ms_step * 1e-3) / 1e9 / peak,
                         "note": "the wavefront is bound by its dependency chain (latency), not by HBM; the largest share of the step is the wait "
                                 "for the four coefficient Huffman trees (exact emulation of the reference's sequential heap, two warps per tree, side stream): "
                                 "%.2f ms of %.2f ms" % (side, ms_step)},
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": len(s_host), "ms_per_step": e_ms},
            "gpu_launches": int(launches), "clocks": clocks}
    if world == 1:
        line["cpu_baseline"] = cpu_baseline(harr, s_host)
    print(json.dumps(line), file=_JSON_OUT or sys.stdout, flush=True)
    if dist is not None:
        dist.destroy_process_group()


def cpu_baseline(harr, gpu_stream):
    """the reference (oracle/_ref, compiled from its own sources) timed on ONE host core on the same field; also the parity
    check at full size: the GPU stream must equal the reference stream byte for byte."""
    lib, szref = ref_lib()
    import hashlib
    if lib is not None:
        kind = "reference"
        lib.init(dict(errorBoundMode="REL", relBoundRatio=REL_BOUND, absErrBound=1e-4))
        t0 = time.perf_counter()
        s = lib.compress(harr, szref.REL, 1e-4, REL_BOUND)
        dt = time.perf_counter() - t0
        t0 = time.perf_counter()
        lib.decompress(s, np.float32, harr.shape)
        ddt = time.perf_counter() - t0
        lib.finalize()
    elif szref.have_oracle():
        kind = "port"
        orc = szref.Oracle()
        orc.init(dict(errorBoundMode="REL", relBoundRatio=REL_BOUND, absErrBound=1e-4))
        t0 = time.perf_counter()
        s = orc.compress(harr, szref.REL, 1e-4, REL_BOUND)
        dt = time.perf_counter() - t0
        ddt = None
    else:
        return None
    return {"value": harr.nbytes / dt / 1e9, "unit": UNIT, "cores": 1, "kind": kind, "sample": "the full 512^3 field, one serial SZ_compress_args call",
            "decompress_value": None if ddt is None else harr.nbytes / ddt / 1e9, "stream_bytes": len(s),
            "gpu_stream_equals_reference": bool(bytes(s) == bytes(gpu_stream)), "stream_md5": hashlib.md5(s).hexdigest()}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
