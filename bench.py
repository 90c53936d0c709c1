#!/usr/bin/env python
"""bench.py -- headline benchmark of the SZ2 prediction/quantisation/entropy hot path on B200.

    python bench.py --gpus N --steps K --warmup W            (N>1: launched by torchrun, one rank per GPU)
    python bench.py --impl reference --gpus N --steps K --warmup W

Workload (BASELINE.json configs[1]): NYX-shape 512x512x512 fp32 synthetic smooth field, REL 1e-3, blocked regression +
Lorenzo, szMode=SZ_BEST_SPEED.  One "step" = one SZ_compress_args of one field per GPU (weak scaling: every rank owns one
512^3 slab of a (N*512)x512x512 array; for N>1 the step ends with the NCCL all-gather of stream sizes + streams).
`value` = uncompressed GB/s of the whole job with the input already resident in HBM (CUDA events on the library's stream,
max over ranks); `e2e` = the same through the host-pointer C API (pinned host input, H2D + kernels + D2H inside the timer);
`e2e_pageable` = the same from malloc'ed memory, which is what the reference's own callers pass.  `roofline` is reported for
the kernel with the largest share of the step, measured live.  `extra` holds the other BASELINE configs: the 13- and 77-field
batches (configs[2], [3]) through the batch call and the 11 GB fp64 array (configs[4]) cut into one slab per rank (strong scaling).
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
WORKLOAD = "NYX-shape 512x512x512 fp32 synthetic smooth field, REL 1e-3, SZ_BEST_SPEED (BASELINE.json configs[1])"


def shared_config():
    """the `config` object, identical in both arms (what differs between the arms is said in `arm`)"""
    return {"workload": WORKLOAD, "per_gpu_bytes": int(np.prod(SHAPE)) * 4, "l2": "input (512 MiB) larger than L2, no flush needed",
            "parallelism": "independent SZ slabs, one 512^3 slab per GPU (the reference's example/test_threadsafe.c pattern)"}


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
    times, walls = [], []
    sizes = None
    ctx = mp.get_context("spawn")
    with ctx.Pool(len(bounds)) as pool:
        jobs = [(tmp.name, lo, hi, SHAPE, REL_BOUND) for lo, hi in bounds]
        for it in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            res = pool.map(_ref_worker, jobs)
            dt = time.perf_counter() - t0
            if it >= args.warmup:
                times.append(max(r[0] for r in res))          # the slowest slab, timed around SZ_compress_args only
                walls.append(dt)                               # with the workers' load of their slab, config file and SZ_Init
            sizes = [r[1] for r in res]
    os.unlink(tmp.name)
    ms = 1e3 * sum(times) / len(times)
    gbps = a.nbytes / (ms * 1e-3) / 1e9
    omp = _omp_variant(a, ncores) if lib is not None else None
    line = {"impl": "reference", "metric": METRIC, "value": gbps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": shared_config(),
            "arm": "%d slab processes on %d host cores, one 512^3 field per step; the step time is the slowest worker's SZ_compress_args" % (len(bounds), ncores),
            "wall_ms_per_step_with_worker_setup": 1e3 * sum(walls) / len(walls),
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

    # N > 1: the gather of step i (NCCL, its own stream) runs beside the compress of step i + 1; stream and gather buffers alternate,
    # and a buffer is reused only after the gather that read it has finished (the library orders itself after torch's current stream).
    # The gather's host side (one read-back of the sizes, N broadcast launches) runs on a helper thread while the main thread is inside
    # the next compress call (ctypes releases the interpreter lock there); all collectives are issued by that one thread, in step order.
    stream_bufs = [stream_buf, torch.empty(cap, dtype=torch.uint8, device="cuda")] if world > 1 else [stream_buf]
    gather_bufs = [torch.empty(world * (cap // 8), dtype=torch.uint8, device="cuda") for _ in range(2)] if world > 1 else None
    comm_stream = torch.cuda.Stream() if world > 1 else None
    gather_done = [None, None]
    gather_fut = [None, None]
    step_no = [0]
    pool = None
    if world > 1:
        from concurrent.futures import ThreadPoolExecutor
        pool = ThreadPoolExecutor(max_workers=1)

    def gather_job(b, n):
        torch.cuda.set_device(local)
        with torch.cuda.stream(comm_stream):
            slabs.gather_streams(stream_bufs[b], n, out=gather_bufs[b])
            ev = torch.cuda.Event()
            ev.record(comm_stream)
        gather_done[b] = ev

    def drain():
        for b in range(2):
            if gather_fut[b] is not None:
                gather_fut[b].result()
                gather_fut[b] = None

    def step():
        b = step_no[0] & 1 if world > 1 else 0
        step_no[0] += 1
        if world > 1 and gather_fut[b] is not None:
            gather_fut[b].result()                                  # the gather that last read this buffer has been queued ...
            gather_fut[b] = None
            torch.cuda.current_stream().wait_event(gather_done[b])   # ... and the library waits for it on the device
        n = sz_b200.compress_device(x, stream_bufs[b], sz_b200.REL, 1e-4, REL_BOUND)        # returns when the stream is complete
        if world > 1:
            gather_fut[b] = pool.submit(gather_job, b, n)
        return n

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        t = torch.tensor([float(v)], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    for _ in range(args.warmup):
        n = step()
    drain()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stage_ms, launches = {}, 0
    ev0.record(lib_stream)
    for _ in range(args.steps):
        n = step()
    drain()                                                         # every gather is queued: the closing event goes behind the last one
    # (the gather of an N>1 step is enqueued on the communication stream after the library's job has finished: the closing event goes
    #  where the last work of the last step went)
    ev1.record(lib_stream if world == 1 else comm_stream)
    barrier()
    ms_dev = ev0.elapsed_time(ev1)
    # launch count and per-stage times: the library's own record of a step, read outside the timed region (every step launches the same)
    st = sz_b200.last_stats()
    launches = st.launches * args.steps
    for k, v in st.as_dict()["ms_stage"].items():
        stage_ms[k] = v * args.steps
    clocks = sampler.stop() if rank == 0 else None
    ms_step = max_over_ranks(ms_dev) / args.steps
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
    dms = max_over_ranks(d0.elapsed_time(d1)) / dsteps
    eb = REL_BOUND * float((x.max() - x.min()).item())
    max_err = float((recon.double() - x.double()).abs().max().item())

    # ---- e2e through the host-pointer C API (H2D + kernels + D2H inside the timer): pinned input, pageable (malloc'ed) input,
    #      and the way back (SZ_decompress of a host stream into a malloc'ed array) ----
    harr = host.numpy()
    hpage = np.array(harr, copy=True)                        # ordinary pageable memory, as example/sz.c reads its input file into
    e_steps = max(2, min(args.steps, 5))

    def timed(fn):
        for _ in range(2):
            r = fn()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e_steps):
            r = fn()
        torch.cuda.synchronize()
        return max_over_ranks((time.perf_counter() - t0) * 1e3 / e_steps), r

    # compress_view = SZ_compress_args on the host array; the returned numpy array wraps the malloc'ed stream of the C call
    e_ms, s_host = timed(lambda: sz_b200.compress_view(harr, sz_b200.REL, 1e-4, REL_BOUND))
    ep_ms, s_page = timed(lambda: sz_b200.compress_view(hpage, sz_b200.REL, 1e-4, REL_BOUND))
    s_bytes = bytes(s_host)
    ed_ms, rec_host = timed(lambda: sz_b200.decompress(s_bytes, np.float32, SHAPE))
    e2e = world * nbytes / (e_ms * 1e-3) / 1e9

    extra = run_extras(args, torch, dist, sz_b200, slabs, rank, world, barrier, max_over_ranks)

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    peak, peak_src = peak_hbm()
    algo_bytes = nbytes + n                                  # N*sizeof(T) read + outSize written
    per_step = {k: v / args.steps for k, v in stage_ms.items() if v > 0}
    # the kernels (or launch groups) a stage consists of, for the roofline line: the one with the largest measured share
    kernels = {"coeff_trees_kernel_side_stream": "huff_build_kernel (four coefficient Huffman trees: the reference's binary heap replayed operation by operation, side stream)",
               "coeff_chain": "coeff_lattice_kernel (+ reg_gather: the coefficient DPCM recurrence)",
               "wavefront_quant": "prepass_compress_kernel + wavefront_kernel<float,3> (predict + quantise)",
               "encode": "enc_len_kernel + scans + enc_write_kernel (Huffman bit packing)",
               "fit_select": "fit_select_kernel (regression fit + predictor selection)",
               "opt_intervals": "sample_mean / opt_sample / opt_finish kernels (interval optimiser)",
               "histogram": "hist_kernel"}
    dom = max((k for k in kernels if k in per_step), key=lambda k: per_step[k])
    dom_ms = per_step[dom]
    traffic = traffic_step = None
    traffic_src = None
    tj = os.path.join(ROOT, "profiles", "r02_dram_traffic_512cube.json")
    if os.path.exists(tj):
        try:
            with open(tj) as f:
                tinfo = json.load(f)
            traffic = tinfo.get("by_stage", {}).get(dom)
            traffic_step = tinfo.get("all_launches_of_one_step")
            traffic_src = "profiles/r02_dram_traffic_512cube.json (ncu dram__bytes_read.sum + dram__bytes_write.sum, " + tinfo.get("source", "") + ")"
        except Exception:  # noqa: BLE001
            pass
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": shared_config(),
            "arm": "one device-resident call per step and GPU%s" % ("; gather of sizes + exact-length streams over NCCL inside the timed region, on its own stream: the gather of step i runs beside the compress of step i + 1 (alternating buffers), the last one is waited for before the clock stops" if world > 1 else ""),
            "compressed_bytes": int(n), "ratio": nbytes / float(n),
            "decompress": {"value": world * nbytes / (dms * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": dms, "max_abs_err": max_err, "eb": eb,
                           "within_bound": bool(max_err <= eb), "stage_ms": {k: v / dsteps for k, v in dstage.items() if v > 0}},
            "stage_ms": per_step,
            "roofline": {"bound": "hbm", "kernel": kernels[dom], "stage": dom,
                         "achieved": algo_bytes / (dom_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s", "frac": algo_bytes / (dom_ms * 1e-3) / 1e9 / peak,
                         "traffic": traffic, "traffic_all_launches_of_one_step": traffic_step,
                         "traffic_over_algorithmic": None if not traffic_step else traffic_step / float(algo_bytes),
                         "traffic_source": traffic_src, "peak_source": peak_src,
                         "algorithmic_bytes": algo_bytes, "kernel_ms": dom_ms,
                         "kernel_ms_source": ("the library's per-stage record: CUDA events on the launching stream; the coefficient-tree kernel is queued before the chain "
                                              "kernel and waits on its SMs for the chain's flags, so its duration is read with %globaltimer inside the kernel, behind "
                                              "that wait (events around the launch would add the chain's 1 ms)"),
                         "job_frac": algo_bytes / (ms_step * 1e-3) / 1e9 / peak,
                         "note": "the dominant kernel is a serial dependency chain of the stream format (exact heap order decides the Huffman "
                                 "tree), bound by instruction latency on a few SMs, not by HBM: its fraction measures how far the step is from "
                                 "the bandwidth bound, not how well it uses bandwidth"},
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": len(s_host), "ms_per_step": e_ms, "host_memory": "pinned"},
            "e2e_pageable": {"value": world * nbytes / (ep_ms * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": ep_ms, "host_memory": "malloc"},
            "e2e_decompress": {"value": world * nbytes / (ed_ms * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": ed_ms, "h2d_bytes_per_step": len(s_host),
                               "d2h_bytes_per_step": nbytes, "host_memory": "malloc (stream bytes in, numpy array out)"},
            "extra": extra,
            "gpu_launches": int(launches), "clocks": clocks}
    ok = True
    if world == 1:
        line["cpu_baseline"] = cpu_baseline(harr, s_host)
        ok = line["cpu_baseline"] is None or bool(line["cpu_baseline"]["gpu_stream_equals_reference"])
        ok = ok and bytes(s_page) == s_bytes and rec_host.tobytes() == recon.cpu().numpy().tobytes()
    print(json.dumps(line), file=_JSON_OUT or sys.stdout, flush=True)
    if dist is not None:
        dist.destroy_process_group()
    if not ok:
        sys.stderr.write("bench.py: the GPU stream differs from the reference stream (or host and device paths disagree): the numbers above are void\n")
        sys.exit(3)


def field_on_device(torch, shape, field_id, dtype, dev="cuda"):
    """the SURVEY 8d smooth field (formula of make_field) evaluated on the device, for the large configurations"""
    g = [torch.linspace(0, 8 * np.pi, n, dtype=torch.float64, device=dev) for n in shape]
    ph = 0.37 * field_id
    if len(shape) == 2:
        y, x = g[0][:, None], g[1][None, :]
        f = torch.sin(x + ph) * torch.cos(1.3 * y) + 0.25 * torch.sin(3.1 * x + 1.7 * y)
    else:
        z, y, x = g[0][:, None, None], g[1][None, :, None], g[2][None, None, :]
        f = torch.sin(x + ph) * torch.cos(1.3 * y) * torch.sin(0.7 * z + 0.3) + 0.25 * torch.sin(3.1 * x + 1.7 * y) * torch.cos(2.3 * z)
    f += 1e-3 * torch.randn(shape, dtype=torch.float64, device=dev, generator=torch.Generator(device=dev).manual_seed(1234 + field_id))
    return f.to(dtype).contiguous()


def run_extras(args, torch, dist, sz_b200, slabs, rank, world, barrier, max_over_ranks):
    """The other BASELINE configs, device resident, as the lines of `extra`:
      configs[2]  13 fields 100x500x500 fp32, ABS 1e-2  -- batch call, the fields dealt round-robin to the ranks
      configs[3]  77 fields 1800x3600 fp32 (2D), REL 1e-4 -- the same
      configs[4]  500x500x500x11 fp64, ABS_AND_REL -- ONE array of 11 GB, cut into one slab per rank along the slowest dimension
                  (strong scaling: the same array at every N; at N = 1 it is one call), streams gathered over NCCL"""
    out = {}
    reps = 2

    def batch(name, nfields, shape, dtype, mode, ab, rl, cfg):
        mine = list(range(rank, nfields, world))
        sz_b200.init(cfg)
        xs = [field_on_device(torch, shape, i, dtype) for i in mine]
        outs = [torch.empty(sz_b200.compress_bound(np.float32 if dtype == torch.float32 else np.float64, shape), dtype=torch.uint8, device="cuda") for _ in mine]
        times, sizes = [], []
        for it in range(1 + reps):
            barrier()
            t0 = time.perf_counter()
            sizes = sz_b200.compress_batch_device(xs, outs, mode, ab, rl, lanes=4) if mine else []
            torch.cuda.synchronize()
            if it:
                times.append(time.perf_counter() - t0)
        dt = max_over_ranks(min(times))
        total = nfields * int(np.prod(shape)) * xs[0].element_size() if xs else nfields * int(np.prod(shape)) * 4
        recs = [torch.empty_like(v) for v in xs]
        barrier()
        t0 = time.perf_counter()
        if mine:
            sz_b200.decompress_batch_device(outs, sizes, recs, lanes=4)
        torch.cuda.synchronize()
        ddt = max_over_ranks(time.perf_counter() - t0)
        csum = torch.tensor([float(sum(sizes))], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(csum)
        out[name] = {"compress_GBps": total / dt / 1e9, "decompress_GBps": total / ddt / 1e9, "ms": dt * 1e3, "fields": nfields, "lanes_per_gpu": 4,
                     "ratio": total / float(csum[0]), "scaling": "strong (the batch is dealt to the ranks)"}
        del xs, outs, recs

    batch("config3_isabel_13x100x500x500_f32_abs1e-2", 13, (100, 500, 500), torch.float32, sz_b200.ABS, 1e-2, 1e-4,
          dict(errorBoundMode="ABS", absErrBound=1e-2))
    batch("config4_cesm_77x1800x3600_f32_rel1e-4", 77, (1800, 3600), torch.float32, sz_b200.REL, 1e-4, 1e-4,
          dict(errorBoundMode="REL", relBoundRatio=1e-4))

    # configs[4]: rows [lo, hi) of the flattened (5500, 500, 500) array
    cfg5 = dict(errorBoundMode="ABS_AND_REL", absErrBound=1e-3, relBoundRatio=1e-3)
    sz_b200.init(cfg5)
    lo, hi = slabs.slab_bounds(5500, world)[rank]
    free, _ = torch.cuda.mem_get_info()
    need = (hi - lo) * 250000 * 8 * 3.6
    if free < need:
        out["config5_s3d_500x500x500x11_f64_absrel"] = {"skipped": "needs %.0f GB of device memory per rank, %.0f free" % (need / 1e9, free / 1e9)}
    else:
        g = [torch.linspace(0, 8 * np.pi, 500, dtype=torch.float64, device="cuda") for _ in range(3)]
        x5 = torch.empty((hi - lo, 500, 500), dtype=torch.float64, device="cuda")
        for c0 in range(lo, hi, 250):                             # in pieces, to bound temporary memory
            c1 = min(hi, c0 + 250)
            rows = torch.arange(c0, c1, device="cuda")
            s4, z = (rows // 500).double()[:, None, None], g[0][rows % 500][:, None, None]
            y, xx = g[1][None, :, None], g[2][None, None, :]
            x5[c0 - lo:c1 - lo] = torch.sin(xx) * torch.cos(1.3 * y) * torch.sin(0.7 * z + 0.3 + 0.37 * s4) + 0.25 * torch.sin(3.1 * xx + 1.7 * y) * torch.cos(2.3 * z)
            x5[c0 - lo:c1 - lo] += 1e-3 * torch.randn((c1 - c0, 500, 500), dtype=torch.float64, device="cuda", generator=torch.Generator(device="cuda").manual_seed(99 + c0))
        buf5 = torch.empty(sz_b200.compress_bound(np.float64, tuple(x5.shape)), dtype=torch.uint8, device="cuda")
        gat5 = torch.empty(int(1.2e9), dtype=torch.uint8, device="cuda") if world > 1 else None
        times = []
        for it in range(1 + reps):
            barrier()
            t0 = time.perf_counter()
            n5 = sz_b200.compress_device(x5, buf5, sz_b200.ABS_AND_REL, 1e-3, 1e-3)
            if world > 1:
                sizes5, _ = slabs.gather_streams(buf5, n5, out=gat5)
            else:
                sizes5 = [n5]
            torch.cuda.synchronize()
            if it:
                times.append(time.perf_counter() - t0)
        dt = max_over_ranks(min(times))
        rec5 = torch.empty_like(x5)
        barrier()
        t0 = time.perf_counter()
        sz_b200.decompress_device(buf5, n5, rec5)
        torch.cuda.synchronize()
        ddt = max_over_ranks(time.perf_counter() - t0)
        eb5 = min(1e-3, 1e-3 * float(x5.max() - x5.min()))
        err5 = max_over_ranks(float((rec5 - x5).abs().max()) / eb5)
        total = 5500 * 250000 * 8
        out["config5_s3d_500x500x500x11_f64_absrel"] = {"compress_GBps": total / dt / 1e9, "decompress_GBps": total / ddt / 1e9, "ms": dt * 1e3,
                                                        "slabs": world, "ratio": total / float(sum(sizes5)), "max_err_over_bound": err5,
                                                        "scaling": "strong (one 11 GB array, one slab per rank, streams gathered over NCCL inside the timer)"}
        del x5, buf5, rec5
    sz_b200.init(dict(errorBoundMode="REL", relBoundRatio=REL_BOUND, absErrBound=1e-4))
    torch.cuda.empty_cache()
    return out


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
