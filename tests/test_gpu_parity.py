"""GPU parity tests (run on the B200 box: python -m pytest tests -m gpu).  Every call goes through the C-ABI of
libSZ_b200.so (SZ_Init / SZ_compress_args / SZ_decompress and the szb200_* device-pointer variants); the oracle and the
golden digests are only the checkers.  Bar: bit-exact streams and bit-exact reconstructions."""
import hashlib
import json
import os
import sys

import numpy as np
import pytest

import cases
import szref

pytestmark = pytest.mark.gpu

GOLD = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "streams.json")))


@pytest.fixture(scope="module")
def sz():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import sz_b200
    assert os.path.exists(sz_b200.LIB_PATH), "libSZ_b200.so missing: the CUDA extension must be built in-tree (no fallback)"
    sz_b200.set_device(0)
    return sz_b200


@pytest.fixture(scope="module")
def orc():
    return szref.Oracle()


def _norm(stream, kind):
    return stream if kind != "constant" else stream[:19] + b"\0" + stream[20:]


@pytest.mark.parametrize("case", cases.CASES, ids=cases.CASE_IDS)
def test_stream_and_reconstruction_match_reference_digests(sz, case):
    name, g, mode, ab, rl, extra = case
    a = g()
    gold = GOLD[name]
    sz.init(cases.make_cfg(mode, ab, rl, extra))
    s = sz.compress(a, mode, ab, rl)
    assert len(s) == gold["size"]
    assert hashlib.md5(_norm(s, gold["kind"])).hexdigest() == gold["md5"], "GPU stream differs from the reference stream"
    rec = sz.decompress(s, a.dtype, a.shape)
    if gold["dec_md5"] is not None:
        assert hashlib.md5(rec.tobytes()).hexdigest() == gold["dec_md5"], "GPU reconstruction differs from the reference's"
    else:
        assert np.array_equal(rec, a)
    sz.finalize()


@pytest.mark.parametrize("case", cases.CASES[::4], ids=cases.CASE_IDS[::4])
def test_cross_decode_with_oracle(sz, orc, case):
    """oracle decodes the GPU stream; GPU decodes the oracle stream; identical bits; every point within the bound."""
    name, g, mode, ab, rl, extra = case
    a = g()
    cfg = cases.make_cfg(mode, ab, rl, extra)
    sz.init(cfg)
    orc.init(cfg)
    s_gpu = sz.compress(a, mode, ab, rl)
    s_orc, tr = orc.compress(a, mode, ab, rl, trace=True)
    d1 = orc.decompress(s_gpu, a.dtype, a.shape)
    d2 = sz.decompress(s_orc, a.dtype, a.shape)
    assert d1.tobytes() == d2.tobytes()
    if tr["kind"] == 0:
        assert np.max(np.abs(d2.astype(np.float64) - a.astype(np.float64))) <= tr["eb"]
    sz.finalize()


@pytest.mark.skipif(not szref.have_ref(), reason="compiled reference (oracle/_ref) not present")
@pytest.mark.parametrize("case", cases.CASES[1::5], ids=cases.CASE_IDS[1::5])
def test_live_reference_decodes_gpu_stream(sz, case):
    name, g, mode, ab, rl, extra = case
    a = g()
    if a.size <= 20:
        pytest.skip("the reference cannot decode its own headerless N<=20 streams")
    cfg = cases.make_cfg(mode, ab, rl, extra)
    ref = szref.SZLib(szref.REF_SO)
    ref.init(cfg)
    sz.init(cfg)
    s_gpu = sz.compress(a, mode, ab, rl)
    s_ref = ref.compress(a, mode, ab, rl)
    assert _norm(s_gpu, GOLD[name]["kind"]) == _norm(s_ref, GOLD[name]["kind"])
    d_ref = ref.decompress(s_gpu, a.dtype, a.shape)
    d_gpu = sz.decompress(s_ref, a.dtype, a.shape)
    assert d_ref.tobytes() == d_gpu.tobytes()
    ref.finalize()
    sz.finalize()


BIG = [
    # BASELINE.json config shapes, at sizes the oracle finishes in seconds
    ("cfg2_nyx_256c_rel", lambda: szref.smooth_field((256, 256, 256), np.float32, span=4 * np.pi), szref.REL, 1e-4, 1e-3),
    ("cfg3_isabel_100x500x500_abs", lambda: szref.smooth_field((100, 500, 500), np.float32, span=8 * np.pi, field_id=3), szref.ABS, 1e-2, 1e-3),
    ("cfg4_cesm_1800x3600_rel", lambda: szref.smooth_field((1800, 3600), np.float32, span=8 * np.pi, field_id=5), szref.REL, 1e-4, 1e-4),
    ("cfg5_s3d_4d_f64_absrel", lambda: szref.smooth_field((11, 40, 50, 60), np.float64, span=2 * np.pi), szref.ABS_AND_REL, 1e-3, 1e-3),
    ("plane_120c_regression_heavy", lambda: szref.plane_field((120, 120, 120), np.float32), szref.ABS, 1e-1, 1e-3),
    ("background_150c_use_mean", lambda: szref.background_field((150, 150, 150), np.float32), szref.REL, 1e-4, 1e-3),
    # partial pencils / chunks in every direction, fp64 in 3D, fp32 in 4D, a long thin 2D field
    ("odd_f64_131x77x203_rel", lambda: szref.smooth_field((131, 77, 203), np.float64, span=6 * np.pi), szref.REL, 1e-4, 1e-4),
    ("odd_f32_4d_7x33x41x59_abs", lambda: szref.smooth_field((7, 33, 41, 59), np.float32, span=4 * np.pi), szref.ABS, 1e-3, 1e-3),
    ("odd_f64_2d_611x1031_rel", lambda: szref.smooth_field((611, 1031), np.float64, span=10 * np.pi), szref.REL, 1e-4, 1e-5),
    ("thin_f32_2d_5000x37_abs", lambda: szref.smooth_field((5000, 37), np.float32, span=40 * np.pi), szref.ABS, 1e-3, 1e-3),
    # coefficient Huffman trees with 30-40 thousand leaves each (beyond the shared-memory heap)
    ("blocky_240c_f32_huge_coefficient_trees", lambda: szref.blocky_field(240, np.float32), szref.ABS, 1e-2, 1e-3),
    ("blocky_210c_f64_large_coefficient_trees", lambda: szref.blocky_field(210, np.float64, seed=11), szref.ABS, 1e-2, 1e-3),
    ("noisy_f32_90x100x110_rel", lambda: (szref.smooth_field((90, 100, 110), np.float32, span=4 * np.pi)
                                          + 0.02 * szref.random_field((90, 100, 110), np.float32)).astype(np.float32), szref.REL, 1e-4, 1e-3),
]


@pytest.mark.parametrize("case", BIG, ids=[c[0] for c in BIG])
def test_baseline_shapes_bit_exact_vs_oracle(sz, orc, case):
    name, g, mode, ab, rl = case
    a = g()
    cfg = cases.make_cfg(mode, ab, rl, {})
    sz.init(cfg)
    orc.init(cfg)
    s_gpu = sz.compress(a, mode, ab, rl)
    s_orc = orc.compress(a, mode, ab, rl)
    assert len(s_gpu) == len(s_orc)
    assert s_gpu == s_orc
    d_gpu = sz.decompress(s_gpu, a.dtype, a.shape)
    d_orc = orc.decompress(s_orc, a.dtype, a.shape)
    assert d_gpu.tobytes() == d_orc.tobytes()
    sz.finalize()


def _fuzz_case(seed):
    """seeded odd-shaped field: smooth + a noisy slab + a constant slab + a few spikes (regression, Lorenzo, mean-free and
    unpredictable points all occur), random dtype / rank / mode / bound"""
    rng = np.random.RandomState(1000 + seed)
    rank = [2, 3, 3, 3, 4][seed % 5]
    if rank == 2:
        shape = (int(rng.randint(17, 300)), int(rng.randint(17, 300)))
    elif rank == 3:
        shape = tuple(int(v) for v in rng.randint(6, 70, size=3))
    else:
        shape = tuple(int(v) for v in rng.randint(3, 24, size=4))
    dt = np.float32 if seed % 2 == 0 else np.float64
    a = szref.smooth_field(shape, dt, span=float(rng.uniform(1.0, 12.0)), field_id=seed).astype(np.float64)
    flat = a.reshape(shape[0], -1)
    k = shape[0] // 3
    flat[k:k + max(1, shape[0] // 5)] += 0.05 * rng.standard_normal(flat[k:k + max(1, shape[0] // 5)].shape)
    flat[-max(1, shape[0] // 6):] = 0.75
    idx = rng.randint(0, a.size, size=max(1, a.size // 500))
    a.reshape(-1)[idx] += rng.uniform(-50, 50, size=idx.size)
    mode = [szref.REL, szref.ABS, szref.ABS_AND_REL, szref.ABS_OR_REL][seed % 4]
    rel = float(10.0 ** rng.uniform(-5, -2))
    ab = float(10.0 ** rng.uniform(-4, -1))
    return np.ascontiguousarray(a.astype(dt)), mode, ab, rel


@pytest.mark.parametrize("seed", range(24))
def test_fuzzed_shapes_bit_exact_vs_oracle(sz, orc, seed):
    a, mode, ab, rl = _fuzz_case(seed)
    cfg = cases.make_cfg(mode, ab, rl, {})
    sz.init(cfg)
    orc.init(cfg)
    s_gpu = sz.compress(a, mode, ab, rl)
    s_orc = orc.compress(a, mode, ab, rl)
    assert s_gpu == s_orc, "stream differs for shape %s dtype %s mode %d" % (a.shape, a.dtype, mode)
    d_gpu = sz.decompress(s_gpu, a.dtype, a.shape)
    d_orc = orc.decompress(s_orc, a.dtype, a.shape)
    assert d_gpu.tobytes() == d_orc.tobytes()
    sz.finalize()


def test_batch_of_fields_equals_individual_calls(sz):
    """BASELINE configs 3 / 4 in miniature: a batch of equally shaped fields on concurrent worker lanes gives, item by item,
    the stream of the single call (which the other tests pin to the reference), and decodes back bit-exactly."""
    import torch
    for shape, dt, mode, ab, rl in (((40, 60, 50), np.float32, szref.ABS, 1e-2, 1e-3), ((200, 360), np.float32, szref.REL, 1e-4, 1e-4),
                                    ((24, 30, 36), np.float64, szref.ABS_AND_REL, 1e-3, 1e-3)):
        sz.init(cases.make_cfg(mode, ab, rl, {}))
        fields = [szref.smooth_field(shape, dt, span=6 * np.pi, field_id=k) for k in range(7)]
        fields[3] = szref.background_field(shape, dt) if len(shape) == 3 else fields[3]          # a use_mean field inside the batch
        singles = [sz.compress(a, mode, ab, rl) for a in fields]
        xs = [torch.from_numpy(a).cuda() for a in fields]
        cap = sz.compress_bound(dt, shape)
        outs = [torch.empty(cap, dtype=torch.uint8, device="cuda") for _ in fields]
        for lanes in (1, 3):
            sizes = sz.compress_batch_device(xs, outs, mode, ab, rl, lanes=lanes)
            for k, (n, o) in enumerate(zip(sizes, outs)):
                assert bytes(o[:n].cpu().numpy()) == singles[k], "batch item %d differs (lanes=%d)" % (k, lanes)
            recs = [torch.empty_like(x) for x in xs]
            sz.decompress_batch_device(outs, sizes, recs, lanes=lanes)
            for k, r in enumerate(recs):
                assert r.cpu().numpy().tobytes() == sz.decompress(singles[k], dt, shape).tobytes()
        sz.finalize()


def test_batch_reports_a_bad_item_without_spoiling_the_others(sz):
    """one truncated stream inside a batch: its item fails loudly (negative rc), the other items still decode bit-exactly"""
    import ctypes as C
    import torch
    from sz_b200.api import BatchItem
    shape, dt, mode, ab, rl = (40, 44, 48), np.float32, szref.ABS, 1e-3, 1e-3
    sz.init(cases.make_cfg(mode, ab, rl, {}))
    fields = [szref.smooth_field(shape, dt, span=5 * np.pi, field_id=k) for k in range(4)]
    streams = [sz.compress(a, mode, ab, rl) for a in fields]
    dev = [torch.from_numpy(np.frombuffer(s, dtype=np.uint8).copy()).cuda() for s in streams]
    outs = [torch.empty(shape, dtype=torch.float32, device="cuda") for _ in fields]
    items = (BatchItem * 4)()
    for i in range(4):
        items[i].d_stream, items[i].stream_capacity, items[i].d_out = dev[i].data_ptr(), len(streams[i]), outs[i].data_ptr()
    items[2].stream_capacity = len(streams[2]) // 3            # truncated
    L = sz.lib()
    L.szb200_decompress_batch_device.argtypes = [C.c_int, C.POINTER(BatchItem), C.c_int, C.c_int] + [C.c_size_t] * 5
    L.szb200_decompress_batch_device.restype = C.c_int
    rc = L.szb200_decompress_batch_device(sz.SZ_FLOAT, items, 4, 2, 0, 0, *shape)
    assert rc != 0 and items[2].rc != 0
    for i in (0, 1, 3):
        assert items[i].rc == 0
        assert outs[i].cpu().numpy().tobytes() == sz.decompress(streams[i], dt, shape).tobytes()
    sz.finalize()


_DROPIN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "dropin")


@pytest.mark.skipif(not os.path.exists(os.path.join(_DROPIN, "testfloat_compress")), reason="relinked reference examples (make -C oracle dropin) not present")
@pytest.mark.parametrize("kind", ["float", "double"])
def test_reference_example_programs_run_unmodified_on_the_library(sz, tmp_path, kind):
    """the reference's own example/test{float,double}_{compress,decompress}.c, compiled unmodified and linked
    against libSZ_b200.so instead of libSZ, write the reference's .sz file and decode it back (digests of the reference)."""
    import hashlib
    import shutil
    import subprocess
    dat = "test%s_8_8_128.dat" % kind
    shutil.copy(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", dat), tmp_path / dat)
    szref.write_config(str(tmp_path / "sz.config"), cases.make_cfg(szref.ABS, 1e-4, 1e-3, {}))
    gold = GOLD["test%s_8_8_128_abs" % kind] if kind == "float" else GOLD["cfg1_testdouble_8_8_128_abs"]
    r = subprocess.run([os.path.join(_DROPIN, "test%s_compress" % kind), "sz.config", dat, "8", "8", "128"], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "done" in r.stdout, r.stdout + r.stderr
    stream = (tmp_path / (dat + ".sz")).read_bytes()
    assert len(stream) == gold["size"] and hashlib.md5(stream).hexdigest() == gold["md5"]
    r = subprocess.run([os.path.join(_DROPIN, "test%s_decompress" % kind), dat + ".sz", "8", "8", "128"], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    out = (tmp_path / (dat + ".sz.out")).read_bytes()
    assert hashlib.md5(out).hexdigest() == gold["dec_md5"], r.stdout


@pytest.mark.skipif(not os.path.exists(os.path.join(_DROPIN, "sz")), reason="relinked reference command-line tool (make -C oracle dropin) not present")
@pytest.mark.parametrize("kind", ["float", "double"])
def test_reference_command_line_tool_runs_unmodified_on_the_library(sz, tmp_path, kind):
    """the reference's command-line tool (example/sz.c), compiled unmodified and linked against libSZ_b200.so:
    `sz -z` writes the reference's stream, `sz -x` the reference's reconstruction, `-a` reports the bound as held, `-p` prints
    the stream's metadata."""
    import hashlib
    import shutil
    import subprocess
    exe = os.path.join(_DROPIN, "sz")
    flag = "-f" if kind == "float" else "-d"
    dat = "test%s_8_8_128.dat" % kind
    shutil.copy(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", dat), tmp_path / dat)
    szref.write_config(str(tmp_path / "sz.config"), cases.make_cfg(szref.ABS, 1e-4, 1e-3, {}))
    gold = GOLD["test%s_8_8_128_abs" % kind] if kind == "float" else GOLD["cfg1_testdouble_8_8_128_abs"]
    r = subprocess.run([exe, "-z", flag, "-c", "sz.config", "-i", dat, "-3", "8", "8", "128"], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "compressed data file" in r.stdout, r.stdout + r.stderr
    stream = (tmp_path / (dat + ".sz")).read_bytes()
    assert len(stream) == gold["size"] and hashlib.md5(stream).hexdigest() == gold["md5"]
    r = subprocess.run([exe, "-x", flag, "-s", dat + ".sz", "-i", dat, "-3", "8", "8", "128", "-a"], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "decompressed data file" in r.stdout, r.stdout + r.stderr
    out = (tmp_path / (dat + ".sz.out")).read_bytes()
    assert hashlib.md5(out).hexdigest() == gold["dec_md5"], r.stdout
    line = [l for l in r.stdout.splitlines() if l.startswith("Max absolute error")]
    assert line and float(line[0].split("=")[1]) <= 1e-4, r.stdout
    r = subprocess.run([exe, "-p", "-s", dat + ".sz"], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "ABS" in r.stdout, r.stdout + r.stderr
    # the -M / -A overrides of the tool write into confparams_cpr directly: the library must honour them
    r = subprocess.run([exe, "-z", flag, "-c", "sz.config", "-i", dat, "-M", "REL", "-R", "1E-3", "-3", "8", "8", "128"], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    a = np.fromfile(tmp_path / dat, dtype=np.float32 if kind == "float" else np.float64).reshape(128, 8, 8)
    sz.init(dict(errorBoundMode="REL", relBoundRatio=1e-3))
    want = sz.compress(a, sz.REL, 1e-4, 1e-3)
    sz.finalize()
    assert (tmp_path / (dat + ".sz")).read_bytes() == bytes(want)


def test_device_pointer_api_roundtrip_and_determinism(sz):
    import torch
    a = szref.smooth_field((200, 210, 220), np.float32, span=6 * np.pi)
    sz.init(dict(errorBoundMode="REL", relBoundRatio=1e-3))
    s_host = sz.compress(a, sz.REL, 1e-4, 1e-3)
    x = torch.from_numpy(a).cuda()
    buf = torch.empty(sz.compress_bound(np.float32, a.shape), dtype=torch.uint8, device="cuda")
    n1 = sz.compress_device(x, buf, sz.REL, 1e-4, 1e-3)
    s_dev = bytes(buf[:n1].cpu().numpy().tobytes())
    assert s_dev == s_host, "device-pointer and host-pointer entry points must produce the same stream"
    n2 = sz.compress_device(x, buf, sz.REL, 1e-4, 1e-3)
    assert n2 == n1 and bytes(buf[:n2].cpu().numpy().tobytes()) == s_host, "compression must be deterministic"
    out = torch.empty_like(x)
    sz.decompress_device(buf, n1, out)
    rec = sz.decompress(s_host, np.float32, a.shape)
    assert out.cpu().numpy().tobytes() == rec.tobytes()
    eb = np.float32(1e-3 * float(np.float32(a.max()) - np.float32(a.min())))
    assert float(np.max(np.abs(rec.astype(np.float64) - a.astype(np.float64)))) <= float(eb)
    sz.finalize()


FULL = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "full_size.json")))


def test_full_size_nyx_512_stream_equals_reference(sz):
    """BASELINE.json configs[1] at full size (512^3 fp32, REL 1e-3, the field bench.py times): the stream and the reconstruction
    equal the reference's, byte for byte -- against the digests tests/golden/make_golden.py --full took from the compiled
    reference when this box generates the very same input bytes, against the compiled reference run here otherwise."""
    import torch
    import bench
    a = bench.make_field((512, 512, 512))
    gold = FULL["cfg2_nyx_512_rel"]
    sz.init(dict(errorBoundMode="REL", relBoundRatio=1e-3))
    x = torch.from_numpy(a).cuda()
    buf = torch.empty(sz.compress_bound(np.float32, a.shape), dtype=torch.uint8, device="cuda")
    n = sz.compress_device(x, buf, sz.REL, 1e-4, 1e-3)
    st = sz.last_stats()
    assert st.num_blocks == 85 ** 3 and st.intervals >= 32
    s_gpu = buf[:n].cpu().numpy().tobytes()
    out = torch.empty_like(x)
    sz.decompress_device(buf, n, out)
    dec_md5 = hashlib.md5(out.cpu().numpy().tobytes()).hexdigest()
    if hashlib.md5(a.tobytes()).hexdigest() == gold["input_md5"]:
        assert n == gold["size"] and hashlib.md5(s_gpu).hexdigest() == gold["md5"], "GPU stream differs from the reference stream at 512^3"
        assert dec_md5 == gold["dec_md5"], "GPU reconstruction differs from the reference's at 512^3"
    else:
        assert szref.have_ref(), "this box's numpy makes a different field and the compiled reference is not here to compare with"
        ref = szref.SZLib(szref.REF_SO)
        ref.init(dict(errorBoundMode="REL", relBoundRatio=1e-3))
        s_ref = ref.compress(a, szref.REL, 1e-4, 1e-3)
        assert s_gpu == s_ref
        assert dec_md5 == hashlib.md5(ref.decompress(s_ref, a.dtype, a.shape).tobytes()).hexdigest()
        ref.finalize()
    rng = float((x.max() - x.min()).item())
    eb = float(np.float32(1e-3 * np.float64(np.float32(rng))))
    assert float((out.double() - x.double()).abs().max().item()) <= eb
    n2 = sz.compress_device(x, buf, sz.REL, 1e-4, 1e-3)          # deterministic
    assert n2 == n and buf[:n2].cpu().numpy().tobytes() == s_gpu
    sz.finalize()


def _cfg5_slab(r, lo, hi, dev):
    """rows [lo, hi) of the flattened (5500, 500, 500) fp64 S3D-shape field of BASELINE.json configs[4] (4D: 11 x 500 x 500 x 500;
    the reference flattens r4 * r3 itself, sz_float.c:3010), generated on the device"""
    import torch
    g = [torch.linspace(0, 8 * np.pi, 500, dtype=torch.float64, device=dev) for _ in range(3)]
    rows = torch.arange(lo, hi, device=dev)
    s4, z = (rows // 500).double()[:, None, None], g[0][rows % 500][:, None, None]
    y, xx = g[1][None, :, None], g[2][None, None, :]
    x = torch.sin(xx) * torch.cos(1.3 * y) * torch.sin(0.7 * z + 0.3 + 0.37 * s4) + 0.25 * torch.sin(3.1 * xx + 1.7 * y) * torch.cos(2.3 * z)
    x += 1e-3 * torch.randn(x.shape, dtype=torch.float64, device=dev, generator=torch.Generator(device=dev).manual_seed(4321 + r))
    return x.contiguous()


@pytest.mark.slow
@pytest.mark.skipif(not szref.have_ref(), reason="compiled reference (oracle/_ref) not present")
def test_config5_full_size_slabs_equal_reference(sz, tmp_path):
    """BASELINE.json configs[4] at its own size and dtype: the 11 GB fp64 ABS_AND_REL array cut into 8 slabs along the slowest
    dimension (sz_b200.slabs.slab_bounds, the reference's own pattern: example/test_threadsafe.c:40-58).  Every slab's GPU stream
    equals the stream of the compiled reference called on that slab (one host process per slab), the GPU reconstructs every slab
    within the bound, and the reference's reconstruction of one slab equals the GPU's bit for bit.  First evidence at this scale
    for the fp64 coefficient trees (tens of thousands of leaves) and a 688-row slow axis."""
    import concurrent.futures as cf
    import multiprocessing as mp
    import shutil
    import torch
    from sz_b200 import slabs
    G = 8
    free, _ = torch.cuda.mem_get_info()
    if free < 12e9:
        pytest.skip("needs 12 GB of device memory")
    shm = "/dev/shm" if os.path.isdir("/dev/shm") and shutil.disk_usage("/dev/shm").free > 13e9 else str(tmp_path)
    if shutil.disk_usage(shm).free < 13e9:
        pytest.skip("needs 13 GB of scratch space for the slabs handed to the reference processes")
    cfg = dict(errorBoundMode="ABS_AND_REL", absErrBound=1e-3, relBoundRatio=1e-3)
    sz.init(cfg)
    bounds = slabs.slab_bounds(5500, G)
    paths, futs, gpu = [], [], []
    workers = max(1, min(G, (os.cpu_count() or 8)))
    with cf.ProcessPoolExecutor(max_workers=workers, mp_context=mp.get_context("spawn")) as pool:
        for r, (lo, hi) in enumerate(bounds):
            x = _cfg5_slab(r, lo, hi, "cuda")
            shape = (hi - lo, 500, 500)
            path = os.path.join(shm, "szb200_cfg5_slab%d.npy" % r)
            np.save(path, x.cpu().numpy())
            paths.append(path)
            futs.append(pool.submit(szref.ref_compress_file, path, np.float64, shape, cfg, szref.ABS_AND_REL, 1e-3, 1e-3, r == 0))
            buf = torch.empty(sz.compress_bound(np.float64, shape), dtype=torch.uint8, device="cuda")
            n = sz.compress_device(x, buf, sz.ABS_AND_REL, 1e-3, 1e-3)
            out = torch.empty_like(x)
            sz.decompress_device(buf, n, out)
            eb = min(1e-3, 1e-3 * float(x.max() - x.min()))
            assert float((out - x).abs().max()) <= eb
            gpu.append((n, hashlib.md5(buf[:n].cpu().numpy().tobytes()).hexdigest(),
                        hashlib.md5(out.cpu().numpy().tobytes()).hexdigest() if r == 0 else None))
            del x, buf, out
        try:
            for r, f in enumerate(futs):
                size, md5, dec, _ = f.result(timeout=600)
                assert gpu[r][0] == size and gpu[r][1] == md5, "slab %d: GPU stream differs from the reference stream" % r
                if dec is not None:
                    assert gpu[r][2] == dec, "slab %d: GPU reconstruction differs from the reference's" % r
        finally:
            for p in paths:
                if os.path.exists(p):
                    os.remove(p)
    sz.finalize()


def test_thread_safe_calls_run_side_by_side_and_equal_the_reference(sz):
    """The reference's slab-parallel pattern (example/test_threadsafe.c:40-58): five threads, each SZ_compress_customize_threadsafe
    on its slab of one array with its own REL bound, then SZ_decompress_customize_threadsafe.  Here every such call has a worker
    context of its own, so the threads overlap on the GPU; every stream must equal the reference's SZ_compress_args on that slab."""
    import ctypes as C
    import threading
    L = sz.lib()
    sz.init(dict(errorBoundMode="ABS", absErrBound=1e-3))
    full = szref.smooth_field((100, 120, 140), np.float32, span=6 * np.pi)
    nthr, rows = 5, 20

    cpr = C.c_void_p.in_dll(L, "confparams_cpr")
    results = [None] * nthr
    stsz = C.c_size_t
    L.SZ_compress_customize_threadsafe.restype = C.c_void_p
    L.SZ_compress_customize_threadsafe.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_void_p] + [stsz] * 5 + [C.POINTER(stsz), C.POINTER(C.c_int)]
    L.SZ_decompress_customize_threadsafe.restype = C.c_void_p
    L.SZ_decompress_customize_threadsafe.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_void_p, stsz] + [stsz] * 5 + [C.POINTER(C.c_int)]

    def worker(i):
        slab = np.ascontiguousarray(full[i * rows:(i + 1) * rows])
        para = sz.SzParams()                                     # memcpy(&params, confparams_cpr, sizeof(sz_params)), as the example does
        C.memmove(C.byref(para), cpr.value, C.sizeof(para))
        para.errorBoundMode, para.relBoundRatio = sz.REL, 1e-4 * (i + 1)
        n, status = stsz(0), C.c_int(0)
        p = L.SZ_compress_customize_threadsafe(b"SZ", C.byref(para), 0, slab.ctypes.data, 0, 0, rows, 120, 140, C.byref(n), C.byref(status))
        stream = C.string_at(p, n.value) if p else None
        q = L.SZ_decompress_customize_threadsafe(b"SZ", C.byref(para), 0, p, n.value, 0, 0, rows, 120, 140, C.byref(status)) if p else None
        rec = np.frombuffer(C.string_at(q, slab.nbytes), dtype=np.float32).reshape(slab.shape).copy() if q else None
        results[i] = (stream, rec, slab)

    for rep in range(2):
        th = [threading.Thread(target=worker, args=(i,)) for i in range(nthr)]
        for t in th:
            t.start()
        for t in th:
            t.join()
        orc = szref.Oracle()
        for i in range(nthr):
            stream, rec, slab = results[i]
            assert stream is not None and rec is not None
            orc.init(dict(errorBoundMode="ABS", absErrBound=1e-3))      # the global configuration (the header's rel field comes from it, ByteToolkit.c:924-929)
            want = orc.compress(slab, szref.REL, 1e-3, 1e-4 * (i + 1))     # mode and bound from the caller's block
            assert stream == want, "thread %d: stream differs from the oracle's on its slab" % i
            assert rec.tobytes() == orc.decompress(want, np.float32, slab.shape).tobytes()
    sz.finalize()


def test_slab_entry_point_of_the_c_boundary(sz):
    """szb200_compress_slabs / szb200_decompress_slabs: one host array, eight slabs, worker threads over the visible devices (all
    on one device here).  Every slab stream in the container equals the oracle's stream for that slab; the array comes back."""
    from sz_b200 import slabs
    for dtype, shape, mode, ab, rl in ((np.float32, (130, 64, 80), szref.REL, 1e-4, 1e-3), (np.float64, (11, 12, 20, 24), szref.ABS_AND_REL, 1e-3, 1e-3),
                                       (np.float32, (200, 333), szref.ABS, 1e-3, 1e-4)):
        a = szref.smooth_field(shape, dtype, span=4 * np.pi)
        cfg = cases.make_cfg(mode, ab, rl, {})
        sz.init(cfg)
        box = sz.compress_slabs(a, mode, ab, rl, n_slabs=8)
        parts = slabs.unframe(box)
        n0 = shape[0] if len(shape) < 4 else shape[0] * shape[1]
        flat = a.reshape((n0,) + shape[-2:]) if len(shape) >= 3 else a
        bounds = slabs.slab_bounds(n0, len(parts))
        assert len(parts) == min(8, max(1, n0 // 12))
        orc = szref.Oracle()
        orc.init(cfg)
        for (lo, hi), s in zip(bounds, parts):
            assert s == orc.compress(np.ascontiguousarray(flat[lo:hi]), mode, ab, rl)
        back = sz.decompress_slabs(box, dtype, shape)
        for (lo, hi), s in zip(bounds, parts):
            assert back.reshape(flat.shape)[lo:hi].tobytes() == orc.decompress(s, dtype, flat[lo:hi].shape).tobytes()
        sz.finalize()


@pytest.mark.skipif(not os.path.exists(os.path.join(_DROPIN, "test_threadsafe")), reason="relinked example/test_threadsafe.c (make -C oracle dropin) not present")
def test_reference_thread_safe_example_runs_unmodified_on_the_library(sz, tmp_path):
    """example/test_threadsafe.c, compiled unmodified with OpenMP and linked against libSZ_b200.so: five threads, five slabs of a
    100 x 500 x 500 field, REL 1e-4 .. 5e-4 per slab; every slab's maximum relative error stays within its bound.  (The example
    never sets szMode and the built-in default is the zstd mode, which this library refuses: SZB200_SZMODE overrides it.)"""
    import re
    import subprocess
    a = szref.smooth_field((100, 500, 500), np.float32, span=8 * np.pi)
    a.tofile(tmp_path / "CLOUDf48.bin.f32")
    r = subprocess.run([os.path.join(_DROPIN, "test_threadsafe")], cwd=tmp_path, capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, OMP_NUM_THREADS="5", SZB200_SZMODE="SZ_BEST_SPEED"))     # the example takes SZ_Init(NULL)'s mode, the zstd one
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    rel = [float(x) for x in re.findall(r"Max relative error = ([0-9.eE+-]+)", r.stdout)]
    assert len(rel) == 5 and max(rel) <= 5e-4 * 1.0001 and min(rel) > 0, r.stdout[-2000:]
    assert re.search(r"max_error=", r.stdout)


@pytest.mark.parametrize("form", ["default", "exact", "narrow_band"])
def test_coefficient_chain_forms_give_the_reference_stream(form):
    """The coefficient chain (szb_chain.cuh) guesses codes and proves them; its fall-backs must give the same bytes: the default band,
    every step with the reference's complete formula (SZB200_CHAIN=exact), and a band so narrow that the guesses near midpoints are
    wrong and windows are rerun (SZB200_LAT_BAND=0.0005: the proof and both rerun levels run).  Fresh process per form: the
    library reads the variables once."""
    import re
    import subprocess
    env = dict(os.environ, SZB200_TREE_PROFILE="1")
    if form == "exact":
        env["SZB200_CHAIN"] = "exact"
    if form == "narrow_band":
        env["SZB200_LAT_BAND"] = "0.0005"
    r = subprocess.run([sys.executable, os.path.join(os.path.dirname(os.path.abspath(__file__)), "chain_modes_check.py")],
                       capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0 and "chain forms check: ok" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
    reruns = [int(x) for x in re.findall(r"reruns: (\d+)", r.stderr)]
    assert reruns, r.stderr[-2000:]
    if form == "narrow_band":
        assert max(reruns) > 0, "the narrow band was meant to make some window fail its proof: " + r.stderr[-2000:]


@pytest.mark.parametrize("kind", ["bell_1024", "two_symbols", "one_symbol", "wide_65536", "skewed_64"])
def test_huffman_container_encode_with_tree_equals_the_reference(sz, orc, kind):
    """encode_withTree / decode_withTree (Huffman.c:790-816, :865-889), the [nodeCount][stateNum/2][tree][bits]
    container of the SZ 1.4 storage, stand-alone through the C-ABI: byte-equal to the compiled reference's encode_withTree where it
    is present, to the container put together from the oracle's tree and bit stream otherwise; and back."""
    import ctypes as C
    import struct
    import torch
    from sz_b200 import api
    rng = np.random.default_rng(11)
    n = 200000
    if kind == "bell_1024":
        state, s = 1024, np.clip(np.rint(rng.normal(256, 40, n)), 0, 511).astype(np.int32)
    elif kind == "two_symbols":
        state, s = 64, rng.integers(7, 9, n).astype(np.int32)
    elif kind == "one_symbol":
        state, s = 64, np.full(n, 5, dtype=np.int32)
    elif kind == "wide_65536":
        state, s = 65536, np.clip(np.rint(rng.normal(16384, 3000, n)), 0, 32767).astype(np.int32)
    else:
        state, s = 128, np.minimum(rng.geometric(0.3, n), 63).astype(np.int32)
    tree, nodes, code, clen = orc.huffman_build(s, state)
    want = struct.pack(">II", nodes, state // 2) + tree + orc.huffman_encode(s, code, clen)
    if szref.have_ref():                                        # the reference itself
        R = C.CDLL(szref.REF_SO)
        R.createHuffmanTree.restype = C.c_void_p
        R.createHuffmanTree.argtypes = [C.c_int]
        R.encode_withTree.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]
        R.SZ_ReleaseHuffman.argtypes = [C.c_void_p]
        t = R.createHuffmanTree(state)
        out, osz = C.c_void_p(), C.c_size_t(0)
        R.encode_withTree(t, s.ctypes.data, s.size, C.byref(out), C.byref(osz))
        ref = C.string_at(out, osz.value)
        R.SZ_ReleaseHuffman(t)
        assert ref == want, "the oracle's container differs from the reference's"
    d_sym = torch.from_numpy(s.astype(np.int16)).cuda()
    got = api.encode_with_tree(d_sym, state)
    assert got == want
    back = api.decode_with_tree(got, n)
    assert torch.equal(back, d_sym)


def test_incompressible_field_fits_a_raw_sized_device_buffer(sz):
    """white noise with a tight bound: the reference stores the original data (lossless container, sz_float.c:526-559, :2979).  A
    device buffer of just that size must be enough -- the Huffman body that does not fit it is the very trigger of that container --
    and the stream equals the host-pointer call's."""
    import torch
    a = szref.random_field((40, 36, 44), np.float32)
    sz.init(dict(errorBoundMode="ABS", absErrBound=1e-7))
    s_host = sz.compress(a, sz.ABS, 1e-7, 0)
    assert s_host[3] == 80 and len(s_host) == a.nbytes + 3 + 28 + 8 + 1
    x = torch.from_numpy(a).cuda()
    buf = torch.empty(len(s_host), dtype=torch.uint8, device="cuda")
    n = sz.compress_device(x, buf, sz.ABS, 1e-7, 0)
    assert n == len(s_host) and buf[:n].cpu().numpy().tobytes() == s_host
    sz.finalize()


def test_corrupt_section_headers_are_refused_without_damage(sz):
    """every count and size a stream carries is checked before it sizes a copy: a stream with a damaged section header is refused
    (or, if the damage happens to be harmless, decoded), and the next valid call still gives the right answer"""
    a = szref.plane_field((30, 26, 40), np.float32)              # regression blocks: coefficient sections present
    sz.init(dict(errorBoundMode="ABS", absErrBound=1e-1))
    s = sz.compress(a, sz.ABS, 1e-1, 0)
    good = sz.decompress(s, a.dtype, a.shape)
    rng = np.random.default_rng(5)
    head = 4 + 28 + 8                                            # version + flags, parameter block, element count
    refused = 0
    for trial in range(60):
        b = bytearray(s)
        off = head + int(rng.integers(0, min(len(s) - head, 400 + 40 * trial)))
        for q in range(int(rng.integers(1, 5))):
            if off + q < len(b):
                b[off + q] = 0xff if trial % 2 else int(rng.integers(0, 256))
        try:
            sz.decompress(bytes(b), a.dtype, a.shape)
        except sz.SZError:
            refused += 1
    assert refused > 0
    assert sz.decompress(s, a.dtype, a.shape).tobytes() == good.tobytes()
    sz.finalize()


def test_unsupported_requests_fail_loudly(sz):
    a = szref.smooth_field((32, 32, 32), np.float32)
    sz.init(dict(szMode="SZ_BEST_COMPRESSION"))
    with pytest.raises(sz.SZError):
        sz.compress(a, sz.ABS, 1e-3, 0)
    sz.init(dict())
    with pytest.raises(sz.SZError):
        sz.compress(a, 10, 1e-3, 0)             # PW_REL is outside this path
    with pytest.raises(sz.SZError):
        sz.compress(a.reshape(-1), sz.ABS, 1e-3, 0)   # 1D = SZ 1.4 path, not built
    with pytest.raises(sz.SZError):
        sz.decompress(b"\x02\x01\x0c" + b"\0" * 60, np.float32, (32, 32, 32))
    sz.finalize()
