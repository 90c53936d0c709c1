"""Test-side ctypes bindings for (a) the UNMODIFIED reference library compiled from /root/reference
(oracle/_ref/libSZ_ref.so) and (b) our C restatement (oracle/libszoracle.so).

TEST INFRASTRUCTURE ONLY: nothing under sz_b200/ imports this module.
"""
import ctypes as C
import os
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libSZ_ref.so")
REF_OMP_SO = os.path.join(ROOT, "oracle", "_ref", "libSZ_ref_omp.so")
ORACLE_SO = os.path.join(ROOT, "oracle", "libszoracle.so")

SZ_FLOAT, SZ_DOUBLE = 0, 1
ABS, REL, ABS_AND_REL, ABS_OR_REL = 0, 1, 2, 3
PSNR, NORM = 4, 5
MODE_NAMES = {ABS: "ABS", REL: "REL", ABS_AND_REL: "ABS_AND_REL", ABS_OR_REL: "ABS_OR_REL", PSNR: "PSNR", NORM: "NORM"}


def have_ref():
    return os.path.exists(REF_SO)


def have_oracle():
    return os.path.exists(ORACLE_SO)


def dtype_code(a):
    if a.dtype == np.float32:
        return SZ_FLOAT
    if a.dtype == np.float64:
        return SZ_DOUBLE
    raise TypeError(a.dtype)


def dims5(shape):
    """numpy C-order shape -> (r5, r4, r3, r2, r1) with r1 the fastest dimension."""
    r = list(reversed(shape)) + [0] * (5 - len(shape))
    return r[4], r[3], r[2], r[1], r[0]


DEFAULT_CFG = dict(
    max_quant_intervals=65536, quantization_intervals=0, sampleDistance=100, predThreshold=0.99,
    szMode="SZ_BEST_SPEED", protectValueRange="NO", withLinearRegression="YES",
    errorBoundMode="ABS", absErrBound=1e-4, relBoundRatio=1e-4, psnr=80.0, normErr=0.05,
)


def write_config(path, cfg):
    c = dict(DEFAULT_CFG)
    c.update(cfg or {})
    with open(path, "w") as f:
        f.write("[ENV]\ndataEndianType = LITTLE_ENDIAN_DATA\nsol_name = SZ\n[PARAMETER]\n")
        f.write("snapshotCmprStep = 5\n")
        f.write("withLinearRegression = %s\n" % c["withLinearRegression"])
        f.write("protectValueRange = %s\n" % c["protectValueRange"])
        f.write("sampleDistance = %d\n" % c["sampleDistance"])
        f.write("quantization_intervals = %d\n" % c["quantization_intervals"])
        f.write("max_quant_intervals = %d\n" % c["max_quant_intervals"])
        f.write("predThreshold = %s\n" % repr(c["predThreshold"]))
        f.write("szMode = %s\n" % c["szMode"])
        f.write("losslessCompressor = ZSTD_COMPRESSOR\ngzipMode = Gzip_BEST_SPEED\nzstdMode = Zstd_HIGH_SPEED\nrandomAccess = 0\n")
        f.write("errorBoundMode = %s\n" % c["errorBoundMode"])
        f.write("absErrBound = %s\n" % repr(float(c["absErrBound"])))
        f.write("relBoundRatio = %s\n" % repr(float(c["relBoundRatio"])))
        f.write("psnr = %s\nnormErr = %s\npw_relBoundRatio = 1E-2\naccelerate_pw_rel_compression = 1\n" % (repr(float(c["psnr"])), repr(float(c["normErr"]))))
    return c


class SZLib:
    """Binding of the reference's public C API (sz/include/sz.h:253-334). Also used by the product tests
    to drive OUR drop-in library through the very same entry points (pass its path)."""

    def __init__(self, path=REF_SO):
        self.lib = C.CDLL(path)
        L = self.lib
        L.SZ_Init.argtypes = [C.c_char_p]
        L.SZ_Init.restype = C.c_int
        L.SZ_Finalize.argtypes = []
        L.SZ_Finalize.restype = None
        L.SZ_compress_args.argtypes = [C.c_int, C.c_void_p, C.POINTER(C.c_size_t), C.c_int, C.c_double, C.c_double, C.c_double] + [C.c_size_t] * 5
        L.SZ_compress_args.restype = C.c_void_p
        L.SZ_compress.argtypes = [C.c_int, C.c_void_p, C.POINTER(C.c_size_t)] + [C.c_size_t] * 5
        L.SZ_compress.restype = C.c_void_p
        L.SZ_decompress.argtypes = [C.c_int, C.c_void_p, C.c_size_t] + [C.c_size_t] * 5
        L.SZ_decompress.restype = C.c_void_p
        L.SZ_decompress_args.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_void_p] + [C.c_size_t] * 5
        L.SZ_decompress_args.restype = C.c_size_t
        self.libc = C.CDLL(None)
        self.libc.free.argtypes = [C.c_void_p]
        self.cfg = None

    def init(self, cfg=None):
        fd, path = tempfile.mkstemp(suffix=".config")
        os.close(fd)
        try:
            self.cfg = write_config(path, cfg)
            rc = self.lib.SZ_Init(path.encode())
        finally:
            os.unlink(path)
        if rc != 0:
            raise RuntimeError("SZ_Init failed rc=%d" % rc)
        return rc

    def finalize(self):
        self.lib.SZ_Finalize()

    def compress(self, a, mode, abs_eb=0.0, rel=0.0):
        a = np.ascontiguousarray(a)
        n = C.c_size_t(0)
        p = self.lib.SZ_compress_args(dtype_code(a), a.ctypes.data, C.byref(n), mode, abs_eb, rel, 0.0, *dims5(a.shape))
        if not p:
            raise RuntimeError("SZ_compress_args returned NULL")
        out = C.string_at(p, n.value)
        self.libc.free(p)
        return out

    def decompress(self, stream, dtype, shape):
        dt = np.dtype(dtype)
        code = SZ_FLOAT if dt == np.float32 else SZ_DOUBLE
        buf = (C.c_ubyte * len(stream)).from_buffer_copy(stream)
        p = self.lib.SZ_decompress(code, buf, len(stream), *dims5(shape))
        if not p:
            raise RuntimeError("SZ_decompress returned NULL")
        n = int(np.prod(shape))
        out = np.frombuffer(C.string_at(p, n * dt.itemsize), dtype=dt).reshape(shape).copy()
        self.libc.free(p)
        return out


class szo_config(C.Structure):
    _fields_ = [("max_quant_intervals", C.c_uint), ("quantization_intervals", C.c_uint), ("sampleDistance", C.c_int),
                ("predThreshold", C.c_float), ("szMode", C.c_int), ("gzipMode", C.c_int), ("protectValueRange", C.c_int),
                ("sol_ID", C.c_int), ("dataEndianType", C.c_int), ("conf_absErrBound", C.c_double), ("conf_relBoundRatio", C.c_double),
                ("psnr", C.c_double), ("normErr", C.c_double)]


class szo_trace(C.Structure):
    _fields_ = [("dtype", C.c_int), ("dim", C.c_int), ("n0", C.c_size_t), ("n1", C.c_size_t), ("n2", C.c_size_t),
                ("N", C.c_size_t), ("num_blocks", C.c_size_t), ("reg_count", C.c_size_t), ("kind", C.c_int),
                ("min", C.c_double), ("max", C.c_double), ("range", C.c_double), ("eb_d", C.c_double), ("eb", C.c_double),
                ("intervals", C.c_uint), ("use_mean", C.c_int),
                ("mean", C.c_double), ("dense_pos", C.c_double), ("max_freq", C.c_double), ("mean_freq", C.c_double), ("sample_mean", C.c_double),
                ("sample_count", C.c_size_t),
                ("reg_params", C.POINTER(C.c_double)), ("indicator", C.POINTER(C.c_ubyte)), ("coeff_type", C.POINTER(C.c_int)),
                ("coeff_last", C.POINTER(C.c_double)), ("type", C.POINTER(C.c_int)), ("unpred", C.POINTER(C.c_double)),
                ("total_unpred", C.c_size_t), ("recon", C.POINTER(C.c_double)),
                ("main_tree_bytes", C.c_size_t), ("main_node_count", C.c_size_t), ("main_bits_bytes", C.c_size_t)]


class Oracle:
    """Binding of oracle/libszoracle.so (our CPU restatement)."""

    def __init__(self, path=ORACLE_SO):
        self.lib = C.CDLL(path)
        L = self.lib
        L.szo_default_config.argtypes = [C.POINTER(szo_config)]
        L.szo_compress.argtypes = [C.c_int, C.c_void_p] + [C.c_size_t] * 5 + [C.c_int, C.c_double, C.c_double, C.POINTER(szo_config),
                                   C.POINTER(C.c_void_p), C.POINTER(C.c_size_t), C.POINTER(szo_trace)]
        L.szo_compress.restype = C.c_int
        L.szo_decompress.argtypes = [C.c_int, C.c_void_p, C.c_size_t] + [C.c_size_t] * 5 + [C.POINTER(C.c_void_p)]
        L.szo_decompress.restype = C.c_int
        L.szo_trace_free.argtypes = [C.POINTER(szo_trace)]
        L.szo_free.argtypes = [C.c_void_p]
        L.szo_filter_dims.argtypes = [C.c_size_t] * 5 + [C.POINTER(C.c_size_t)]
        L.szo_filter_dims.restype = C.c_int
        L.szo_huffman_build.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_uint), C.POINTER(C.c_uint), C.c_void_p, C.c_void_p]
        L.szo_huffman_build.restype = C.c_int
        L.szo_huffman_encode.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
        L.szo_huffman_encode.restype = C.c_size_t
        L.szo_huffman_decode.argtypes = [C.c_void_p, C.c_uint, C.c_void_p, C.c_size_t, C.c_void_p]
        L.szo_huffman_decode.restype = C.c_int
        self.cfg = szo_config()
        L.szo_default_config(C.byref(self.cfg))

    def init(self, cfg=None):
        c = dict(DEFAULT_CFG)
        c.update(cfg or {})
        self.lib.szo_default_config(C.byref(self.cfg))
        self.cfg.max_quant_intervals = c["max_quant_intervals"]
        self.cfg.quantization_intervals = c["quantization_intervals"]
        if c["quantization_intervals"] > 0:
            self.cfg.max_quant_intervals = c["quantization_intervals"]   # conf.c:196
        self.cfg.sampleDistance = c["sampleDistance"]
        self.cfg.predThreshold = c["predThreshold"]
        self.cfg.protectValueRange = 1 if c["protectValueRange"] == "YES" else 0
        self.cfg.conf_absErrBound = c["absErrBound"]
        self.cfg.conf_relBoundRatio = c["relBoundRatio"]
        self.cfg.psnr = c["psnr"]
        self.cfg.normErr = c["normErr"]

    def compress(self, a, mode, abs_eb=0.0, rel=0.0, trace=False):
        a = np.ascontiguousarray(a)
        out = C.c_void_p()
        n = C.c_size_t(0)
        tr = szo_trace() if trace else None
        rc = self.lib.szo_compress(dtype_code(a), a.ctypes.data, *dims5(a.shape), mode, abs_eb, rel, C.byref(self.cfg),
                                   C.byref(out), C.byref(n), C.byref(tr) if trace else None)
        if rc != 0:
            raise RuntimeError("szo_compress rc=%d" % rc)
        s = C.string_at(out.value, n.value)
        self.lib.szo_free(out)
        if not trace:
            return s
        t = self._trace_to_dict(tr, a.dtype)
        self.lib.szo_trace_free(C.byref(tr))
        return s, t

    @staticmethod
    def _trace_to_dict(tr, dt):
        d = {k: getattr(tr, k) for k, _ in szo_trace._fields_ if not k.startswith(("reg_params", "indicator", "coeff_", "type", "unpred", "recon"))}
        if tr.kind not in (0, 2):
            return d
        ncoef = 4 if tr.dim == 3 else 3
        nb, rc, N = tr.num_blocks, tr.reg_count, tr.N
        d["reg_params"] = np.ctypeslib.as_array(tr.reg_params, (ncoef, nb)).astype(dt)
        d["coeff_last"] = np.ctypeslib.as_array(tr.coeff_last, (ncoef, nb)).astype(dt)
        d["indicator"] = np.ctypeslib.as_array(tr.indicator, (nb,)).copy()
        d["coeff_type"] = np.ctypeslib.as_array(tr.coeff_type, (ncoef, nb)).copy()[:, :rc]
        d["type"] = np.ctypeslib.as_array(tr.type, (N,)).copy()
        d["unpred"] = np.ctypeslib.as_array(tr.unpred, (max(tr.total_unpred, 1),)).astype(dt)[:tr.total_unpred]
        shape = (tr.n0, tr.n1, tr.n2) if tr.dim == 3 else (tr.n0, tr.n1)
        d["recon"] = np.ctypeslib.as_array(tr.recon, (N,)).astype(dt).reshape(shape)
        return d

    def decompress(self, stream, dtype, shape):
        dt = np.dtype(dtype)
        code = SZ_FLOAT if dt == np.float32 else SZ_DOUBLE
        buf = (C.c_ubyte * len(stream)).from_buffer_copy(stream)
        out = C.c_void_p()
        rc = self.lib.szo_decompress(code, buf, len(stream), *dims5(shape), C.byref(out))
        if rc != 0:
            raise RuntimeError("szo_decompress rc=%d" % rc)
        n = int(np.prod(shape))
        a = np.frombuffer(C.string_at(out.value, n * dt.itemsize), dtype=dt).reshape(shape).copy()
        self.lib.szo_free(out)
        return a

    def filter_dims(self, r5, r4, r3, r2, r1):
        r = (C.c_size_t * 5)()
        rc = self.lib.szo_filter_dims(r5, r4, r3, r2, r1, r)
        return rc, list(r)

    def huffman_build(self, s, stateNum):
        s = np.ascontiguousarray(s, dtype=np.int32)
        tb = C.c_void_p()
        tbs, nc = C.c_uint(0), C.c_uint(0)
        code = np.zeros(stateNum, dtype=np.uint64)
        clen = np.zeros(stateNum, dtype=np.uint8)
        rc = self.lib.szo_huffman_build(s.ctypes.data, s.size, stateNum, C.byref(tb), C.byref(tbs), C.byref(nc), code.ctypes.data, clen.ctypes.data)
        if rc != 0:
            raise RuntimeError("huffman_build rc=%d" % rc)
        tree = C.string_at(tb.value, tbs.value)
        self.lib.szo_free(tb)
        return tree, nc.value, code, clen

    def huffman_encode(self, s, code, clen):
        s = np.ascontiguousarray(s, dtype=np.int32)
        out = np.zeros(s.size * 8 + 16, dtype=np.uint8)
        n = self.lib.szo_huffman_encode(s.ctypes.data, s.size, code.ctypes.data, clen.ctypes.data, out.ctypes.data)
        return out[:n].tobytes()


# ---------------------------------------------------------------------------------------------------
# synthetic inputs
def smooth_field(shape, dtype=np.float32, seed=1234, span=8 * np.pi, field_id=0, noise=1e-3):
    rng = np.random.default_rng(seed + field_id)
    axes = [np.linspace(0, span, n) for n in shape]
    ph = 0.37 * field_id
    if len(shape) == 3:
        z, y, x = np.meshgrid(*axes, indexing="ij", sparse=True)
        f = np.sin(x + ph) * np.cos(1.3 * y) * np.sin(0.7 * z + 0.3) + 0.25 * np.sin(3.1 * x + 1.7 * y) * np.cos(2.3 * z)
    elif len(shape) == 2:
        y, x = np.meshgrid(*axes, indexing="ij", sparse=True)
        f = np.sin(x + ph) * np.cos(1.3 * y) + 0.25 * np.sin(3.1 * x + 1.7 * y)
    else:
        w, z, y, x = np.meshgrid(*axes, indexing="ij", sparse=True)
        f = np.sin(x + ph) * np.cos(1.3 * y) * np.sin(0.7 * z + 0.3) * np.cos(0.5 * w) + 0.25 * np.sin(3.1 * x + 1.7 * y) * np.cos(2.3 * z)
    f = f + noise * rng.standard_normal(shape)
    return np.ascontiguousarray(f.astype(dtype))


def background_field(shape, dtype=np.float32, seed=7):
    """>=65 % of points on one plateau -> use_mean fires."""
    f = smooth_field(shape, np.float64, seed=seed, span=4 * np.pi, noise=1e-4)
    thr = np.percentile(f, 35)
    return np.ascontiguousarray(np.minimum(f, thr).astype(dtype))


def plane_field(shape, dtype=np.float32, seed=11):
    """plane + noise -> (nearly) all blocks choose regression."""
    rng = np.random.default_rng(seed)
    idx = np.meshgrid(*[np.arange(n, dtype=np.float64) for n in shape], indexing="ij", sparse=True)
    if len(shape) == 3:
        z, y, x = idx
        f = 0.31 * x + 0.17 * y - 0.23 * z + 3 + 2 * np.sin(0.02 * x + 0.03 * y + 0.01 * z)
    else:
        y, x = idx
        f = 0.31 * x + 0.17 * y + 3 + 2 * np.sin(0.02 * x + 0.03 * y)
    f = f + 0.05 * rng.standard_normal(shape)
    return np.ascontiguousarray(f.astype(dtype))


def random_field(shape, dtype=np.float32, seed=0):
    rng = np.random.default_rng(seed)
    return np.ascontiguousarray(rng.random(shape).astype(dtype))


def poly_field(shape, dtype=np.float32, order=2):
    """deterministic polynomial in the flat index (test/cunit-test/RegressionTest.hpp:108-123 family)."""
    n = int(np.prod(shape))
    i = np.arange(n, dtype=np.float64) / n
    f = sum(((-1) ** k) * (k + 1) * i ** k for k in range(order + 1))
    return np.ascontiguousarray(f.reshape(shape).astype(dtype))


def blocky_field(n, dtype, seed=7):
    """n^3 field (n a multiple of 6) that is an independent random plane inside every 6^3 block: every block is a regression
    block and the quantised coefficients are spread over tens of thousands of distinct codes -- the case in which the four
    coefficient Huffman trees are as large as they get."""
    rng = np.random.RandomState(seed)
    nb = n // 6
    A, B, Cc, D = [rng.uniform(-1, 1, size=(nb, nb, nb)) for _ in range(4)]
    ii = np.arange(n) % 6
    up = lambda X: np.repeat(np.repeat(np.repeat(X, 6, 0), 6, 1), 6, 2)
    f = up(A) * ii[:, None, None] + up(B) * ii[None, :, None] + up(Cc) * ii[None, None, :] + 10 * up(D)
    return np.ascontiguousarray(f.astype(dtype))


def noisy_quadratic(shape, dtype=np.float32, seed=3):
    """a smooth quadratic + oscillation with noise a few error bounds wide (REL 1e-3): every block is a regression block and the quantised
    coefficients spread over a thousand and more codes -- the field for the coefficient chain, its Huffman trees and their bit streams"""
    rng = np.random.default_rng(seed)
    g = np.meshgrid(*[np.linspace(-1, 1, n) for n in shape], indexing="ij")
    f = sum((i + 1) * x for i, x in enumerate(g)) + 0.7 * sum(x * x for x in g) + np.sin(3 * g[0]) * np.cos(2 * g[-1])
    f = f + rng.normal(0, 4e-3 * (f.max() - f.min()), size=shape)
    return np.ascontiguousarray(f.astype(dtype))


def ref_compress_file(path, dtype, shape, cfg, mode, abs_eb, rel, want_dec=False):
    """Worker for process pools: run the compiled reference on an array stored as .npy; returns (size, md5, dec_md5 or None,
    seconds inside SZ_compress_args).  The array is memory-mapped, the reference gets a private contiguous copy."""
    import hashlib
    import time
    a = np.ascontiguousarray(np.load(path, mmap_mode="r")).astype(dtype, copy=False).reshape(shape)
    ref = SZLib(REF_SO)
    ref.init(cfg)
    t0 = time.perf_counter()
    s = ref.compress(a, mode, abs_eb, rel)
    dt = time.perf_counter() - t0
    dec = None
    if want_dec:
        dec = hashlib.md5(ref.decompress(s, a.dtype, a.shape).tobytes()).hexdigest()
    ref.finalize()
    return len(s), hashlib.md5(s).hexdigest(), dec, dt
