"""ctypes binding of libSZ_b200.so.  Mirrors the reference's operator interface for this path
(SZ_Init / SZ_compress_args / SZ_decompress, ref: sz/include/sz.h:253-334) plus the device-pointer variants
declared in include/szb200.h.  numpy arrays are HOST buffers; torch CUDA tensors are DEVICE buffers."""
import ctypes as C
import os
import tempfile

import numpy as np

SZ_FLOAT, SZ_DOUBLE = 0, 1
ABS, REL, ABS_AND_REL, ABS_OR_REL = 0, 1, 2, 3
_MODE_NAMES = {ABS: "ABS", REL: "REL", ABS_AND_REL: "ABS_AND_REL", ABS_OR_REL: "ABS_OR_REL"}

LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", os.environ.get("SZB200_LIB", "libSZ_b200.so"))


class SZError(RuntimeError):
    pass


class Stats(C.Structure):
    _fields_ = [("intervals", C.c_uint), ("use_mean", C.c_int), ("mean", C.c_double), ("num_blocks", C.c_size_t), ("reg_count", C.c_size_t),
                ("total_unpred", C.c_size_t), ("tree_bytes", C.c_uint), ("node_count", C.c_uint), ("type_bits_bytes", C.c_size_t),
                ("body_bytes", C.c_size_t), ("ms_total", C.c_float), ("ms_stage", C.c_float * 16), ("launches", C.c_uint)]

    def as_dict(self):
        d = {k: getattr(self, k) for k, _ in self._fields_ if k != "ms_stage"}
        d["ms_stage"] = {n: self.ms_stage[i] for i, n in enumerate(stage_names()) if n}
        return d


_lib = None


def lib():
    """Load libSZ_b200.so (built in-tree by `__graft_entry__.build()` / `make -C sz_b200/csrc`). Fails loudly if absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise SZError("libSZ_b200.so not built: run `python -c 'import __graft_entry__ as g; g.build()'` (no CPU fallback exists)")
    L = C.CDLL(LIB_PATH)
    sz = C.c_size_t
    L.SZ_Init.argtypes = [C.c_char_p]; L.SZ_Init.restype = C.c_int
    L.SZ_Finalize.argtypes = []; L.SZ_Finalize.restype = None
    L.SZ_compress_args.argtypes = [C.c_int, C.c_void_p, C.POINTER(sz), C.c_int, C.c_double, C.c_double, C.c_double] + [sz] * 5
    L.SZ_compress_args.restype = C.c_void_p
    L.SZ_decompress.argtypes = [C.c_int, C.c_void_p, sz] + [sz] * 5; L.SZ_decompress.restype = C.c_void_p
    L.SZ_decompress_args.argtypes = [C.c_int, C.c_void_p, sz, C.c_void_p] + [sz] * 5; L.SZ_decompress_args.restype = sz
    L.szb200_compress_bound.argtypes = [C.c_int] + [sz] * 5; L.szb200_compress_bound.restype = sz
    L.szb200_compress_args_device.argtypes = [C.c_int, C.c_void_p, C.c_void_p, sz, C.POINTER(sz), C.c_int, C.c_double, C.c_double] + [sz] * 5
    L.szb200_compress_args_device.restype = C.c_int
    L.szb200_decompress_device.argtypes = [C.c_int, C.c_void_p, sz, C.c_void_p] + [sz] * 5; L.szb200_decompress_device.restype = C.c_int
    L.szb200_last_stats.argtypes = [C.POINTER(Stats)]; L.szb200_last_stats.restype = C.c_int
    L.szb200_last_error.argtypes = []; L.szb200_last_error.restype = C.c_char_p
    L.szb200_stage_name.argtypes = [C.c_int]; L.szb200_stage_name.restype = C.c_char_p
    L.szb200_set_device.argtypes = [C.c_int]; L.szb200_set_device.restype = C.c_int
    L.szb200_device_count.argtypes = []; L.szb200_device_count.restype = C.c_int
    L.szb200_debug_fetch.argtypes = [C.c_int, C.c_void_p, sz]; L.szb200_debug_fetch.restype = C.c_longlong
    L.szb200_release.argtypes = []; L.szb200_release.restype = None
    L.szb200_sync.argtypes = []; L.szb200_sync.restype = C.c_int
    L.szb200_stream.argtypes = []; L.szb200_stream.restype = C.c_void_p
    L.szb200_range.argtypes = [C.c_int, C.c_void_p, sz, C.POINTER(C.c_double), C.POINTER(C.c_double)]; L.szb200_range.restype = C.c_int
    _lib = L
    return L


_libc = C.CDLL(None)
_libc.free.argtypes = [C.c_void_p]


def stage_names():
    L = lib()
    return [L.szb200_stage_name(i).decode() for i in range(16)]


def _dims5(shape):
    r = list(reversed(shape)) + [0] * (5 - len(shape))
    return r[4], r[3], r[2], r[1], r[0]


def _code(dtype):
    dt = np.dtype(dtype)
    if dt == np.float32:
        return SZ_FLOAT
    if dt == np.float64:
        return SZ_DOUBLE
    raise SZError("only float32 / float64 are on this path (got %s)" % dt)


_DEFAULT_CFG = dict(max_quant_intervals=65536, quantization_intervals=0, sampleDistance=100, predThreshold=0.99, szMode="SZ_BEST_SPEED",
                    protectValueRange="NO", withLinearRegression="YES", errorBoundMode="ABS", absErrBound=1e-4, relBoundRatio=1e-4)


def init(cfg=None, path=None):
    """SZ_Init with an sz.config written from `cfg` (same keys as the reference's example/sz.config), or with `path`."""
    L = lib()
    if path is not None:
        rc = L.SZ_Init(path.encode())
    else:
        c = dict(_DEFAULT_CFG)
        c.update(cfg or {})
        fd, p = tempfile.mkstemp(suffix=".config")
        os.close(fd)
        try:
            with open(p, "w") as f:
                f.write("[ENV]\ndataEndianType = LITTLE_ENDIAN_DATA\nsol_name = SZ\n[PARAMETER]\n")
                for k in ("withLinearRegression", "protectValueRange", "sampleDistance", "quantization_intervals", "max_quant_intervals", "szMode", "errorBoundMode"):
                    f.write("%s = %s\n" % (k, c[k]))
                f.write("predThreshold = %r\nabsErrBound = %r\nrelBoundRatio = %r\n" % (c["predThreshold"], float(c["absErrBound"]), float(c["relBoundRatio"])))
                f.write("psnr = %r\nnormErr = %r\n" % (float(c.get("psnr", 80.0)), float(c.get("normErr", 0.05))))
            rc = L.SZ_Init(p.encode())
        finally:
            os.unlink(p)
    if rc != 0:
        raise SZError("SZ_Init failed (rc=%d)" % rc)


def finalize():
    lib().SZ_Finalize()


def set_device(dev):
    if lib().szb200_set_device(int(dev)) != 0:
        raise SZError(lib().szb200_last_error().decode())


def compress(a, mode, abs_eb=0.0, rel=0.0):
    """SZ_compress_args on a HOST numpy array -> bytes (H2D, kernels, D2H all inside the call)."""
    L = lib()
    a = np.ascontiguousarray(a)
    n = C.c_size_t(0)
    p = L.SZ_compress_args(_code(a.dtype), a.ctypes.data, C.byref(n), int(mode), float(abs_eb), float(rel), 0.0, *_dims5(a.shape))
    if not p:
        raise SZError("SZ_compress_args failed: %s" % L.szb200_last_error().decode())
    out = C.string_at(p, n.value)
    _libc.free(p)
    return out


class _Malloced(object):
    """owner of a malloc'ed block returned by the C API; frees it when the last numpy view dies"""

    def __init__(self, ptr, nbytes):
        self.ptr, self.nbytes = ptr, nbytes
        self.__array_interface__ = {"data": (ptr, False), "shape": (nbytes,), "typestr": "|u1", "version": 3}

    def __del__(self):
        if self.ptr:
            _libc.free(self.ptr)
            self.ptr = 0


def compress_view(a, mode, abs_eb=0.0, rel=0.0):
    """SZ_compress_args on a HOST numpy array -> uint8 numpy array that WRAPS the buffer the C call returned (no second copy;
    the buffer is freed with the array).  This is the call bench.py times as `e2e`."""
    L = lib()
    a = np.ascontiguousarray(a)
    n = C.c_size_t(0)
    p = L.SZ_compress_args(_code(a.dtype), a.ctypes.data, C.byref(n), int(mode), float(abs_eb), float(rel), 0.0, *_dims5(a.shape))
    if not p:
        raise SZError("SZ_compress_args failed: %s" % L.szb200_last_error().decode())
    return np.asarray(_Malloced(p, n.value))


def decompress(stream, dtype, shape, out=None):
    """SZ_decompress_args of a HOST byte stream -> numpy array (written straight into `out` when given)."""
    L = lib()
    dt = np.dtype(dtype)
    buf = np.frombuffer(stream, dtype=np.uint8)
    if out is None:
        out = np.empty(shape, dtype=dt)
    assert out.dtype == dt and out.flags.c_contiguous and out.size == int(np.prod(shape))
    if out.size == 0:
        return out
    got = L.SZ_decompress_args(_code(dt), buf.ctypes.data, buf.size, out.ctypes.data, *_dims5(shape))
    if got != out.size:
        raise SZError("SZ_decompress_args failed: %s" % L.szb200_last_error().decode())
    return out


def compress_bound(dtype, shape):
    return int(lib().szb200_compress_bound(_code(dtype), *_dims5(shape)))


def _order_after_current_stream():
    """The library works on its own stream: make it wait (on the device) for everything already enqueued on torch's current
    stream -- the kernel that is still writing the input, a collective that still reads the output buffer."""
    import torch
    ev = torch.cuda.Event()
    ev.record(torch.cuda.current_stream())
    L = lib()
    L.szb200_wait_event.argtypes = [C.c_void_p]
    L.szb200_wait_event.restype = C.c_int
    rc = L.szb200_wait_event(C.c_void_p(ev.cuda_event))
    if rc != 0:
        raise SZError("szb200_wait_event rc=%d: %s" % (rc, L.szb200_last_error().decode()))


def compress_device(x, out, mode, abs_eb=0.0, rel=0.0):
    """Device-resident compress: x is a contiguous torch CUDA tensor (float32/float64), out a uint8 CUDA tensor of at least
    compress_bound() bytes.  Returns the stream length."""
    import torch
    assert x.is_cuda and out.is_cuda and x.is_contiguous() and out.dtype == torch.uint8
    code = SZ_FLOAT if x.dtype == torch.float32 else SZ_DOUBLE
    n = C.c_size_t(0)
    L = lib()
    _order_after_current_stream()
    rc = L.szb200_compress_args_device(code, x.data_ptr(), out.data_ptr(), out.numel(), C.byref(n), int(mode), float(abs_eb), float(rel), *_dims5(tuple(x.shape)))
    if rc != 0:
        raise SZError("szb200_compress_args_device rc=%d: %s" % (rc, L.szb200_last_error().decode()))
    return n.value


def decompress_device(stream, nbytes, out):
    """Device-resident decompress: stream is a uint8 CUDA tensor, out a contiguous CUDA tensor of the original shape/dtype."""
    import torch
    assert stream.is_cuda and out.is_cuda and out.is_contiguous()
    code = SZ_FLOAT if out.dtype == torch.float32 else SZ_DOUBLE
    L = lib()
    _order_after_current_stream()
    rc = L.szb200_decompress_device(code, stream.data_ptr(), int(nbytes), out.data_ptr(), *_dims5(tuple(out.shape)))
    if rc != 0:
        raise SZError("szb200_decompress_device rc=%d: %s" % (rc, L.szb200_last_error().decode()))
    return out


class SzParams(C.Structure):
    """sz_params (include/sz.h:87-118 == reference sz/include/sz.h:164-198), for the customize / thread-safe calls"""
    _fields_ = [("dataType", C.c_int), ("max_quant_intervals", C.c_uint), ("quantization_intervals", C.c_uint), ("maxRangeRadius", C.c_uint),
                ("sol_ID", C.c_int), ("losslessCompressor", C.c_int), ("sampleDistance", C.c_int), ("predThreshold", C.c_float), ("szMode", C.c_int),
                ("gzipMode", C.c_int), ("errorBoundMode", C.c_int), ("absErrBound", C.c_double), ("relBoundRatio", C.c_double), ("psnr", C.c_double),
                ("normErr", C.c_double), ("pw_relBoundRatio", C.c_double), ("segment_size", C.c_int), ("pwr_type", C.c_int), ("protectValueRange", C.c_int),
                ("fmin", C.c_float), ("fmax", C.c_float), ("dmin", C.c_double), ("dmax", C.c_double), ("snapshotCmprStep", C.c_int), ("predictionMode", C.c_int),
                ("accelerate_pw_rel_compression", C.c_int), ("plus_bits", C.c_int), ("randomAccess", C.c_int), ("withRegression", C.c_int)]


class BatchItem(C.Structure):
    """szb200_batch_item (include/szb200.h)"""
    _fields_ = [("d_data", C.c_void_p), ("d_stream", C.c_void_p), ("stream_capacity", C.c_size_t), ("out_size", C.c_size_t),
                ("d_out", C.c_void_p), ("rc", C.c_int)]


def compress_batch_device(xs, outs, mode, abs_eb=0.0, rel=0.0, lanes=4):
    """Compress a batch of equally shaped CUDA tensors (BASELINE configs 3 / 4) on `lanes` concurrent worker lanes.
    outs[i] is a uint8 CUDA tensor of at least compress_bound() bytes; returns the list of stream lengths."""
    import torch
    assert len(xs) == len(outs) and len(xs) > 0
    shape, dt = tuple(xs[0].shape), xs[0].dtype
    items = (BatchItem * len(xs))()
    for i, (x, o) in enumerate(zip(xs, outs)):
        assert x.is_cuda and o.is_cuda and x.is_contiguous() and tuple(x.shape) == shape and x.dtype == dt and o.dtype == torch.uint8
        items[i].d_data, items[i].d_stream, items[i].stream_capacity = x.data_ptr(), o.data_ptr(), o.numel()
    L = lib()
    L.szb200_compress_batch_device.argtypes = [C.c_int, C.POINTER(BatchItem), C.c_int, C.c_int, C.c_int, C.c_double, C.c_double] + [C.c_size_t] * 5
    L.szb200_compress_batch_device.restype = C.c_int
    code = SZ_FLOAT if dt == torch.float32 else SZ_DOUBLE
    torch.cuda.current_stream().synchronize()        # the worker lanes have streams of their own: inputs must be complete
    rc = L.szb200_compress_batch_device(code, items, len(xs), int(lanes), int(mode), float(abs_eb), float(rel), *_dims5(shape))
    if rc != 0:
        raise SZError("szb200_compress_batch_device rc=%d (items: %s)" % (rc, [it.rc for it in items]))
    return [int(it.out_size) for it in items]


def decompress_batch_device(streams, sizes, outs, lanes=4):
    """Decompress a batch of device-resident streams into equally shaped CUDA tensors."""
    import torch
    assert len(streams) == len(sizes) == len(outs) and len(outs) > 0
    shape, dt = tuple(outs[0].shape), outs[0].dtype
    items = (BatchItem * len(outs))()
    for i, (st, n, o) in enumerate(zip(streams, sizes, outs)):
        assert st.is_cuda and o.is_cuda and o.is_contiguous() and tuple(o.shape) == shape and o.dtype == dt
        items[i].d_stream, items[i].stream_capacity, items[i].d_out = st.data_ptr(), int(n), o.data_ptr()
    L = lib()
    L.szb200_decompress_batch_device.argtypes = [C.c_int, C.POINTER(BatchItem), C.c_int, C.c_int] + [C.c_size_t] * 5
    L.szb200_decompress_batch_device.restype = C.c_int
    code = SZ_FLOAT if dt == torch.float32 else SZ_DOUBLE
    torch.cuda.current_stream().synchronize()
    rc = L.szb200_decompress_batch_device(code, items, len(outs), int(lanes), *_dims5(shape))
    if rc != 0:
        raise SZError("szb200_decompress_batch_device rc=%d (items: %s)" % (rc, [it.rc for it in items]))
    return outs


def compress_slabs(a, mode, abs_eb=0.0, rel=0.0, n_slabs=0, n_devices=0):
    """szb200_compress_slabs: one HOST array cut into slabs along its slowest dimension, the slabs compressed side by side on the
    GPUs of this node (one process, worker threads); returns the slab container as bytes."""
    L = lib()
    sz = C.c_size_t
    L.szb200_compress_slabs.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p), C.POINTER(sz), C.c_int, C.c_double, C.c_double] + [sz] * 5
    L.szb200_compress_slabs.restype = C.c_int
    a = np.ascontiguousarray(a)
    out, n = C.c_void_p(), sz(0)
    rc = L.szb200_compress_slabs(_code(a.dtype), a.ctypes.data, int(n_slabs), int(n_devices), C.byref(out), C.byref(n), int(mode), float(abs_eb), float(rel), *_dims5(a.shape))
    if rc != 0 or not out:
        raise SZError("szb200_compress_slabs rc=%d: %s" % (rc, L.szb200_last_error().decode()))
    res = C.string_at(out, n.value)
    _libc.free(out)
    return res


def decompress_slabs(container, dtype, shape, n_devices=0):
    """szb200_decompress_slabs: the inverse, into a fresh HOST array."""
    L = lib()
    sz = C.c_size_t
    L.szb200_decompress_slabs.argtypes = [C.c_int, C.c_char_p, sz, C.c_int, C.c_void_p] + [sz] * 5
    L.szb200_decompress_slabs.restype = C.c_int
    out = np.empty(shape, dtype=dtype)
    rc = L.szb200_decompress_slabs(_code(np.dtype(dtype)), container, len(container), int(n_devices), out.ctypes.data, *_dims5(tuple(shape)))
    if rc != 0:
        raise SZError("szb200_decompress_slabs rc=%d: %s" % (rc, L.szb200_last_error().decode()))
    return out


def encode_with_tree(sym, stateNum):
    """szb200_encode_with_tree on a uint16 CUDA tensor -> container bytes ([nodeCount][stateNum/2][tree][bits], Huffman.c:790-816)."""
    import torch
    assert sym.is_cuda and sym.dtype in (torch.uint16, torch.int16) and sym.is_contiguous()
    L = lib()
    sz = C.c_size_t
    L.szb200_encode_with_tree.argtypes = [C.c_void_p, sz, C.c_uint, C.c_void_p, C.c_int, sz, C.POINTER(sz)]
    L.szb200_encode_with_tree.restype = C.c_int
    cap = 8 + 16 + 13 * 2 * int(stateNum) + 8 * sym.numel() + 64
    out = np.empty(cap, dtype=np.uint8)
    n = sz(0)
    _order_after_current_stream()
    rc = L.szb200_encode_with_tree(sym.data_ptr(), sym.numel(), int(stateNum), out.ctypes.data, 0, cap, C.byref(n))
    if rc != 0:
        raise SZError("szb200_encode_with_tree rc=%d: %s" % (rc, L.szb200_last_error().decode()))
    return out[:n.value].tobytes()


def decode_with_tree(container, n):
    """szb200_decode_with_tree: container bytes -> uint16 CUDA tensor of n symbols (Huffman.c:865-889)."""
    import torch
    L = lib()
    sz = C.c_size_t
    L.szb200_decode_with_tree.argtypes = [C.c_char_p, sz, C.c_int, sz, C.c_void_p]
    L.szb200_decode_with_tree.restype = C.c_int
    out = torch.empty(int(n), dtype=torch.int16, device="cuda")
    _order_after_current_stream()
    rc = L.szb200_decode_with_tree(container, len(container), 0, int(n), out.data_ptr())
    if rc != 0:
        raise SZError("szb200_decode_with_tree rc=%d: %s" % (rc, L.szb200_last_error().decode()))
    return out


def last_stats():
    s = Stats()
    if lib().szb200_last_stats(C.byref(s)) != 0:
        return None
    return s


def debug_fetch(which, dtype, count):
    buf = np.zeros(count, dtype=dtype)
    n = lib().szb200_debug_fetch(which, buf.ctypes.data, buf.nbytes)
    if n < 0:
        raise SZError("debug_fetch(%d) failed" % which)
    return buf[: n // buf.itemsize]
