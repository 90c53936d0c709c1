"""CPU tests (no GPU): the C-ABI library loads, exports every symbol include/*.h declares, and its host-side logic
(config parsing, dimension filter, params block) matches the reference's when oracle/_ref is present."""
import ctypes as C
import itertools
import os
import re

import numpy as np
import pytest

import szref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "sz_b200", "lib", "libSZ_b200.so")

pytestmark = pytest.mark.skipif(not os.path.exists(LIB), reason="libSZ_b200.so not built (python -c 'import __graft_entry__ as g; g.build()')")


class sz_params(C.Structure):      # ref: sz/include/sz.h:164-198
    _fields_ = [("dataType", C.c_int), ("max_quant_intervals", C.c_uint), ("quantization_intervals", C.c_uint), ("maxRangeRadius", C.c_uint),
                ("sol_ID", C.c_int), ("losslessCompressor", C.c_int), ("sampleDistance", C.c_int), ("predThreshold", C.c_float), ("szMode", C.c_int),
                ("gzipMode", C.c_int), ("errorBoundMode", C.c_int), ("absErrBound", C.c_double), ("relBoundRatio", C.c_double), ("psnr", C.c_double),
                ("normErr", C.c_double), ("pw_relBoundRatio", C.c_double), ("segment_size", C.c_int), ("pwr_type", C.c_int),
                ("protectValueRange", C.c_int), ("fmin", C.c_float), ("fmax", C.c_float), ("dmin", C.c_double), ("dmax", C.c_double),
                ("snapshotCmprStep", C.c_int), ("predictionMode", C.c_int), ("accelerate_pw_rel_compression", C.c_int), ("plus_bits", C.c_int),
                ("randomAccess", C.c_int), ("withRegression", C.c_int)]


class sz_exedata(C.Structure):     # ref: sz/include/sz.h:211-217
    _fields_ = [("optQuantMode", C.c_char), ("intvCapacity", C.c_int), ("intvRadius", C.c_int), ("SZ_SIZE_TYPE", C.c_uint)]


def declared_symbols():
    names = set()
    for h in ("sz.h", "szb200.h"):
        txt = open(os.path.join(ROOT, "include", h)).read()
        txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
        for m in re.finditer(r"\b(SZ_\w+|szb200_\w+|computeDataLength|computeDimension|filterDimension|convertSZParamsToBytes|convertBytesToSZParams|"
                             r"readByteData|readFloatData|readDoubleData|writeByteData|writeFloatData_inBytes|writeDoubleData_inBytes|checkFileExistance|"
                             r"writeStrings|writeFloatData|writeDoubleData|detransposeData|is_lossless_compressed_data|sz_lossless_decompress65536bytes)\s*\(", txt):
            names.add(m.group(1))
    return sorted(names)


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(LIB)
    missing = [n for n in declared_symbols() if not hasattr(lib, n)]
    assert not missing, missing
    for g in ("confparams_cpr", "confparams_dec", "exe_params", "versionNumber", "dataEndianType", "sysEndianType"):
        C.c_void_p.in_dll(lib, g)
    v = (C.c_int * 4).in_dll(lib, "versionNumber")
    assert list(v)[:3] == [2, 1, 12]


def _state(lib):
    p = C.POINTER(sz_params).in_dll(lib, "confparams_cpr").contents
    e = C.POINTER(sz_exedata).in_dll(lib, "exe_params").contents
    d = {k: getattr(p, k) for k, _ in sz_params._fields_}
    d.update({"exe_" + k: getattr(e, k) for k, _ in sz_exedata._fields_})
    return d


CFGS = [dict(), dict(errorBoundMode="REL", relBoundRatio=1e-3), dict(errorBoundMode="ABS_AND_REL", absErrBound=3e-5, relBoundRatio=2e-4),
        dict(quantization_intervals=256), dict(max_quant_intervals=1024, sampleDistance=7, predThreshold=0.97), dict(protectValueRange="YES"),
        dict(szMode="SZ_BEST_COMPRESSION"), dict(withLinearRegression="NO")]


@pytest.mark.skipif(not szref.have_ref(), reason="needs oracle/_ref/libSZ_ref.so")
@pytest.mark.parametrize("cfg", CFGS)
def test_config_state_matches_reference(cfg, tmp_path):
    path = str(tmp_path / "sz.config")
    szref.write_config(path, cfg)
    ref, ours = C.CDLL(szref.REF_SO), C.CDLL(LIB)
    for L in (ref, ours):
        L.SZ_Init.argtypes = [C.c_char_p]
        assert L.SZ_Init(path.encode()) == 0
    a, b = _state(ref), _state(ours)
    assert a == b
    # NULL config -> built-in defaults
    assert ref.SZ_Init(None) == 0 and ours.SZ_Init(None) == 0
    assert _state(ref) == _state(ours)
    # params block of the stream header (ByteToolkit.c:874-972)
    for L in (ref, ours):
        L.SZ_Init(path.encode())
    ra, rb = (C.c_ubyte * 36)(), (C.c_ubyte * 36)()
    for L, buf in ((ref, ra), (ours, rb)):
        p = C.POINTER(sz_params).in_dll(L, "confparams_cpr")
        p.contents.fmin, p.contents.fmax, p.contents.dmin, p.contents.dmax = 1.5, 2.5, -3.0, 9.0
        for dt in (0, 1):
            p.contents.dataType = dt
            L.convertSZParamsToBytes(p, buf)
    assert bytes(ra) == bytes(rb)
    for L in (ref, ours):
        L.SZ_Finalize()


def test_missing_config_file_is_an_error():
    ours = C.CDLL(LIB)
    ours.SZ_Init.argtypes = [C.c_char_p]
    assert ours.SZ_Init(b"/nonexistent/sz.config") == -1


@pytest.mark.skipif(not szref.have_ref(), reason="needs oracle/_ref/libSZ_ref.so")
def test_filter_dimension_matches_reference():
    ref, ours = C.CDLL(szref.REF_SO), C.CDLL(LIB)
    for L in (ref, ours):
        L.filterDimension.argtypes = [C.c_size_t] * 5 + [C.POINTER(C.c_size_t)]
        L.computeDataLength.argtypes = [C.c_size_t] * 5
        L.computeDataLength.restype = C.c_size_t
    for dims in itertools.product([0, 1, 2, 5], repeat=5):
        seq = list(reversed(dims))
        nz = [v != 0 for v in seq]
        if nz != sorted(nz, reverse=True) or seq[0] == 0:
            continue
        a, b = (C.c_size_t * 5)(), (C.c_size_t * 5)()
        ra, rb = ref.filterDimension(*dims, a), ours.filterDimension(*dims, b)

        def eff(x):
            o = []
            for v in x:
                if v == 0:
                    break
                o.append(v)
            return o
        assert eff(a) == eff(b) and ra == rb, (dims, list(a), list(b))
        assert ref.computeDataLength(*dims) == ours.computeDataLength(*dims)


def test_no_gpu_means_loud_failure_not_a_cpu_fallback():
    """Without a CUDA device the product must fail (NULL), never silently compute on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    ours = szref.SZLib(LIB)
    ours.init(dict(errorBoundMode="ABS", absErrBound=1e-3))
    a = szref.smooth_field((16, 16, 16), np.float32)
    with pytest.raises(RuntimeError):
        ours.compress(a, szref.ABS, 1e-3, 0)
    ours.finalize()


def _io_lib(path):
    L = C.CDLL(path)
    L.detransposeData.restype = C.c_void_p
    L.detransposeData.argtypes = [C.c_void_p, C.c_int] + [C.c_size_t] * 5
    L.is_lossless_compressed_data.argtypes = [C.c_char_p, C.c_size_t]
    L.writeFloatData.argtypes = [C.c_void_p, C.c_size_t, C.c_char_p, C.POINTER(C.c_int)]
    L.writeDoubleData.argtypes = [C.c_void_p, C.c_size_t, C.c_char_p, C.POINTER(C.c_int)]
    L.writeStrings.argtypes = [C.c_int, C.POINTER(C.c_char_p), C.c_char_p, C.POINTER(C.c_int)]
    L.checkFileExistance.argtypes = [C.c_char_p]
    return L


def test_command_line_helpers_standalone(tmp_path):
    """The rw.h / utility.h helpers the reference's command-line tool links against (sz_io.c): text writers, existence test,
    magic-number test, and detransposeData as the transpose it is (ref: sz/src/utility.c:236-330)."""
    L = _io_lib(LIB)
    libc = C.CDLL(None)
    libc.free.argtypes = [C.c_void_p]
    st = C.c_int(-7)
    a = np.array([1.5, -2.25, 3.0e-8], dtype=np.float32)
    L.writeFloatData(a.ctypes.data, a.size, str(tmp_path / "f.txt").encode(), C.byref(st))
    assert st.value == 0 and (tmp_path / "f.txt").read_text().split() == ["%.30G" % float(v) for v in a]
    d = np.array([0.1, -1e300], dtype=np.float64)
    L.writeDoubleData(d.ctypes.data, d.size, str(tmp_path / "d.txt").encode(), C.byref(st))
    assert st.value == 0 and (tmp_path / "d.txt").read_text().split() == ["%.20G" % float(v) for v in d]
    strs = (C.c_char_p * 2)(b"first line", b"second")
    L.writeStrings(2, strs, str(tmp_path / "s.txt").encode(), C.byref(st))
    assert st.value == 0 and (tmp_path / "s.txt").read_text() == "first line\nsecond\n"
    assert L.checkFileExistance(str(tmp_path / "s.txt").encode()) == 1 and L.checkFileExistance(str(tmp_path / "nope").encode()) == 0
    assert L.is_lossless_compressed_data(bytes([0x28, 0xB5, 0x2F, 0xFD, 0, 0]), 6) == 1       # a zstd frame
    assert L.is_lossless_compressed_data(bytes([120, 156, 1, 2]), 4) == 0                    # a zlib stream
    assert L.is_lossless_compressed_data(bytes([2, 1, 12, 0xC0]), 4) == -1                   # an SZ_BEST_SPEED stream: neither
    for dtype, code in ((np.float32, 0), (np.float64, 1)):
        for dims in ((0, 0, 0, 3, 5), (0, 0, 4, 3, 5), (0, 2, 4, 3, 5), (0, 0, 0, 0, 7)):
            r5, r4, r3, r2, r1 = dims
            n = int(np.prod([v for v in dims if v]))
            x = np.arange(n, dtype=dtype)
            ptr = L.detransposeData(x.ctypes.data, code, *dims)
            got = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_float if code == 0 else C.c_double)), (n,)).copy()
            libc.free(ptr)
            fast = {1: 1, 2: r1, 3: r3, 4: r4}[sum(1 for v in dims if v)]   # extent of the input's fastest axis
            want = x.reshape(n // fast, fast).T.reshape(-1) if fast > 1 else x
            assert np.array_equal(got, want), (dtype, dims)


@pytest.mark.skipif(not szref.have_ref(), reason="needs oracle/_ref/libSZ_ref.so")
def test_command_line_helpers_match_reference(tmp_path):
    ref, ours = _io_lib(szref.REF_SO), _io_lib(LIB)
    libc = C.CDLL(None)
    libc.free.argtypes = [C.c_void_p]
    rng = np.random.default_rng(5)
    for dims in ((0, 0, 0, 6, 4), (0, 0, 5, 6, 4), (0, 3, 5, 6, 4)):
        n = int(np.prod([v for v in dims if v]))
        for dtype, code, ct in ((np.float32, 0, C.c_float), (np.float64, 1, C.c_double)):
            x = rng.standard_normal(n).astype(dtype)
            outs = []
            for L in (ref, ours):
                ptr = L.detransposeData(x.ctypes.data, code, *dims)
                outs.append(np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ct)), (n,)).copy())
                libc.free(ptr)
            assert outs[0].tobytes() == outs[1].tobytes(), (dims, dtype)
    for head in ([0x28, 0xB5, 0x2F, 0xFD], [104, 5], [104, 129], [104, 222], [120, 1], [120, 94], [120, 156], [120, 218], [120, 2], [2, 1, 12, 0xC0], [0, 0, 0, 0]):
        b = bytes(head + [0] * 8)
        assert ref.is_lossless_compressed_data(b, len(b)) == ours.is_lossless_compressed_data(b, len(b)), head
    st = C.c_int(0)
    a = rng.standard_normal(50).astype(np.float32)
    d = rng.standard_normal(50)
    for nm, L in (("ref", ref), ("ours", ours)):
        L.writeFloatData(a.ctypes.data, a.size, str(tmp_path / (nm + "_f.txt")).encode(), C.byref(st))
        L.writeDoubleData(d.ctypes.data, d.size, str(tmp_path / (nm + "_d.txt")).encode(), C.byref(st))
    assert (tmp_path / "ref_f.txt").read_bytes() == (tmp_path / "ours_f.txt").read_bytes()
    assert (tmp_path / "ref_d.txt").read_bytes() == (tmp_path / "ours_d.txt").read_bytes()
