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
        for m in re.finditer(r"\b(SZ_\w+|szb200_\w+|computeDataLength|computeDimension|filterDimension|convertSZParamsToBytes|convertBytesToSZParams)\s*\(", txt):
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
