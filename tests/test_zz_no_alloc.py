"""GPU test of szb200_range_noalloc(), the one entry point of libSZ_b200.so that allocates nothing (it is meant for host
codes that own all of HBM through their own arena, or that must not trigger cudaMalloc's device-wide synchronisation).
Called through the C-ABI with plain device pointers and checked against szb200_range() -- the allocating entry point that
runs the same reduction -- exactly (integer/bit-exact bar of the other tests; both inputs avoid a zero extreme, whose sign
only szb200_range() pins down).  Two calls in a row on different inputs also show that the scratch left behind by the first
call does not leak into the second.  Kept in its own file so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

import szref

pytestmark = pytest.mark.gpu


def _lib():
    import sz_b200
    L = sz_b200.lib()
    L.szb200_range_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
    L.szb200_range_noalloc.restype = C.c_int
    return sz_b200, L


def _reference_range(L, code, x):
    vmin, vmax = C.c_double(0), C.c_double(0)
    assert L.szb200_range(code, x.data_ptr(), x.numel(), C.byref(vmin), C.byref(vmax)) == 0
    return vmin.value, vmax.value


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_noalloc_range_matches_allocating_entry_point_twice_in_a_row(dtype):
    import torch
    assert torch.cuda.is_available()
    sz_b200, L = _lib()
    sz_b200.set_device(0)
    code = 0 if dtype == np.float32 else 1
    # first call: many CTAs; second call: a smaller, shifted field on the same scratch
    first = torch.from_numpy(szref.smooth_field((96, 100, 104), dtype, span=4 * np.pi) + dtype(3.0)).cuda()
    second = torch.from_numpy(szref.smooth_field((17, 19, 23), dtype, seed=99) - dtype(7.5)).cuda()
    for x in (first, second):
        out = torch.zeros(2, dtype=x.dtype, device="cuda")
        torch.cuda.synchronize()
        rc = L.szb200_range_noalloc(code, x.data_ptr(), x.numel(), out.data_ptr(), None)
        assert rc == 0
        torch.cuda.synchronize()
        got = out.cpu().numpy()
        want = _reference_range(L, code, x)
        assert float(got[0]) == want[0] and float(got[1]) == want[1]
        assert float(got[0]) == float(x.min().item()) and float(got[1]) == float(x.max().item())


def test_noalloc_range_rejects_oversize_and_bad_arguments():
    import torch
    sz_b200, L = _lib()
    sz_b200.set_device(0)
    x = torch.ones(64, dtype=torch.float32, device="cuda")
    out = torch.zeros(2, dtype=torch.float32, device="cuda")
    limit = 512 * 1024 * 1024
    assert L.szb200_range_noalloc(0, x.data_ptr(), limit // 4 + 1, out.data_ptr(), None) == -4      # refused before any launch
    assert L.szb200_range_noalloc(1, x.data_ptr(), limit // 8 + 1, out.data_ptr(), None) == -4
    assert L.szb200_range_noalloc(7, x.data_ptr(), 64, out.data_ptr(), None) == -3
    assert L.szb200_range_noalloc(0, x.data_ptr(), 0, out.data_ptr(), None) == -1
    torch.cuda.synchronize()
    assert L.szb200_range_noalloc(0, x.data_ptr(), 64, out.data_ptr(), None) == 0
    torch.cuda.synchronize()
    assert out.cpu().tolist() == [1.0, 1.0]
