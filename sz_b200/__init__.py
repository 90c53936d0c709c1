"""sz_b200 -- Python plumbing around libSZ_b200.so (the B200-native SZ2 hot path).

The product is the shared library (C host layer + hand-written sm_100a kernels, see include/sz.h and
include/szb200.h).  This package only loads it through ctypes and offers numpy / torch convenience wrappers for
tests, bench.py and multi-GPU slab sharding.  There is NO CPU fallback: if the library is missing, or no CUDA device
is present, the calls fail loudly.
"""
from .api import (ABS, ABS_AND_REL, ABS_OR_REL, REL, SZ_DOUBLE, SZ_FLOAT, LIB_PATH, SZError, Stats, compress, compress_device,
                  compress_view, compress_bound, compress_batch_device, decompress_batch_device, decompress, decompress_device, finalize, init, last_stats, lib, set_device, stage_names,
                  compress_slabs, decompress_slabs, SzParams)

__all__ = ["ABS", "REL", "ABS_AND_REL", "ABS_OR_REL", "SZ_FLOAT", "SZ_DOUBLE", "LIB_PATH", "SZError", "Stats", "compress", "compress_device",
           "compress_view", "compress_bound", "compress_batch_device", "decompress_batch_device", "decompress", "decompress_device", "finalize", "init", "last_stats", "lib", "set_device", "stage_names", "compress_slabs", "decompress_slabs", "SzParams"]
