"""Shared parity-case table: (name, generator, mode, abs, rel, config overrides).
Sizes are kept small enough for the CPU oracle / reference to finish in well under a second each."""
import os

import numpy as np

import szref
from szref import ABS, REL, ABS_AND_REL, ABS_OR_REL, PSNR, NORM

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def intree(name, dtype, shape):
    """The reference's own example inputs (example/testdata/x86/*.dat), copied as small fixtures."""
    return lambda: np.fromfile(os.path.join(_DATA, name), dtype=dtype).reshape(shape)


def gen(fn, shape, dtype, **kw):
    return lambda: fn(shape, dtype, **kw)


F32, F64 = np.float32, np.float64
TWO_PI = 2 * np.pi

CASES = []


def add(name, g, mode, ab, rl, **cfg):
    CASES.append((name, g, mode, ab, rl, cfg))


# BASELINE.json configs[0] and the other Appendix-D known answers (in-tree files)
add("cfg1_testdouble_8_8_128_abs", intree("testdouble_8_8_128.dat", F64, (128, 8, 8)), ABS, 1e-4, 1e-3)
add("testdouble_8_8_128_rel", intree("testdouble_8_8_128.dat", F64, (128, 8, 8)), REL, 1e-4, 1e-3)
add("testdouble_8_8_128_absrel", intree("testdouble_8_8_128.dat", F64, (128, 8, 8)), ABS_AND_REL, 1e-4, 1e-3)
add("testdouble_2d_64_128_abs", intree("testdouble_8_8_128.dat", F64, (128, 64)), ABS, 1e-4, 1e-3)
add("testfloat_8_8_128_abs", intree("testfloat_8_8_128.dat", F32, (128, 8, 8)), ABS, 1e-4, 1e-3)
add("testfloat_8_8_128_rel", intree("testfloat_8_8_128.dat", F32, (128, 8, 8)), REL, 1e-4, 1e-3)
add("testfloat_8_8_128_absrel", intree("testfloat_8_8_128.dat", F32, (128, 8, 8)), ABS_AND_REL, 1e-4, 1e-3)
add("testfloat_2d_64_128_rel", intree("testfloat_8_8_128.dat", F32, (128, 64)), REL, 1e-4, 1e-3)
add("testdouble_4d_8_8_8_128_absrel", intree("testdouble_8_8_8_128.dat", F64, (128, 8, 8, 8)), ABS_AND_REL, 1e-4, 1e-3)

for _dt, _tag in ((F32, "f32"), (F64, "f64")):
    add("smooth_30x26x40_rel_" + _tag, gen(szref.smooth_field, (30, 26, 40), _dt, span=TWO_PI), REL, 1e-4, 1e-3)
    add("smooth_64c_rel_" + _tag, gen(szref.smooth_field, (64, 64, 64), _dt, span=TWO_PI), REL, 1e-4, 1e-3)
    add("smooth_64c_abs_" + _tag, gen(szref.smooth_field, (64, 64, 64), _dt, span=TWO_PI), ABS, 1e-2, 1e-3)
    add("smooth_64c_absrel_" + _tag, gen(szref.smooth_field, (64, 64, 64), _dt, span=TWO_PI), ABS_AND_REL, 1e-3, 1e-3)
    add("smooth_64c_absorrel_" + _tag, gen(szref.smooth_field, (64, 64, 64), _dt, span=TWO_PI), ABS_OR_REL, 1e-3, 1e-3)
    add("smooth_hf_50x70x90_rel_" + _tag, gen(szref.smooth_field, (50, 70, 90), _dt, span=8 * np.pi), REL, 1e-4, 1e-3)
    add("background_30x26x40_" + _tag, gen(szref.background_field, (30, 26, 40), _dt), REL, 1e-4, 1e-3)
    add("background_48x50x52_" + _tag, gen(szref.background_field, (48, 50, 52), _dt), REL, 1e-4, 1e-3)
    add("plane_30x26x40_" + _tag, gen(szref.plane_field, (30, 26, 40), _dt), ABS, 1e-1, 1e-3)
    add("plane_150x203_" + _tag, gen(szref.plane_field, (150, 203), _dt), ABS, 1e-1, 1e-3)
    add("random_32c_" + _tag, gen(szref.random_field, (32, 32, 32), _dt), ABS, 1e-4, 1e-3)
    add("random_32c_loose_" + _tag, gen(szref.random_field, (32, 32, 32), _dt), ABS, 1e-2, 1e-3)
    add("random_64x70_" + _tag, gen(szref.random_field, (64, 70), _dt), ABS, 1e-4, 1e-3)
    add("random_64x70_loose_" + _tag, gen(szref.random_field, (64, 70), _dt), ABS, 2e-2, 1e-3)
    add("poly_64c_" + _tag, gen(szref.poly_field, (64, 64, 64), _dt, order=3), ABS, 1e-4, 1e-3)
    add("smooth2d_150x203_" + _tag, gen(szref.smooth_field, (150, 203), _dt, span=TWO_PI), REL, 1e-4, 1e-4)
    add("smooth2d_97x333_" + _tag, gen(szref.smooth_field, (97, 333), _dt, span=TWO_PI), ABS, 1e-3, 1e-4)
    add("smooth4d_5x12x13x14_" + _tag, gen(szref.smooth_field, (5, 12, 13, 14), _dt, span=TWO_PI), ABS_AND_REL, 1e-3, 1e-3)
    add("const_10c_" + _tag, lambda _dt=_dt: np.full((10, 10, 10), 3.25, _dt), ABS, 1e-3, 1e-3)
    add("tiny_2x3x3_" + _tag, lambda _dt=_dt: np.arange(18, dtype=_dt).reshape(2, 3, 3), ABS, 1e-3, 1e-3)
    add("small_5c_" + _tag, gen(szref.smooth_field, (5, 5, 5), _dt), ABS, 1e-3, 1e-3)
    add("thin_3x40x7_" + _tag, gen(szref.smooth_field, (3, 40, 7), _dt), ABS, 1e-3, 1e-3)
    add("ragged_13x19x45_" + _tag, gen(szref.smooth_field, (13, 19, 45), _dt, span=TWO_PI), ABS, 1e-3, 1e-3)
    add("squeeze_1x40x50_" + _tag, gen(szref.smooth_field, (1, 40, 50), _dt), ABS, 1e-3, 1e-3)
    add("lossless_16c_" + _tag, gen(szref.random_field, (16, 16, 16), _dt), ABS, 1e-12, 1e-3)
    add("fixed_itv256_40c_" + _tag, gen(szref.smooth_field, (40, 40, 40), _dt, span=TWO_PI), ABS, 1e-3, 1e-3, quantization_intervals=256)
    add("sd7_40c_" + _tag, gen(szref.smooth_field, (40, 40, 40), _dt, span=TWO_PI), ABS, 1e-3, 1e-3, sampleDistance=7)
    add("protect_40c_" + _tag, gen(szref.smooth_field, (40, 40, 40), _dt, span=TWO_PI), ABS, 1e-3, 1e-3, protectValueRange="YES")
    add("psnr70_40c_" + _tag, gen(szref.smooth_field, (40, 40, 40), _dt, span=TWO_PI), PSNR, 1e-3, 1e-3, psnr=70.0)
    add("psnr95_2d_90x120_" + _tag, gen(szref.smooth_field, (90, 120), _dt, span=TWO_PI), PSNR, 1e-3, 1e-3, psnr=95.0)
    add("norm_40c_" + _tag, gen(szref.smooth_field, (40, 40, 40), _dt, span=TWO_PI), NORM, 1e-3, 1e-3, normErr=0.8)
    add("maxitv1024_40c_" + _tag, gen(szref.random_field, (40, 40, 40), _dt), ABS, 1e-4, 1e-3, max_quant_intervals=1024)

CASE_IDS = [c[0] for c in CASES]


def make_cfg(mode, ab, rl, extra):
    cfg = dict(errorBoundMode=szref.MODE_NAMES[mode], absErrBound=ab, relBoundRatio=rl)
    cfg.update(extra)
    return cfg
