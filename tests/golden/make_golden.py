"""Generates tests/golden/streams.json from the UNMODIFIED reference compiled out of /root/reference
(oracle/_ref/libSZ_ref.so, recipe: oracle/Makefile).  Run in the build container only:

    make -C oracle ref && python tests/golden/make_golden.py

For every case of tests/cases.py it records the reference stream's size + md5, the md5 of the reference's
SZ_decompress output, and a few header facts.  tests/test_golden.py checks the CPU oracle (and, on the GPU
box, the CUDA path) against these digests without needing /root/reference.
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import cases  # noqa: E402
import szref  # noqa: E402


def main():
    ref = szref.SZLib(szref.REF_SO)
    out = {}
    for name, g, mode, ab, rl, extra in cases.CASES:
        a = g()
        ref.init(cases.make_cfg(mode, ab, rl, extra))
        s = ref.compress(a, mode, ab, rl)
        rec = None
        if a.size > 20:     # the reference cannot decode its own N<=20 raw streams (no header)
            rec = ref.decompress(s, a.dtype, a.shape)
        ref.finalize()
        kind = "raw" if a.size <= 20 else {0x80: "regression", 0x01: "constant", 0x10: "lossless"}[s[3] & 0x91]
        body = s if kind != "constant" else s[:19] + b"\0" + s[20:]   # byte 19 is uninitialised malloc memory in the reference
        out[name] = dict(input_md5=hashlib.md5(a.tobytes()).hexdigest(), size=len(s), md5=hashlib.md5(body).hexdigest(), kind=kind,
                         dec_md5=None if rec is None else hashlib.md5(rec.tobytes()).hexdigest(),
                         max_err=None if rec is None else float(np.max(np.abs(rec.astype(np.float64) - a.astype(np.float64)))))
        print(name, out[name]["size"], out[name]["md5"], kind)
    with open(os.path.join(HERE, "streams.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)


def full_size():
    """BASELINE.json configs[1] at its own size (512^3 fp32, REL 1e-3, the field of bench.make_field): digests of the reference's
    stream and reconstruction -> tests/golden/full_size.json (the GPU test compares against them when the box generates the very
    same input bytes, and runs the compiled reference itself otherwise)."""
    sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
    import bench
    ref = szref.SZLib(szref.REF_SO)
    out = {}
    a = bench.make_field((512, 512, 512))
    ref.init(dict(errorBoundMode="REL", relBoundRatio=1e-3))
    s = ref.compress(a, szref.REL, 1e-4, 1e-3)
    rec = ref.decompress(s, a.dtype, a.shape)
    ref.finalize()
    out["cfg2_nyx_512_rel"] = dict(input_md5=hashlib.md5(a.tobytes()).hexdigest(), size=len(s), md5=hashlib.md5(s).hexdigest(),
                                   dec_md5=hashlib.md5(rec.tobytes()).hexdigest(),
                                   max_err=float(np.max(np.abs(rec.astype(np.float64) - a.astype(np.float64)))))
    print(out)
    with open(os.path.join(HERE, "full_size.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--full":
        full_size()
    else:
        main()
