"""Worker of tests/test_gpu_parity.py::test_coefficient_chain_forms_give_the_reference_stream (run in a fresh process because the
library reads SZB200_CHAIN / SZB200_LAT_BAND once): compress fields with many regression blocks through the C-ABI and compare every
stream byte for byte with the CPU oracle (the checker).  Prints one line per field and "chain forms check: ok" / "FAILED"."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))


def main():
    import szref
    import sz_b200
    cfg = dict(errorBoundMode="REL", relBoundRatio=1e-3)
    sz_b200.init(cfg)
    orc = szref.Oracle()
    orc.init(cfg)
    fields = [("noisy quadratic f32 3D", szref.noisy_quadratic((120, 96, 102), np.float32)),      # every block a regression block, ~1500 distinct codes
              ("noisy quadratic f64 2D", szref.noisy_quadratic((640, 700), np.float64)),
              ("blocky f32", szref.blocky_field(120, np.float32)),           # coefficient codes spread over the whole range, unpredictable ones among them
              ("blocky f64", szref.blocky_field(60, np.float64)),
              ("plane f32", szref.plane_field((66, 70, 74), np.float32))]
    bad = 0
    for name, a in fields:
        got = sz_b200.compress(a, sz_b200.REL, 1e-4, 1e-3)
        want = orc.compress(a, szref.REL, 1e-4, 1e-3)
        ok = bytes(got) == want
        st = sz_b200.last_stats()
        print("%s: %d bytes, regression blocks %d of %d, %s" % (name, len(want), st.reg_count, st.num_blocks, "equal" if ok else "DIFFERENT"), flush=True)
        bad += 0 if ok else 1
    sz_b200.finalize()
    print("chain forms check: %s" % ("ok" if bad == 0 else "FAILED"), flush=True)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
