"""Stage-level GPU parity (run on the B200 box): the intermediates of one compress job -- regression fit, predictor selection,
coefficient codes and quantised coefficients, quantisation codes, reconstructed field, interval count, mean -- are fetched through
the C-ABI (szb200_debug_fetch) and compared with the traces of the CPU oracle, stage by stage, so that a divergence names the
stage instead of showing up as "md5 differs".  The oracle is the checker only."""
import numpy as np
import pytest

import cases
import szref
from triage import bm_index

pytestmark = pytest.mark.gpu

# regression-stream cases of every family: 2D / 3D / 4D, float / double, with and without use_mean, regression-heavy, fixed intervals
PICK = [c for c in cases.CASES if any(k in c[0] for k in ("smooth_30x26x40", "smooth_hf_50x70x90", "background_48x50x52", "plane_30x26x40", "plane_150x203",
                                                          "smooth2d_150x203", "testdouble_4d", "fixed_itv256", "random_32c_loose", "poly_64c"))]


@pytest.fixture(scope="module")
def sz():
    import torch
    assert torch.cuda.is_available()
    import sz_b200
    sz_b200.set_device(0)
    return sz_b200


def _bits(a):
    a = np.ascontiguousarray(a)
    return a.view("u%d" % a.itemsize) if a.dtype.kind == "f" else a


@pytest.mark.parametrize("case", PICK, ids=[c[0] for c in PICK])
def test_every_stage_equals_the_oracle_trace(sz, case, monkeypatch):
    from sz_b200 import api
    monkeypatch.setenv("SZB200_DEBUG_RECON", "1")               # keep the compressor-side reconstructed field for the fetch
    name, g, mode, ab, rl, extra = case
    a = g()
    cfg = cases.make_cfg(mode, ab, rl, extra)
    orc = szref.Oracle()
    orc.init(cfg)
    so, tr = orc.compress(a, mode, ab, rl, trace=True)
    if tr["kind"] != 0:
        pytest.skip("not a regression stream")
    sz.init(cfg)
    sg = sz.compress(a, mode, ab, rl)
    st = sz.last_stats()
    dim, n0, n1, n2 = tr["dim"], tr["n0"], tr["n1"], tr["n2"]
    if dim == 2:
        n0, n1, n2 = 1, tr["n0"], tr["n1"]                   # the oracle's trace names the two dimensions n0, n1
    ncoef, nb, N, rc = (4 if dim == 3 else 3), tr["num_blocks"], tr["N"], tr["reg_count"]
    assert (st.intervals, st.use_mean, st.reg_count, st.total_unpred) == (tr["intervals"], tr["use_mean"], rc, tr["total_unpred"]), "interval optimiser / mean decision"
    assert np.array([st.mean], dtype=a.dtype).tobytes() == np.array([tr["mean"]], dtype=a.dtype).tobytes(), "mean"
    reg = api.debug_fetch(0, a.dtype, ncoef * nb).reshape(ncoef, nb)
    assert np.array_equal(_bits(reg), _bits(np.asarray(tr["reg_params"]).reshape(ncoef, nb).astype(a.dtype))), "regression fit (a7)"
    ind = api.debug_fetch(1, np.uint8, nb)
    assert np.array_equal(ind, np.asarray(tr["indicator"]).astype(np.uint8)), "predictor selection (a8)"
    if rc:
        ct = api.debug_fetch(5, np.uint16, ncoef * nb).reshape(ncoef, nb)[:, :rc]
        assert np.array_equal(ct.astype(np.int64), np.asarray(tr["coeff_type"]).reshape(ncoef, rc)), "coefficient codes (a9)"
        cl = api.debug_fetch(2, a.dtype, ncoef * nb).reshape(ncoef, nb)
        m = np.asarray(tr["indicator"]) == 0
        assert np.array_equal(_bits(cl[:, m]), _bits(np.asarray(tr["coeff_last"]).reshape(ncoef, nb)[:, m].astype(a.dtype))), "quantised coefficients (a9)"
    codes = api.debug_fetch(3, np.uint16, N)
    p = bm_index(n0, n1, n2, dim)
    assert np.array_equal(codes.astype(np.int64), np.asarray(tr["type"])[p]), "quantisation codes (a10 / a11)"
    rec = api.debug_fetch(4, a.dtype, N)
    assert np.array_equal(_bits(rec), _bits(np.asarray(tr["recon"]).reshape(-1).astype(a.dtype))), "reconstructed field (a10 / a11)"
    assert st.tree_bytes == tr["main_tree_bytes"] and st.node_count == tr["main_node_count"] and st.type_bits_bytes == tr["main_bits_bytes"], "Huffman tree / bit stream sizes (a13-a15)"
    assert sg == so, "stream assembly (a16)"
    sz.finalize()
