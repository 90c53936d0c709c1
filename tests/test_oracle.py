"""CPU tests of the oracle (oracle/libszoracle.so):
  * against the golden digests generated from the unmodified reference (tests/golden/streams.json);
  * byte-for-byte against the compiled reference itself when oracle/_ref/libSZ_ref.so is present;
  * round-trip / error-bound properties.
"""
import hashlib
import json
import os

import numpy as np
import pytest

import cases
import szref

GOLD = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "streams.json")))

pytestmark = pytest.mark.skipif(not szref.have_oracle(), reason="oracle/libszoracle.so not built (python -c 'import __graft_entry__ as g; g.build()')")


@pytest.fixture(scope="module")
def orc():
    return szref.Oracle()


def _norm(stream, kind):
    return stream if kind != "constant" else stream[:19] + b"\0" + stream[20:]


@pytest.mark.parametrize("case", cases.CASES, ids=cases.CASE_IDS)
def test_oracle_matches_golden(orc, case):
    name, g, mode, ab, rl, extra = case
    a = g()
    gold = GOLD[name]
    assert hashlib.md5(a.tobytes()).hexdigest() == gold["input_md5"], "input generator drifted"
    orc.init(cases.make_cfg(mode, ab, rl, extra))
    s = orc.compress(a, mode, ab, rl)
    assert len(s) == gold["size"]
    assert hashlib.md5(_norm(s, gold["kind"])).hexdigest() == gold["md5"]
    rec = orc.decompress(s, a.dtype, a.shape)
    if gold["dec_md5"] is not None:
        assert hashlib.md5(rec.tobytes()).hexdigest() == gold["dec_md5"]
    else:
        assert np.array_equal(rec, a)


@pytest.mark.skipif(not szref.have_ref(), reason="oracle/_ref/libSZ_ref.so not built (needs /root/reference)")
@pytest.mark.parametrize("case", cases.CASES[::3], ids=cases.CASE_IDS[::3])
def test_oracle_matches_reference_live(orc, case):
    name, g, mode, ab, rl, extra = case
    a = g()
    ref = szref.SZLib(szref.REF_SO)
    cfg = cases.make_cfg(mode, ab, rl, extra)
    ref.init(cfg)
    orc.init(cfg)
    s_ref = ref.compress(a, mode, ab, rl)
    s_orc, tr = orc.compress(a, mode, ab, rl, trace=True)
    kind = GOLD[name]["kind"]
    assert _norm(s_orc, kind) == _norm(s_ref, kind)
    if a.size > 20:
        d_ref = ref.decompress(s_ref, a.dtype, a.shape)
        d_orc = orc.decompress(s_ref, a.dtype, a.shape)
        assert d_ref.tobytes() == d_orc.tobytes()
        if kind == "regression" and not extra.get("protectValueRange"):
            # compressor-side reconstruction == decompressor output, and within the bound
            assert tr["recon"].reshape(a.shape).tobytes() == d_ref.tobytes()
            assert np.max(np.abs(d_ref.astype(np.float64) - a.astype(np.float64))) <= tr["eb"]
    ref.finalize()


@pytest.mark.skipif(not szref.have_ref(), reason="needs the compiled reference")
def test_filter_dims_matches_reference(orc):
    import ctypes as C
    ref = szref.SZLib(szref.REF_SO)
    ref.lib.filterDimension.argtypes = [C.c_size_t] * 5 + [C.POINTER(C.c_size_t)]
    ref.lib.filterDimension.restype = C.c_int
    vals = [0, 1, 2, 3]
    import itertools
    for dims in itertools.product(vals, repeat=5):
        r5, r4, r3, r2, r1 = dims
        # only well-formed dimension lists: zeros only as a suffix (r1 first)
        seq = [r1, r2, r3, r4, r5]
        nz = [v != 0 for v in seq]
        if nz != sorted(nz, reverse=True) or r1 == 0:
            continue
        out = (C.c_size_t * 5)()
        ref.lib.filterDimension(r5, r4, r3, r2, r1, out)
        _, mine = orc.filter_dims(r5, r4, r3, r2, r1)
        n_ref = [v for v in list(out)]
        # compare the effective dimension lists (what computeDataLength/dispatch see)
        def eff(lst):
            o = []
            for v in lst:
                if v == 0:
                    break
                o.append(v)
            return o
        assert eff(n_ref) == eff(mine), (dims, n_ref, mine)


def test_huffman_roundtrip_distributions(orc):
    # test ideas from the reference's stale test_Huffman.c:17-193 (uniform / linear / extreme; 256 and 65536 states)
    rng = np.random.default_rng(5)
    for states in (256, 65536):
        for kind in ("uniform", "linear", "extreme", "single"):
            n = 20000
            if kind == "uniform":
                s = rng.integers(0, states, n)
            elif kind == "linear":
                p = np.arange(1, states + 1, dtype=np.float64)
                s = rng.choice(states, n, p=p / p.sum())
            elif kind == "extreme":
                s = np.where(rng.random(n) < 0.999, states // 2, rng.integers(0, states, n))
            else:
                s = np.full(n, 7)
            s = s.astype(np.int32)
            tree, nc, code, clen = orc.huffman_build(s, states)
            bits = orc.huffman_encode(s, code, clen)
            import ctypes as C
            out = np.zeros(n, dtype=np.int32)
            buf = np.frombuffer(bits + b"\0" * 8, dtype=np.uint8)
            tb = np.frombuffer(tree, dtype=np.uint8)
            rc = orc.lib.szo_huffman_decode(tb.ctypes.data, nc, buf.ctypes.data, n, out.ctypes.data)
            assert rc == 0
            assert np.array_equal(out, s)
            assert len(bits) == (int(clen[s].astype(np.int64).sum()) + 7) // 8
