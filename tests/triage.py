"""GPU parity triage (run on the B200 box): for each case, compare the CUDA path with the CPU oracle STAGE BY STAGE
and print where they first diverge.  Test infrastructure (imports the oracle); never part of the product path.

    python tests/triage.py [substring-of-case-name ...]
"""
import os
import sys
import time
import traceback

import numpy as np

os.environ.setdefault("SZB200_DEBUG_RECON", "1")
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))
import cases  # noqa: E402
import szref  # noqa: E402
import sz_b200  # noqa: E402
from sz_b200 import api  # noqa: E402


def axis(n, B):
    nb = 1 if n <= B else n // B
    late = n // nb
    split = n % nb
    early = late + (1 if split else 0)
    coords = np.arange(n)
    cut = split * early
    b = np.where(coords < cut, coords // early, split + (coords - cut) // late)
    inner = np.where(coords < cut, coords - b * early, (coords - cut) - (b - split) * late)
    edge = np.where(b < split, early, late)
    off = np.where(b < split, b * early, b * late + split)
    return dict(nb=nb, b=b, inner=inner, edge=edge, off=off)


def bm_index(n0, n1, n2, dim):
    """block-major position p of every raster point (SURVEY.md Appendix B / sz_float.c:7064)."""
    B = 6 if dim == 3 else 16
    a0 = axis(n0, B) if dim == 3 else axis(1, 1)
    a1, a2 = axis(n1, B), axis(n2, B)
    sx = a0["edge"][:, None, None].astype(np.int64); sy = a1["edge"][None, :, None].astype(np.int64); szz = a2["edge"][None, None, :].astype(np.int64)
    X = a0["off"][:, None, None].astype(np.int64); Y = a1["off"][None, :, None].astype(np.int64); Z = a2["off"][None, None, :].astype(np.int64)
    ii = a0["inner"][:, None, None].astype(np.int64); jj = a1["inner"][None, :, None].astype(np.int64); kk = a2["inner"][None, None, :].astype(np.int64)
    s0 = n1 * n2
    p = X * s0 + sx * (Y * n2 + sy * Z) + (ii * sy + jj) * szz + kk
    return p.reshape(-1)


def first_diff(a, b):
    a = np.asarray(a).reshape(-1); b = np.asarray(b).reshape(-1)
    n = min(a.size, b.size)
    d = np.nonzero(a[:n].view(np.uint8 if a.dtype.kind == "V" else a.dtype) != b[:n])[0] if a.dtype.kind != "f" else np.nonzero(a[:n].view("u%d" % a.itemsize) != b[:n].view("u%d" % b.itemsize))[0]
    return (int(d[0]), int(d.size)) if d.size else (None, 0)


def run_case(case, orc):
    name, g, mode, ab, rl, extra = case
    a = g()
    cfg = cases.make_cfg(mode, ab, rl, extra)
    orc.init(cfg)
    so, tr = orc.compress(a, mode, ab, rl, trace=True)
    sz_b200.init(cfg)
    t0 = time.time()
    try:
        sg = sz_b200.compress(a, mode, ab, rl)
    except Exception as e:  # noqa: BLE001
        print("  [%s] GPU compress FAILED: %s" % (name, e))
        return False
    dt = time.time() - t0
    ok = (sg == so) or (tr["kind"] == 1 and sg[:19] == so[:19] and sg[20:] == so[20:])
    st = sz_b200.last_stats()
    print("[%s] %s %s kind=%d  oracle=%d B  gpu=%d B  stream %s  (%.1f ms wall)" % (name, a.dtype, a.shape, tr["kind"], len(so), len(sg), "EQ" if ok else "DIFF", dt * 1e3))
    if not ok and tr["kind"] in (0, 2) and st is not None:
        dim = tr["dim"]; n0, n1, n2 = tr["n0"], tr["n1"], tr["n2"]
        ncoef = 4 if dim == 3 else 3
        nb = tr["num_blocks"]; N = tr["N"]
        print("   header: oracle %s | gpu %s" % (so[:4 + 36 + 8 + 28].hex()[:160], sg[:4 + 36 + 8 + 28].hex()[:160]))
        print("   intervals o=%d g=%d | use_mean o=%d g=%d | mean o=%r g=%r | reg_count o=%d g=%d | unpred o=%d g=%d | tree o=%d/%d g=%d/%d | bits o=%d g=%d" % (
            tr["intervals"], st.intervals, tr["use_mean"], st.use_mean, tr["mean"], st.mean, tr["reg_count"], st.reg_count, tr["total_unpred"], st.total_unpred,
            tr["main_tree_bytes"], tr["main_node_count"], st.tree_bytes, st.node_count, tr["main_bits_bytes"], st.type_bits_bytes))
        try:
            reg = api.debug_fetch(0, a.dtype, ncoef * nb).reshape(ncoef, nb)
            fd = first_diff(reg, tr["reg_params"]); print("   reg_params first diff (flat idx, count):", fd)
            ind = api.debug_fetch(1, np.uint8, nb)
            fd = first_diff(ind, tr["indicator"]); print("   indicator  first diff:", fd)
            if tr["reg_count"]:
                rc = tr["reg_count"]
                ct = api.debug_fetch(5, np.uint16, ncoef * nb).reshape(ncoef, nb)[:, :rc]
                fd = first_diff(ct.astype(np.int32), tr["coeff_type"]); print("   coeff_type first diff:", fd)
                cl = api.debug_fetch(2, a.dtype, ncoef * nb).reshape(ncoef, nb)
                regmask = tr["indicator"] == 0
                fd = first_diff(cl[:, regmask], tr["coeff_last"][:, regmask]); print("   coeff_last first diff:", fd)
            codes = api.debug_fetch(3, np.uint16, N)
            p = bm_index(n0, n1, n2, dim)
            o_raster = tr["type"][p]           # oracle codes in raster order
            fd = first_diff(codes.astype(np.int32), o_raster.astype(np.int32))
            print("   codes      first diff (raster idx, count):", fd)
            if fd[0] is not None:
                i = fd[0]; k = i % n2; j = (i // n2) % n1; ii = i // (n1 * n2)
                print("      at (i,j,k)=(%d,%d,%d): gpu=%d oracle=%d  x=%r" % (ii, j, k, codes[i], o_raster[i], a.reshape(-1)[i]))
            rec = api.debug_fetch(4, a.dtype, N)
            fd = first_diff(rec, tr["recon"].reshape(-1)); print("   recon      first diff:", fd)
        except Exception:  # noqa: BLE001
            traceback.print_exc()
        n = min(len(sg), len(so))
        d = [i for i in range(n) if sg[i] != so[i]][:8]
        print("   first differing stream offsets:", d)
    # decode checks (GPU decodes the ORACLE stream; oracle decodes the GPU stream when it matched)
    if a.size > 20:
        try:
            dg = sz_b200.decompress(so, a.dtype, a.shape)
            do = orc.decompress(so, a.dtype, a.shape)
            dok = dg.tobytes() == do.tobytes()
            std = sz_b200.last_stats()
            print("   decompress(oracle stream): %s" % ("EQ" if dok else "DIFF"), "" if dok else "first diff %s" % (first_diff(dg, do),),
                  "" if std is None else "| %.2f ms device" % std.ms_total)
            ok = ok and dok
        except Exception as e:  # noqa: BLE001
            print("   GPU decompress FAILED: %s" % e)
            ok = False
    if st is not None:
        print("   compress device time %.3f ms, launches %d, stages: %s" % (st.ms_total, st.launches, ", ".join("%s=%.3f" % (k, v) for k, v in st.as_dict()["ms_stage"].items() if v > 0)))
    sz_b200.finalize()
    return ok


def main():
    filt = sys.argv[1:]
    orc = szref.Oracle()
    sz_b200.set_device(0)
    bad = []
    for case in cases.CASES:
        if filt and not any(f in case[0] for f in filt):
            continue
        try:
            if not run_case(case, orc):
                bad.append(case[0])
        except Exception:  # noqa: BLE001
            traceback.print_exc()
            bad.append(case[0])
        sys.stdout.flush()
    print("TRIAGE SUMMARY: %d failing: %s" % (len(bad), bad))


if __name__ == "__main__":
    main()
