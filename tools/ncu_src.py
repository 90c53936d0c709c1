"""Summarise an ncu '--page source --csv' export: top SASS lines by stall samples and by executed instructions.
usage: python tools/ncu_src.py file.csv [topN]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
ia, isrc, isamp, iex = hdr.index("Address"), hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")


def num(v):
    try:
        return int(float(v))
    except ValueError:
        return 0


data = [(r[ia], r[isrc], num(r[isamp]), num(r[iex])) for r in rows[hi + 1:] if len(r) > iex and r[ia].startswith("0x")]
ts, te = sum(d[2] for d in data), sum(d[3] for d in data)
print("total samples", ts, "total warp-inst", te, "sass lines", len(data))
print("--- top by samples (idx, samples, %, executed, sass)")
for i, d in sorted(enumerate(data), key=lambda x: -x[1][2])[:top]:
    print(i, d[2], "%.1f%%" % (100.0 * d[2] / max(ts, 1)), d[3], d[1][:100])
print("--- top by executed")
for i, d in sorted(enumerate(data), key=lambda x: -x[1][3])[:top // 2]:
    print(i, d[2], d[3], "%.1f%%" % (100.0 * d[3] / max(te, 1)), d[1][:100])
