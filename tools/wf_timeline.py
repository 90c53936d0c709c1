"""Reads the per-pencil timeline written by the profiling build (SZB200_WF_TIMELINE=path) and prints where the
wavefront's wall time goes: hop time along the anti-diagonals, run time of a pencil once started, idle before start."""
import sys
import numpy as np

path = sys.argv[1]
with open(path) as f:
    pa, pb = map(int, f.readline().split())
    t = np.array([list(map(int, l.split())) for l in f], dtype=np.int64).reshape(pa, pb, 4)
t[..., 1] = np.where(t[..., 1] == 0, t[..., 0], t[..., 1])      # pencils without neighbours never record a first-chunk wait
t0 = t[..., 0].min()
tick, start, end = (t[..., 0] - t0) / 1e3, (t[..., 1] - t0) / 1e3, (t[..., 2] - t0) / 1e3   # us
print("pencil grid %d x %d; kernel span %.1f us" % (pa, pb, end.max()))
d = np.add.outer(np.arange(pa), np.arange(pb))
ds = np.arange(d.max() + 1)
first = np.array([start[d == k].min() for k in ds]); last = np.array([start[d == k].max() for k in ds])
fin = np.array([end[d == k].max() for k in ds])
print("start of diagonal d (earliest pencil), us:", np.round(first[::19], 1))
print("mean hop (start[d+1]-start[d]) %.2f us; median %.2f; p90 %.2f" % (np.diff(first).mean(), np.median(np.diff(first)), np.percentile(np.diff(first), 90)))
run = end - start
print("run time of a pencil after its first chunk was released: mean %.1f us, min %.1f, p50 %.1f, p90 %.1f, max %.1f" % (run.mean(), run.min(), np.median(run), np.percentile(run, 90), run.max()))
idle = start - tick
print("idle between ticket and first chunk: mean %.1f us, p50 %.1f, max %.1f" % (idle.mean(), np.median(idle), idle.max()))
# direct hop: start of (a,b) minus max(start of up, start of left)
hop = np.full((pa, pb), np.nan)
for a in range(pa):
    for b in range(pb):
        p = []
        if a: p.append(start[a - 1, b])
        if b: p.append(start[a, b - 1])
        if p: hop[a, b] = start[a, b] - max(p)
h = hop[~np.isnan(hop)]
print("start lag behind the later predecessor: mean %.2f us, p10 %.2f, p50 %.2f, p90 %.2f" % (h.mean(), np.percentile(h, 10), np.median(h), np.percentile(h, 90)))
late = (tick > np.where(np.isnan(hop), 0, start - hop))
print("pencils whose ticket was taken AFTER their predecessors had started: %d of %d" % (late.sum(), pa * pb))
