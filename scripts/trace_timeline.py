#!/usr/bin/env python
"""Reads the event timeline a -DFMLK_TRACE build dumps (FMLK_TRACE_FILE) and prints, for CTA 0:
per-warp mean phase durations, and for each scheduler (warp % 4) the share of time with 0..4 of
its MMA warps inside the multiply phase.  usage: trace_timeline.py FILE N_MMA_WARPS"""
import sys
import numpy as np

path, W = sys.argv[1], int(sys.argv[2])
ev = np.fromfile(path, dtype=np.uint64).reshape(16, -1)
NAMES = {(0, 1): "wait buffer", (1, 2): "wait rows", (2, 3): "multiply+select", (3, 4): "merge/labels/publish",
         (4, 0): "loop", (5, 6): "acc wait", (6, 5): "acc work", (6, 7): "acc copy-in", (7, 8): "acc barrier", (8, 5): "acc rmw"}
intervals = {}
for w in range(16):
    e = ev[w][ev[w] != 0]
    if e.size == 0:
        continue
    t = (e >> np.uint64(4)).astype(np.int64)
    k = (e & np.uint64(15)).astype(np.int64)
    t -= t[0]
    dur = {}
    for i in range(len(k) - 1):
        dur.setdefault((int(k[i]), int(k[i + 1])), []).append(t[i + 1] - t[i])
    skip = 4  # leave the start-up out
    line = " ".join("%s=%.0f" % (NAMES.get(key, str(key)), np.mean(v[skip:])) for key, v in sorted(dur.items()) if len(v) > skip)
    print("warp %2d (%s) events=%d span=%d cycles: %s" % (w, "mma" if w < W else "acc", len(k), t[-1], line))
    if w < W:
        st = t[k == 2]
        en = t[k == 3]
        n = min(len(st), len(en))
        intervals[w] = (ev[w], st[:n] + int((ev[w][0] >> np.uint64(4))), en[:n] + int((ev[w][0] >> np.uint64(4))))
for s in range(4):
    ws = [w for w in intervals if w % 4 == s]
    if not ws:
        continue
    lo = max(intervals[w][1][4] for w in ws)
    hi = min(intervals[w][2][-1] for w in ws)
    pts = []
    for w in ws:
        _, st, en = intervals[w]
        for a, b in zip(st, en):
            a, b = max(a, lo), min(b, hi)
            if b > a:
                pts.append((a, 1))
                pts.append((b, -1))
    pts.sort()
    hist = np.zeros(len(ws) + 1)
    cur, last = 0, lo
    for tt, dlt in pts:
        hist[cur] += tt - last
        last = tt
        cur += dlt
    hist[cur] += hi - last
    print("scheduler %d (warps %s): share of time with n warps multiplying: %s" % (
        s, ws, " ".join("%d:%.1f%%" % (i, 100 * h / hist.sum()) for i, h in enumerate(hist))))
