#!/usr/bin/env python
"""Reads the event timeline a -DFMLK_TRACE build dumps for assign_fast_kernel (FMLK_TRACE_FILE) and prints,
for every warp of CTA 0, the mean duration of each phase between consecutive events (cycles), steady state.
usage: fast_timeline.py FILE"""
import sys
import numpy as np

NAMES = {(0, 1): "wait rows", (1, 2): "wait A free", (2, 3): "convert+st", (3, 4): "thr+arrive", (4, 0): "loop",
         (5, 6): "wait thr+scores", (6, 7): "ld+top2", (7, 8): "flags+reduce issue", (8, 9): "wait reduce read",
         (7, 9): "flags+labels", (9, 5): "loop", (10, 11): "wait stage free", (11, 10): "issue copies",
         (12, 13): "wait A", (13, 14): "wait D free", (14, 15): "issue mma", (15, 12): "loop"}
ROLE = ["cnv0"] * 4 + ["cnv1"] * 4 + ["epi0"] * 4 + ["epi1"] * 4 + ["load", "mma", "cent"]
NAMES.update({(5, 6): "wait thr + first scores", (6, 6): "one pass: ld+top2 (+ wait next)", (6, 7): "last pass: ld+top2", (13, 14): "wait B/D", (14, 14): "issue pass (+ wait next B/D)", (14, 15): "issue last pass", (12, 13): "wait A", (7, 9): "merge+flags+labels", (7, 8): "merge+flags+reduce issue", (11, 10): "issue copies (half)"})
ev = np.fromfile(sys.argv[1], dtype=np.uint64).reshape(len(ROLE), -1)
for w in range(ev.shape[0]):
    e = ev[w][ev[w] != 0]
    if e.size < 8:
        continue
    t = (e >> np.uint64(5)).astype(np.int64)
    k = (e & np.uint64(31)).astype(np.int64)
    dur = {}
    for i in range(len(k) - 1):
        dur.setdefault((int(k[i]), int(k[i + 1])), []).append(t[i + 1] - t[i])
    skip = 3
    first = int(k[0])
    per_tile = np.diff(t[k == first])
    line = "  ".join("%s=%.0f" % (NAMES.get(key, str(key)), np.mean(v[skip:])) for key, v in sorted(dur.items()) if len(v) > skip)
    print("warp %2d %-4s tiles=%d cycles/tile=%.0f | %s" % (w, ROLE[w], (k == first).sum(), np.mean(per_tile[skip:]), line))
