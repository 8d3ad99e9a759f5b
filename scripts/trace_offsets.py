#!/usr/bin/env python
"""Start offsets of the multiply phases of the MMA warps of scheduler 0 (CTA 0), tile by tile.
usage: trace_offsets.py FILE N_MMA_WARPS"""
import sys
import numpy as np
ev = np.fromfile(sys.argv[1], dtype=np.uint64).reshape(16, -1)
W = int(sys.argv[2])
ws = [w for w in range(W) if w % 4 == 0]
S = {}
t0 = int(ev[0][0] >> np.uint64(4))
for w in ws:
    e = ev[w][ev[w] != 0]
    t = (e >> np.uint64(4)).astype(np.int64) - t0
    k = (e & np.uint64(15)).astype(np.int64)
    S[w] = (t[k == 2], t[k == 3])
n = min(len(S[w][0]) for w in ws)
for i in list(range(0, 12)) + list(range(20, n, 20)):
    print("tile %3d: start %s  dur %s" % (i, [int(S[w][0][i] - S[ws[0]][0][i]) for w in ws],
                                          [int(S[w][1][i] - S[w][0][i]) for w in ws]))
