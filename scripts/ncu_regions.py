"""Aggregates the warp-stall samples of an ncu source page by SASS region.
python scripts/ncu_regions.py source.csv   (csv from: ncu -i rep --page source --csv)
A region is a maximal run of instructions between branch targets/branches; regions that contain
DMMA are 'loop' regions.  Prints samples and the top stall reasons per region."""
import csv
import sys
from collections import Counter

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
data = rows[2:]
regions, cur = [], []
for r in data:
    src = r[ix["Source"]]
    cur.append(r)
    if any(t in src for t in ("BRA", "EXIT", "RET", "CALL", "BSYNC", "WARPSYNC")):
        regions.append(cur)
        cur = []
if cur:
    regions.append(cur)
tot = sum(int(r[ix["# Samples"]] or 0) for r in data)
print("total samples", tot)
out = []
for reg in regions:
    n = sum(int(r[ix["# Samples"]] or 0) for r in reg)
    if n < tot * 0.004:
        continue
    st = Counter()
    for r in reg:
        for c in stall_cols:
            v = r[ix[c]]
            if v:
                st[c] += int(v)
    dm = sum("DMMA" in r[ix["Source"]] for r in reg)
    ex = sum(int(r[ix["Instructions Executed"]] or 0) for r in reg)
    out.append((reg[0][ix["Address"]], len(reg), dm, n, ex, st.most_common(4)))
for a, ln, dm, n, ex, st in out:
    print("%s len=%4d dmma=%3d samples=%6d (%4.1f%%) inst=%9d  %s" % (a[-5:], ln, dm, n, 100.0 * n / tot, ex,
          " ".join("%s=%d" % (k[6:], v) for k, v in st)))
