"""Hot-code footprint per source line: joins `nvdisasm -g` output of one kernel with the per-instruction
execution counts of an ncu source page (same build).
python scripts/hot_lines.py all.sass KERNEL_MANGLED source.csv [min_exec]"""
import csv
import re
import sys
from collections import Counter

sass, kern, src = sys.argv[1:4]
min_exec = float(sys.argv[4]) if len(sys.argv) > 4 else 1e5
lines = open(sass).read().split("\n")
start = next(i for i, l in enumerate(lines) if l.startswith(".text." + kern + ":"))
cur, instr = None, []
for l in lines[start + 1:]:
    if l.startswith(".text.") or l.startswith("\t.section") and instr:
        break
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)), "inlined" in m.group(3))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        instr.append((int(m.group(1), 16), m.group(2), cur))
rows = list(csv.reader(open(src)))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
data = rows[2:]
print("sass instrs", len(instr), "profile instrs", len(data))
n = min(len(instr), len(data))
per_line, per_line_samples = Counter(), Counter()
for (addr, text, loc), r in zip(instr[:n], data[:n]):
    e = int(r[ix["Instructions Executed"]] or 0)
    s = int(r[ix["# Samples"]] or 0)
    if e >= min_exec:
        per_line[loc[:2] if loc else None] += 1
    per_line_samples[loc[:2] if loc else None] += s
tot = sum(per_line.values())
print("hot instrs", tot)
for loc, c in sorted(per_line.items(), key=lambda kv: (kv[0] is None, kv[0])):
    print("%-22s %5d hot instrs  %6d samples" % ("%s:%d" % loc if loc else "?", c, per_line_samples[loc]))
