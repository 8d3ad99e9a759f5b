#!/usr/bin/env python
"""Copies the evidence of the last scripts/gpu_bench_profile.sh + measure_configs.py run from
gpurun_out/ (scratch) into profiles/ (tracked): bench lines, launch list and its summary, the DRAM
traffic of the dominant kernel, the ncu summary with the stall samples by SASS region."""
import csv
import io
import json
import os
import re
import shutil
import subprocess
import sys
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
for src, dst in (("bench_full.json", "r1_bench_n1.json"), ("bench_ref.json", "r1_bench_reference_n1.json"),
                 ("launches_bench.csv", "r1_bench_launches.csv"), ("measure_configs.jsonl", "r1_configs_n1.jsonl")):
    if os.path.exists(os.path.join(G, src)):
        shutil.copy(os.path.join(G, src), os.path.join(P, dst))

rows = list(csv.reader(l for l in open(os.path.join(G, "launches_bench.csv")) if l.startswith('"')))
ix = {h: i for i, h in enumerate(rows[0])}
agg = defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    if r[ix["Metric Name"]] != "gpu__time_duration.sum":
        continue
    name = re.sub(r"\(.*", "", r[ix["Kernel Name"]]).replace("fmlk::", "")
    ns = float(r[ix["Metric Value"]].replace(",", "")) * {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}.get(r[ix["Metric Unit"]], 1)
    agg[name][0] += 1
    agg[name][1] += ns
step = [k for k in agg if not any(t in k for t in ("peak", "generate", "elementwise", "vectorized", "fill", "copy"))]
tot = sum(agg[k][1] for k in step)
with open(os.path.join(P, "r1_bench_launch_list_summary.csv"), "w") as f:
    f.write("# launch list of `python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu` (ncu --metrics gpu__time_duration.sum --clock-control none)\n")
    f.write("# share = share of the training-step kernels (peak microbenchmarks and the data generator run outside the timed region)\n")
    f.write("kernel,launches,total_ns,share_of_step_kernels\n")
    for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write("%s,%d,%d,%s\n" % (k, n, ns, ("%.4f" % (ns / tot)) if k in step else "outside timed region"))

rep = os.path.join(G, "prof_c2full.ncu-rep")
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(io.StringIO(out)))
h, u, v = rr[0], rr[1], rr[2]


def get(name):
    i = h.index(name)
    return float(v[i].replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u[i], 1)


rd, wr = get("dram__bytes_read.sum"), get("dram__bytes_write.sum")
json.dump({"kernel": "fmlk::assign_kernel<1> (assign + accumulate)", "workload": "c2 (10M x 64, k=100), one launch",
           "dram_bytes_read": int(rd), "dram_bytes_write": int(wr), "dram_bytes_per_launch": int(rd + wr),
           "algorithmic_bytes_per_launch": 5120000000,
           "source": "ncu --set full --clock-control none of `python scripts/run_assign.py 10000000 64 100 3`; "
                     "summary in profiles/r1_assign_c2_ncu_summary.txt"},
          open(os.path.join(P, "traffic_c2.json"), "w"), indent=1)
with open(os.path.join(P, "r1_assign_c2_ncu_summary.txt"), "w") as f:
    f.write(subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "ncu_summary.py"), rep], capture_output=True, text=True).stdout)
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    tmp = os.path.join(G, "src_c2full.csv")
    open(tmp, "w").write(src)
    f.write(subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "ncu_regions.py"), tmp], capture_output=True, text=True).stdout)
b = json.load(open(os.path.join(P, "r1_bench_n1.json")))
print("bench value %.4g e2e %.4g roofline %.2f %s = %.3f traffic %s" % (
    b["value"], b["e2e"]["value"], b["roofline"]["achieved"], b["roofline"]["unit"], b["roofline"]["frac"], int(rd + wr)))
