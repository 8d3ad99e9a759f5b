#!/usr/bin/env python
"""gpurun_out/ (scratch) -> profiles/ (tracked) after scripts/gpu_evidence_r2.sh: bench lines, the launch list
and its per-kernel shares, one summary per ncu capture (selected metrics incl. the tensor sub-pipes, stall
samples by SASS region), the DRAM traffic files bench.py reads, and a SASS census of the built library."""
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
for f in ("r2_bench_n1.json", "r2_bench_reference_n1.json", "r2_bench_fast_n1.json", "r2_launches_bench.csv", "r2_configs_n1.jsonl",
          "r2_l2_add_bench.txt"):
    if os.path.exists(os.path.join(G, f)):
        shutil.copy(os.path.join(G, f), os.path.join(P, f))

ll = os.path.join(G, "r2_launches_bench.csv")
if os.path.exists(ll):
    rows = list(csv.reader(l for l in open(ll) if l.startswith('"')))
    ix = {h: i for i, h in enumerate(rows[0])}
    agg = defaultdict(lambda: [0, 0.0])
    for r in rows[1:]:
        if r[ix["Metric Name"]] != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", r[ix["Kernel Name"]]).replace("fmlk::", "").replace("void ", "")
        ns = float(r[ix["Metric Value"]].replace(",", "")) * {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}.get(r[ix["Metric Unit"]], 1)
        agg[name][0] += 1
        agg[name][1] += ns
    with open(os.path.join(P, "r2_bench_launch_list_summary.csv"), "w") as f:
        f.write("# launch list of `python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu --no-north-star --no-extra`\n"
                "# (ncu --metrics gpu__time_duration.sum --clock-control none): the FP64 fits, then the fast-mode fits.\n"
                "# Times under ncu are cold-cache and serialised: compare shares, not absolutes.\n")
        f.write("kernel,launches,total_ns\n")
        for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write("%s,%d,%d\n" % (k, n, ns))

PATS = [r"gpu__time_duration\.sum", r"dram__bytes_(read|write)\.sum$", r"dram__bytes_read.sum.per_second", r"launch__registers_per_thread",
        r"sm__throughput.avg.pct", r"smsp__issue_active.avg.pct", r"sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        r"sm__pipe_tensor_subpipe_(dmma|hmma)_cycles_active.avg.pct", r"sm__inst_executed_pipe_(tensor|fp64|uniform)\b.*sum$",
        r"smsp__inst_executed.sum$", r"l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum", r"lts__t_sector_op_read_hit_rate",
        r"smsp__average_warps?_issue_stalled.*_per_issue_active", r"launch__occupancy_limit", r"sm__cycles_elapsed.max$"]
names = {"c2_train_f64": "assign_kernel<3>, config 2 (10M x 64, k=100), training, FP64 mode",
         "c2_train_fast": "assign_fast_kernel<true>, config 2, training, fast mode",
         "c2_predict_fast": "assign_fast_kernel<false>, config 2 shape, predict, fast mode",
         "c1b_train_f64": "assign_kernel<3>, kmeans-benchmark.json shape (1M x 100, k=10), training, FP64 mode",
         "c1b_predict_f64": "assign_kernel<0>, kmeans-benchmark.json shape, predict, FP64 mode",
         "c4shape_train_f64": "assign_ring_kernel<true,1>, config 4's shape (10M x 64, k=1000), training, FP64 mode",
         "c4shape_train_fast": "assign_fast_kernel<true>, config 4's shape, training, fast mode (streamed centroid terms)",
         "c3shape_predict_f64": "assign_ring_kernel<false,2>, config 3's shape (4M x 128, k=1000), predict, FP64 mode",
         "c5shape_train_f64": "assign_ring_kernel<true,2>, config 5's shape (4M x 128, k=256), training, FP64 mode",
         "c1shape_train_small": "assign_small_kernel<16,false,true>, config 1's shape with 8M rows (d=10, k=2), training: one row per thread",
         "c2_train_ordered": "assign_kernel<3> with FMLK_FLAG_DETERMINISTIC (ordered bulk reductions, one accumulator per warp), config 2, training",
         "knn": "knn_predict_kernel<false>, KnnModel.transform: 65536 rows x 64 against 10000 training points, k=5"}
traffic = {}
for key, title in names.items():
    rep = os.path.join(G, "r2_%s.ncu-rep" % key)
    if not os.path.exists(rep):
        continue
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rr = list(csv.reader(io.StringIO(raw)))
    h, u, v = rr[0], rr[1], rr[2]

    def get(name):
        i = h.index(name)
        return float(v[i].replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u[i], 1)

    with open(os.path.join(P, "r2_%s_ncu_summary.txt" % key), "w") as f:
        f.write("# %s\n# ncu --set full --clock-control none --import-source on, one launch (scripts/gpu_evidence_r2.sh)\n" % title)
        f.write("== kernel: %s grid %s block %s\n" % (v[h.index("Kernel Name")][:70], v[h.index("Grid Size")], v[h.index("Block Size")]))
        for hh, uu, vv in zip(h, u, v):
            if any(re.search(p, hh) for p in PATS):
                f.write("  %-95s %-14s %s\n" % (hh, uu, vv))
        src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
        tmp = os.path.join(G, "src_%s.csv" % key)
        open(tmp, "w").write(src)
        f.write(subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "ncu_regions.py"), tmp], capture_output=True, text=True).stdout)
    traffic[key] = (get("dram__bytes_read.sum"), get("dram__bytes_write.sum"), v[h.index("Kernel Name")])
for key, fname, alg in (("c2_train_f64", "traffic_c2.json", 5120000000), ("c2_train_fast", "traffic_c2_fast.json", 5120000000)):
    if key in traffic:
        rd, wr, kn = traffic[key]
        json.dump({"kernel": kn, "workload": "c2 (10M x 64, k=100), one launch", "dram_bytes_read": int(rd), "dram_bytes_write": int(wr),
                   "dram_bytes_per_launch": int(rd + wr), "algorithmic_bytes_per_launch": alg,
                   "source": "ncu --set full --clock-control none, profiles/r2_%s_ncu_summary.txt" % key},
                  open(os.path.join(P, fname), "w"), indent=1)

# SASS census of the built library: which Blackwell-native instructions each kernel family holds
lib = os.path.join(ROOT, "flink-ml_b200", "lib", "libflinkml_kmeans_cuda.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
census, cur = defaultdict(lambda: defaultdict(int)), None
for line in sass.split("\n"):
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(.*", "", cur).replace("fmlk::", "").replace("void ", "")
        continue
    m = re.search(r"^\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
    if m and cur:
        op = m.group(1)
        for fam in ("UTCHMMA", "UTCBAR", "UTCATOMSWS", "LDTM", "STTM", "UTMALDG", "UBLKCP", "UBLKRED", "DMMA", "SYNCS", "LDGSTS", "F2F.F32.F64"):
            if op.startswith(fam):
                census[cur][fam] += 1
with open(os.path.join(P, "r2_sass_census.txt"), "w") as f:
    f.write("# cuobjdump -sass flink-ml_b200/lib/libflinkml_kmeans_cuda.so, instruction counts per kernel (static)\n"
            "# UTCHMMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTCBAR = tcgen05.commit, UTCATOMSWS = tcgen05.alloc,\n"
            "# UTMALDG = tensor-map TMA load, UBLKCP = cp.async.bulk, UBLKRED = cp.reduce.async.bulk, DMMA = FP64 mma.sync,\n"
            "# SYNCS = mbarrier, LDGSTS = cp.async\n")
    for k in sorted(census):
        f.write("%-60s %s\n" % (k[:60], " ".join("%s=%d" % kv for kv in sorted(census[k].items()))))
print(open(os.path.join(P, "r2_sass_census.txt")).read())
