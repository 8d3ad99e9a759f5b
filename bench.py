#!/usr/bin/env python
"""bench.py -- KMeans point-iterations/sec on B200 (BASELINE.json's metric).

A step is one pass of the hot path over one batch of synthetic input: one KMeans.fit of
maxIter rounds over the rows resident in HBM (DenseVectorGenerator data, seed 2; initial
centroids from the reference sampler with the KMeans default seed).

  python bench.py --gpus 1 --steps K --warmup W            # our arm (one rank per GPU under torchrun)
  python bench.py --impl reference --steps K --warmup W    # the reference's CPU path (oracle port)

Workloads (--workload; default c2 = BASELINE.json configs[1], the config the metric is quoted on
that fits one GPU).  With N > 1 every rank holds the same number of rows (weak scaling) and the
per-round k*d sums + k counts go through one ncclAllReduce.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (rows per GPU, d, k, rounds, description)
    "c1": (10_000, 10, 2, 20, "benchmark-demo.json KMeans-1: 10k x 10, k=2, maxIter=20"),
    "c1b": (1_000_000, 100, 10, 10, "kmeans-benchmark.json: 1M x 100, k=10, maxIter=10"),
    "c2": (10_000_000, 64, 100, 20, "KMeans.fit 10M x 64 FP64, k=100, maxIter=20"),
    "c4": (125_000_000, 64, 1000, 20, "KMeans.fit 125M x 64 per GPU (1B over 8), k=1000, maxIter=20"),
    "c4s": (10_000_000, 64, 1000, 20, "KMeans.fit 10M x 64 per GPU, k=1000, maxIter=20 (config-4 shape, short)"),
}
DATA_SEED = 2
METRIC = "KMeans point-iterations/sec"
UNIT = "point-iterations/s"


def read_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_baseline(d, k, seconds_target=12.0, threads=None):
    """The oracle (a C port of the reference's CPU arithmetic; no JVM exists on this box) timed on
    the host cores over a bounded sample of the same workload: one training round over `rows`
    DenseVectorGenerator rows, rows split over the threads like Flink subtasks."""
    import numpy as np
    from oracle import oracle as O
    threads = threads or O.max_threads()
    probe = 20_000 * threads
    x = O.generate_dense(DATA_SEED, 0, 1, probe, d)
    cent = O.select_random_centroids([x], k, O.KMEANS_DEFAULT_SEED)
    O.train_round_mt(x[:2000], cent, threads)
    t0 = time.perf_counter()
    O.train_round_mt(x, cent, threads)
    rate = probe / (time.perf_counter() - t0)
    # bounded sample: at most 4M rows (2 GB at d=64), repeated for enough rounds to fill the target time
    rows = int(min(max(rate * seconds_target, probe), 4_000_000))
    rounds = max(1, int(round(rate * seconds_target / rows)))
    if rows > probe:
        x = O.generate_dense(DATA_SEED, 0, 1, rows, d)
    t0 = time.perf_counter()
    for _ in range(rounds):
        O.train_round_mt(x[:rows], cent, threads)
    dt = time.perf_counter() - t0
    return {"value": rows * rounds / dt, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": "oracle (C restatement of the reference's Java loop, not the JVM): %d rounds over %d rows x %d,"
                      " k=%d, %d threads, %.1f s" % (rounds, rows, d, k, threads, dt)}, x[:rows], cent


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path.  The JVM reference
    cannot run here (no JDK on the box), so this is the oracle port on all host cores; each step is
    a bounded sample (one round over `rows` rows of the workload)."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    import numpy as np
    from oracle import oracle as O
    n, d, k, rounds, desc = WORKLOADS[args.workload]
    threads = O.max_threads()
    base, x, cent = cpu_baseline(d, k, seconds_target=3.0, threads=threads)
    rows = x.shape[0]
    for _ in range(args.warmup):
        O.train_round_mt(x, cent, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        O.train_round_mt(x, cent, threads)
    dt = time.perf_counter() - t0
    value = rows * args.steps / dt
    base = dict(base, value=value,
                sample="oracle port, %d threads: each step = 1 round over %d rows x %d, k=%d" % (threads, rows, d, k))
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": desc, "sample_rows_per_step": rows, "rounds_per_step": 1, "k": k, "d": d},
        "cpu_baseline": base,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def bind_near_gpu(gpu_index):
    """Pins this process to the CPUs NVML lists as local to the GPU (its NUMA node / PCIe root), so the pinned host
    buffers allocated afterwards sit on that node.  Returns what was done, for the e2e record."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1}
        allowed = os.sched_getaffinity(0)
        cpus &= allowed
        if not cpus:
            return "NVML affinity set is empty; left at %d CPUs" % len(allowed)
        if cpus != allowed:
            os.sched_setaffinity(0, cpus)
        return "%d CPUs local to GPU %d (%d..%d) of %d" % (len(cpus), gpu_index, min(cpus), max(cpus), len(allowed))
    except Exception as e:  # no NVML, no permission: the run goes on unbound
        return "not bound: %s" % e


def run_ours(args):
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    affinity = bind_near_gpu(local)  # before anything allocates pinned memory

    import numpy as np
    import torch
    import torch.distributed as dist
    import flink_ml_b200 as F
    from flink_ml_b200 import kmeans as K
    from flink_ml_b200 import native

    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus %d needs torchrun with --nproc-per-node %d" % (args.gpus, args.gpus))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    host_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        host_group = dist.new_group(backend="gloo")

    # ---- N > 1: every world > 1 code path against the CPU oracle at parallelism = world, before
    # anything is timed (p-shard fit over the NCCL all-reduce, labels, the rank-ordered online merge,
    # allReduceSum).  The oracle is the checker here, never the thing measured.  A mismatch ends the
    # run with a non-zero exit code, so the driver's per-N record carries it.
    parity = None
    if world > 1 and not args.no_parity:
        sys.path.insert(0, os.path.join(ROOT, "scripts"))
        from multi_gpu_parity import run_checks
        pc = K.Cluster(rank, world, local, host_group=host_group)
        parity = run_checks(pc, host_group)
        pc.close()
        if not parity["ok"]:
            if rank == 0:
                print(json.dumps({"metric": METRIC, "value": None, "unit": UNIT, "n_gpus": world, "parity": parity,
                                  "error": "world > 1 parity against the oracle failed; nothing was timed"}))
            dist.barrier()
            dist.destroy_process_group()
            sys.exit(1)

    n, d, k, rounds, desc = WORKLOADS[args.workload]
    if args.rows:
        n = args.rows
    stream = torch.cuda.current_stream()
    peaks, peak_src = read_peaks()
    dmma_peak, dfma_peak = native.measure_fp64_peaks(local, 300.0)
    if rank == 0:
        # the FP64 denominators of this run, with the clock they were measured at (MEASURED_PEAKS.json has no FP64 entry)
        try:
            with open(os.path.join(ROOT, "profiles", "fp64_peaks.json"), "w") as f:
                json.dump({"dmma_m8n8k4_tflops": dmma_peak, "dfma_tflops": dfma_peak, "gpu": torch.cuda.get_device_name(local),
                           "how": "fmlk_measure_fp64_peaks: register-resident DMMA / DFMA loops, best of the repeats in 300 ms",
                           "theoretical_dmma_tflops_at_1965MHz": 148 * 64 * 2 * 1.965e9 / 1e12}, f, indent=1)
        except OSError:
            pass

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def generate(rows_per_gpu, dd):
        total = rows_per_gpu * world
        assert native.rows_on_task(total, rank, world) == rows_per_gpu
        t = torch.empty((rows_per_gpu, dd), dtype=torch.float64, device=dev)
        native.generate_dense_dev(local, stream.cuda_stream, DATA_SEED, rank, world, total, dd, t.data_ptr())
        torch.cuda.synchronize()
        return t

    def timed_fits(flags, xt, kk, nrounds, steps, warmup, sample_clocks=False):
        """`steps` KMeans fits of `nrounds` rounds over the device-resident rows xt, timed on the device
        (CUDA events on the launching stream, barrier + synchronize on both sides, max over ranks)."""
        nn, dd = xt.shape
        cl = K.Cluster(rank, world, local, flags=flags | native.FLAG_PROFILE, host_group=host_group)
        cx = cl.context()
        cx.set_stream(stream.cuda_stream)
        cx.adopt_points_dev(xt.data_ptr(), nn, dd, keep=xt)
        est_ = F.KMeans(cl).set_k(kk).set_max_iter(nrounds)
        init_ = K.select_random_centroids(cx, cl, kk, est_.get_seed())

        def one_step():
            cx.set_centroids(init_)
            cx.fit(nrounds, want_model=False)

        for _ in range(warmup):
            one_step()
        cx.reset_profile()
        sampler = ClockSampler(local) if (sample_clocks and rank == 0) else None
        if sampler:
            sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        t0 = time.perf_counter()
        e0.record(stream)
        for _ in range(steps):
            one_step()
        e1.record(stream)
        barrier()
        wall = time.perf_counter() - t0
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        total_ms = float(ms.item())
        prof = cx.profile()
        launches = max(prof["assign_launches"], 1)
        assign_ms = prof["assign_ms"] / launches
        desc_ = cx.describe_launch(kk, dd, True)
        fast = "fast=1" in desc_ and cx.fast_launches() > 0
        tf = 2.0 * kk * dd * nn / (assign_ms * 1e-3) * 1e-12        # 2*k*d flop per point-iteration
        gbs = 8.0 * dd * nn / (assign_ms * 1e-3) * 1e-9             # 8*d bytes per point-iteration
        hbm_frac = gbs / peaks["hbm_gbs"]
        tensor_frac = tf / dmma_peak if dmma_peak > 0 else 0.0
        hbm = {"bound": "hbm", "achieved": gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": hbm_frac,
               "peak_source": peak_src}
        if fast:
            # the contraction is on the tcgen05 pipe (three TF32 passes: 6*k*d flop per point-iteration);
            # nothing FP64 is left on the critical path, so the row stream from HBM is what bounds it
            tf32_peak = peaks.get("bf16_tflops", 0.0) / 2.0
            tf32 = {"bound": "tensor", "achieved": 3.0 * tf, "peak": tf32_peak, "unit": "TFLOP/s",
                    "frac": 3.0 * tf / tf32_peak if tf32_peak else 0.0,
                    "peak_source": "half of MEASURED_PEAKS.json's burst bf16 figure (TF32 runs at half the bf16 rate); "
                                   "flop = 3 TF32 passes x 2*k*d per point-iteration"}
            # few centroids: the row stream from HBM bounds it; many: the three TF32 passes on the tcgen05 pipe
            roof = dict(tf32) if tf32["frac"] > hbm_frac else dict(hbm)
            roof["tf32_tensor"] = tf32
            kern = "fmlk::assign_fast_kernel<true> (tcgen05 TF32x3 contraction, FP64 exact replay of near-tie rows, sums by bulk reductions)"
        elif tensor_frac >= hbm_frac:
            roof = {"bound": "tensor", "achieved": tf, "peak": dmma_peak, "unit": "TFLOP/s", "frac": tensor_frac,
                    "peak_source": "FP64 DMMA m8n8k4 microbenchmark run live by this bench, written to profiles/fp64_peaks.json "
                                   "(MEASURED_PEAKS.json has no FP64 entry; datasheet 37-40 TFLOP/s)"}
        else:
            roof = dict(hbm)
        if not fast:
            kern = ("fmlk::assign_ring_kernel<true> (d-blocked assign + accumulate, centroid blocks through a 4-stage ring)"
                    if "ring=1" in desc_ else
                    "fmlk::assign_wide_kernel<true> (d-blocked assign + accumulate)" if "wide=1" in desc_ else
                    "fmlk::assign_kernel<2> (assign + warp-private sums)" if "private_sums=1" in desc_ else
                    "fmlk::assign_kernel<3> (assign + sums by bulk reductions into one accumulator per CTA)"
                    if "red_sums=1" in desc_ else "fmlk::assign_kernel<1> (assign + accumulate)")
        roof.update({"traffic": None, "kernel": kern, "ms_per_launch": assign_ms, "launches_timed": prof["assign_launches"],
                     "share_of_step": prof["assign_ms"] / total_ms, "hbm": hbm,
                     "fp64_tensor": {"achieved": tf, "peak": dmma_peak, "unit": "TFLOP/s", "frac": tensor_frac},
                     "fp64_fma_peak_tflops": dfma_peak, "launch": desc_})
        model = cx.get_model()
        res = {"value": float(nn) * world * nrounds * steps / (total_ms * 1e-3), "total_ms": total_ms, "wall": wall,
               "prof": prof, "roofline": roof, "model": model, "clocks": sampler.stop() if sampler else None,
               "fast": fast, "ms_per_round": total_ms / steps / nrounds}
        cx.clear_points()
        cl.close()
        return res

    # ---- value: rows resident in HBM, K fits timed on the device
    x = generate(n, d)
    total_rows = n * world
    main_flags = native.FLAG_FAST_TF32 if args.fast else 0
    r = timed_fits(main_flags, x, k, rounds, args.steps, args.warmup, sample_clocks=True)
    value, total_ms, wall, prof, roofline, clocks = r["value"], r["total_ms"], r["wall"], r["prof"], r["roofline"], r["clocks"]
    model_c, model_w = r["model"]
    ncu_traffic = os.path.join(ROOT, "profiles", "traffic_%s%s.json" % (args.workload, "_fast" if r["fast"] else ""))
    if os.path.exists(ncu_traffic):
        try:
            roofline["traffic"] = json.load(open(ncu_traffic)).get("dram_bytes_per_launch")
            roofline["traffic_source"] = "ncu --set full capture of this kernel on this workload, committed as " + \
                                         os.path.relpath(ncu_traffic, ROOT)
        except Exception:
            pass

    # ---- the same workload in the other mode, beside (never instead of) the headline: the tcgen05 fast mode
    fast_rec = None
    if not args.fast and not args.no_fast:
        fr = timed_fits(native.FLAG_FAST_TF32, x, k, rounds, max(1, min(args.steps, 3)), 3)
        fc, fw = fr["model"]
        fast_rec = {
            "dtype": "tf32x3+f64-recheck", "value": fr["value"], "unit": UNIT, "ms_per_round": fr["ms_per_round"],
            "speedup_over_f64_mode": fr["value"] / value, "took_tcgen05_kernel": fr["fast"],
            "roofline": fr["roofline"], "rechecked_rows": fr["prof"]["rechecked_rows"],
            "rechecked_fraction": fr["prof"]["rechecked_rows"] / float(n * rounds * max(1, min(args.steps, 3))),
            "weights_bit_exact_vs_f64_mode": bool(np.array_equal(fw, model_w)),
            "centroids_within_1e-9_of_f64_mode": bool(np.allclose(fc, model_c, rtol=1e-9, atol=0, equal_nan=True)),
            "gpu_launches": fr["prof"]["total_launches"],
        }
        fast_traffic = os.path.join(ROOT, "profiles", "traffic_%s_fast.json" % args.workload)
        if fr["fast"] and os.path.exists(fast_traffic):
            try:
                fast_rec["roofline"]["traffic"] = json.load(open(fast_traffic)).get("dram_bytes_per_launch")
                fast_rec["roofline"]["traffic_source"] = "ncu --set full capture of this kernel on this workload, committed as " + \
                                                         os.path.relpath(fast_traffic, ROOT)
            except Exception:
                pass

    # ---- e2e: the public API with HOST buffers: KMeans.fit(Table) -> model, H2D and D2H inside
    e2e = None

    def e2e_fits(flags, table_, steps_):
        cl = K.Cluster(rank, world, local, flags=flags, host_group=host_group)
        est_ = F.KMeans(cl).set_k(k).set_max_iter(rounds)
        est_.fit(table_)  # warm-up (allocations, page tables)
        barrier()
        t0_ = time.perf_counter()
        for _ in range(steps_):
            model_ = est_.fit(table_)
        barrier()
        dt_ = torch.tensor([time.perf_counter() - t0_], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt_, op=dist.ReduceOp.MAX)
        md_ = F.KMeansModelData.from_table(model_.get_model_data()[0])
        cl.close()
        sec = float(dt_.item())
        return {"value": float(total_rows) * rounds * steps_ / sec, "unit": UNIT,
                "h2d_bytes_per_step": int(n * d * 8 + k * d * 8), "d2h_bytes_per_step": int(k * d * 8 + (k * d + k) * 8),
                "steps": steps_, "ms_per_step": sec / steps_ * 1e3,
                "h2d_gbs_per_rank_if_copy_only": n * d * 8 / 1e9 / (sec / steps_),
                "api": "flink_ml_b200.KMeans.fit(Table) from pinned host rows; timed on the host clock "
                       "around the calls (max over ranks)",
                "model_matches_device_run": bool(np.allclose(md_.centroids, model_c, rtol=1e-9, atol=0, equal_nan=True))}

    if not args.no_e2e:
        host = torch.empty((n, d), dtype=torch.float64, pin_memory=True)
        host.copy_(x)
        torch.cuda.synchronize()
        del x
        torch.cuda.empty_cache()
        table = F.Table({"features": host.numpy()})
        e2e_steps = max(1, min(args.steps, args.e2e_steps))
        e2e = e2e_fits(main_flags, table, e2e_steps)
        # the ceiling of that number: every rank copies 1 GiB of its pinned rows to its GPU at the same time
        probe_bytes = min(n * d * 8, 1 << 30)
        sink = torch.empty(probe_bytes // 8, dtype=torch.float64, device=dev)
        flat = host.view(-1)[:probe_bytes // 8]
        sink.copy_(flat, non_blocking=True)
        barrier()
        t0 = time.perf_counter()
        for _ in range(3):
            sink.copy_(flat, non_blocking=True)
        torch.cuda.synchronize()
        mine = torch.tensor([3 * probe_bytes / 1e9 / (time.perf_counter() - t0)], dtype=torch.float64, device=dev)
        lo, tot = mine.clone(), mine.clone()
        if world > 1:
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        e2e["host_link"] = {"h2d_gbs_per_rank_min": float(lo.item()), "h2d_gbs_all_ranks": float(tot.item()),
                            "how": "3 x %d MiB pinned host-to-device copies per rank, all ranks at once" % (probe_bytes >> 20),
                            "copy_only_ms_per_step": n * d * 8 / 1e9 / float(lo.item()) * 1e3,
                            "cpu_affinity_rank0": affinity}
        del sink
        if fast_rec is not None:
            fast_rec["e2e"] = e2e_fits(native.FLAG_FAST_TF32, table, e2e_steps)
        del host, table
    else:
        del x
        torch.cuda.empty_cache()

    # ---- the other two entry points north_star names, each device-timed and end to end from pinned host rows:
    # KMeansModel.transform (config 3's shape, a bounded number of rows) and OnlineKMeans mini-batches (config 5)
    transform_rec = online_rec = knn_rec = sgd_rec = None
    if not args.no_extra and args.workload == "c2":
        modes = [("f64", 0)] + ([] if args.no_fast else [("tf32x3+f64-recheck", native.FLAG_FAST_TF32)])
        tn, td, tk = args.transform_rows, 128, 1000
        cent = torch.empty((tk, td), dtype=torch.float64, device=dev)
        native.generate_dense_dev(local, stream.cuda_stream, 1, 0, 1, tk, td, cent.data_ptr())  # model data: seed 1
        xt = generate(tn, td)
        host_t = torch.empty((tn, td), dtype=torch.float64, pin_memory=True)
        host_t.copy_(xt)
        torch.cuda.synchronize()
        cent_h = cent.cpu().numpy()
        table_t = F.Table({"features": host_t.numpy()})
        transform_rec = {"workload": "KMeansModel.transform, config 3's shape: %d x %d FP64 rows per GPU, k=%d" % (tn, td, tk),
                         "unit": "rows/s", "modes": {}}
        ref_labels = None
        for name, fl in modes:
            cl = K.Cluster(rank, world, local, flags=fl | native.FLAG_PROFILE, host_group=host_group)
            cx = cl.context()
            cx.set_stream(stream.cuda_stream)
            cx.set_centroids(cent_h)
            lab = torch.empty((tn,), dtype=torch.int32, device=dev)
            for _ in range(3):
                cx.predict_dev(xt.data_ptr(), tn, td, lab.data_ptr())
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            e0.record(stream)
            for _ in range(3):
                cx.predict_dev(xt.data_ptr(), tn, td, lab.data_ptr())
            e1.record(stream)
            barrier()
            ms = torch.tensor([e0.elapsed_time(e1) / 3], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            dev_ms = float(ms.item())
            model = F.KMeansModel(cl).set_k(tk).set_model_data(F.KMeansModelData(cent_h, np.zeros(tk)).to_table())
            model.transform(table_t)  # warm-up
            barrier()
            t0 = time.perf_counter()
            out_t = model.transform(table_t)[0]
            barrier()
            dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(dt, op=dist.ReduceOp.MAX)
            labels_h = np.asarray(out_t.column("prediction"))
            same = bool(np.array_equal(labels_h, lab.cpu().numpy()))
            if ref_labels is None:
                ref_labels = labels_h
            tf_t = 2.0 * tk * td * tn / (dev_ms * 1e-3) * 1e-12
            transform_rec["modes"][name] = {
                "value": tn * world / (dev_ms * 1e-3), "ms_per_launch": dev_ms, "fp64_equivalent_tflops": tf_t,
                "frac_of_fp64_dmma_peak": tf_t / dmma_peak if dmma_peak else None, "launch": cx.describe_launch(tk, td, False),
                "e2e": {"value": tn * world / float(dt.item()), "unit": "rows/s", "ms": float(dt.item()) * 1e3,
                        "h2d_bytes_per_step": int(tn * td * 8 + tk * td * 8), "d2h_bytes_per_step": int(tn * 4),
                        "api": "KMeansModel.transform(Table) from pinned host rows (copy and kernel overlapped on two streams)",
                        "labels_match_device_run": same},
                "labels_bit_exact_vs_f64_mode": bool(np.array_equal(labels_h, ref_labels))}
            cl.close()
        del xt, host_t, table_t, cent
        torch.cuda.empty_cache()

        bn, bd, bk, nbatch = 100_000, 128, 256, args.online_batches
        per_rank = bn // world
        xb = generate(per_rank * nbatch, bd)
        host_b = torch.empty((per_rank * nbatch, bd), dtype=torch.float64, pin_memory=True)
        host_b.copy_(xb)
        torch.cuda.synchronize()
        del xb
        hb = host_b.numpy()
        c0 = torch.empty((bk, bd), dtype=torch.float64, device=dev)
        native.generate_dense_dev(local, stream.cuda_stream, 0, 0, 1, bk, bd, c0.data_ptr())
        c0h = c0.cpu().numpy()
        online_rec = {"workload": "OnlineKMeans mini-batches, config 5: global batch %d x %d (%d rows per GPU), k=%d, decay 0.5, "
                                  "%d batches" % (bn, bd, per_rank, bk, nbatch), "unit": "rows/s", "modes": {}}
        ref_model = None
        for name, fl in modes:
            cl = K.Cluster(rank, world, local, flags=fl, host_group=host_group)
            cx = cl.context()
            c, w = c0h.copy(), np.zeros(bk)
            for b in range(3):
                cx.online_round(hb[b * per_rank:(b + 1) * per_rank], c, w, 0.5)
            c, w = c0h.copy(), np.zeros(bk)
            barrier()
            t0 = time.perf_counter()
            for b in range(nbatch):
                c, w = cx.online_round(hb[b * per_rank:(b + 1) * per_rank], c, w, 0.5)
            barrier()
            dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(dt, op=dist.ReduceOp.MAX)
            if ref_model is None:
                ref_model = (c, w)
            online_rec["modes"][name] = {
                "value": bn * nbatch / float(dt.item()), "ms_per_batch": float(dt.item()) / nbatch * 1e3,
                "h2d_bytes_per_step": int(per_rank * bd * 8 + (bk * bd + bk) * 8), "d2h_bytes_per_step": int((bk * bd + bk) * 8),
                "api": "fmlk_online_round per batch: host batch and model in, merged model out (what OnlineKMeans.fit calls)",
                "weights_bit_exact_vs_f64_mode": bool(np.array_equal(w, ref_model[1])),
                "centroids_within_1e-9_of_f64_mode": bool(np.allclose(c, ref_model[0], rtol=1e-9, atol=0, equal_nan=True))}
            cl.close()
        del host_b, hb
        torch.cuda.empty_cache()

        # KnnModel.transform: every row against every training point, exact FP64, reference order
        qn, qm, qd, qk = args.knn_rows, 10_000, 64, 5
        feats = generate(qm, qd)
        rows_q = generate(qn, qd)
        feats_h = feats.cpu().numpy()
        host_q = torch.empty((qn, qd), dtype=torch.float64, pin_memory=True)
        host_q.copy_(rows_q)
        torch.cuda.synchronize()
        labels_h = (np.arange(qm) % 10).astype(np.float64)
        cl = K.Cluster(rank, world, local, flags=0, host_group=host_group)
        cx = cl.context()
        cx.set_stream(stream.cuda_stream)
        cx.knn_set_model(feats_h, labels_h)
        pred = torch.empty((qn,), dtype=torch.float64, device=dev)
        for _ in range(2):
            cx.knn_predict_dev(rows_q.data_ptr(), qn, qd, qk, pred.data_ptr())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(stream)
        for _ in range(3):
            cx.knn_predict_dev(rows_q.data_ptr(), qn, qd, qk, pred.data_ptr())
        e1.record(stream)
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1) / 3], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        knn_ms = float(ms.item())
        from flink_ml_b200.knn import KnnModel, KnnModelData
        kmodel = KnnModel(cl).set_k(qk).set_model_data(KnnModelData(feats_h, cx.knn_norm_squares(), labels_h).to_table())
        table_q = F.Table({"features": host_q.numpy()})
        kmodel.transform(table_q)
        barrier()
        t0 = time.perf_counter()
        out_q = kmodel.transform(table_q)[0]
        barrier()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        knn_tf = 2.0 * qn * qm * qd / (knn_ms * 1e-3) * 1e-12
        knn_rec = {"workload": "KnnModel.transform: %d x %d FP64 rows per GPU against %d training points, k=%d" % (qn, qd, qm, qk),
                   "unit": "rows/s", "value": qn * world / (knn_ms * 1e-3), "ms_per_launch": knn_ms,
                   "roofline": {"bound": "fp64 pipe (multiply and add kept apart, as the reference rounds them)",
                                "achieved": knn_tf, "peak": dfma_peak / 2.0, "unit": "TFLOP/s",
                                "frac": knn_tf / (dfma_peak / 2.0) if dfma_peak else None,
                                "peak_is": "half the measured DFMA rate: one DMUL + one DADD per 2 flop"},
                   "e2e": {"value": qn * world / float(dt.item()), "unit": "rows/s", "ms": float(dt.item()) * 1e3,
                           "h2d_bytes_per_step": int(qn * qd * 8 + qm * qd * 8 + qm * 16), "d2h_bytes_per_step": int(qn * 8),
                           "api": "KnnModel.transform(Table) from pinned host rows (model data re-sent per call)",
                           "predictions_match_device_run": bool(np.array_equal(np.asarray(out_q.column("prediction")),
                                                                                pred.cpu().numpy()))}}
        cl.close()
        del feats, rows_q, host_q, table_q
        torch.cuda.empty_cache()

        # SGD (allReduceSum's other caller): logisticregression-benchmark.json's conf per GPU (10 M x 100
        # rows resident in HBM, maxIter 20, a mini-batch of 100 000 rows per GPU and epoch), the whole loop on the device
        sn, sd, sb, se = args.sgd_rows, 100, 100_000, 20
        free_b, _ = torch.cuda.mem_get_info(local)
        # (one GPU only: the N > 1 SGD path is covered by the parity leg above, which fits over p shards before anything is timed)
        if world == 1 and free_b > sn * sd * 8 + (4 << 30):
            xs = generate(sn, sd)
            ys = ((xs @ torch.linspace(-1.0, 1.0, sd, dtype=torch.float64, device=dev)) > 0).to(torch.float64)
            cl = K.Cluster(rank, world, local, flags=0, host_group=host_group)
            cx = cl.context()
            cx.set_stream(stream.cuda_stream)
            cx.sgd_set_data_dev(xs.data_ptr(), ys.data_ptr(), None, sn, sd, keep=(xs, ys))
            for _ in range(2):
                coef, ran, last = cx.sgd_fit(native.LOSS_LOGISTIC, max_iter=se, global_batch_size=sb * world, tol=0.0)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            e0.record(stream)
            for _ in range(3):
                coef, ran, last = cx.sgd_fit(native.LOSS_LOGISTIC, max_iter=se, global_batch_size=sb * world, tol=0.0)
            e1.record(stream)
            barrier()
            ms = torch.tensor([e0.elapsed_time(e1) / 3], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            sgd_ms = float(ms.item())
            sgd_gbs = 8.0 * sd * sb * ran / (sgd_ms * 1e-3) * 1e-9
            sgd_rec = {"workload": "LogisticRegression.fit (SGD): %d x %d FP64 rows per GPU in HBM, maxIter=%d, globalBatchSize=%d"
                                   % (sn, sd, se, sb * world),
                       "unit": "rows/s", "value": sb * world * ran / (sgd_ms * 1e-3), "ms_per_fit": sgd_ms,
                       "us_per_epoch": sgd_ms * 1e3 / max(ran, 1), "epochs": ran, "last_loss": last,
                       "gpu_launches_per_epoch": 1 if world == 1 else 3,
                       "roofline": {"bound": "hbm", "achieved": sgd_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                                    "frac": sgd_gbs / peaks["hbm_gbs"],
                                    "note": "8 d bytes per mini-batch row; at this batch size the epoch is bound by launch and "
                                            "instruction latency, not by the row stream (DESIGN.md)"}}
            cl.close()
            del xs, ys
            torch.cuda.empty_cache()

    # ---- the north-star shape (config 4: 1e9 x 64, k = 1000 over 8 GPUs): this GPU's share, a short run
    north = strong = None
    if not args.no_north_star and args.workload == "c2":
        nn, nd, nk, nrounds = args.north_rows, 64, 1000, 3
        free_b, _ = torch.cuda.mem_get_info(local)
        if free_b > nn * nd * 8 + (8 << 30):
            xs = generate(nn, nd)
            recs, models = {}, {}
            for name, fl in (("f64", 0), ("tf32x3+f64-recheck", native.FLAG_FAST_TF32)):
                if name != "f64" and args.no_fast:
                    continue
                nr = timed_fits(fl, xs, nk, nrounds, 1, 1)
                models[name] = nr["model"]
                recs[name] = {"value": nr["value"], "unit": UNIT, "ms_per_round": nr["ms_per_round"],
                              "per_gpu_fp64_equivalent_tflops": nr["roofline"]["fp64_tensor"]["achieved"],
                              "roofline": {kx: nr["roofline"][kx] for kx in ("bound", "achieved", "peak", "unit", "frac", "kernel", "launch")},
                              "took_tcgen05_kernel": nr["fast"], "rechecked_rows": nr["prof"]["rechecked_rows"]}
            if "tf32x3+f64-recheck" in recs:
                recs["tf32x3+f64-recheck"]["weights_bit_exact_vs_f64_mode"] = bool(
                    np.array_equal(models["f64"][1], models["tf32x3+f64-recheck"][1]))
                recs["tf32x3+f64-recheck"]["centroids_within_1e-9_of_f64_mode"] = bool(
                    np.allclose(models["f64"][0], models["tf32x3+f64-recheck"][0], rtol=1e-9, atol=0, equal_nan=True))
            north = {"workload": "config 4's per-GPU share: %d x %d FP64 rows per GPU (%d in all), k=%d, %d rounds, "
                                 "device-resident, warm-up 1 fit" % (nn, nd, nn * world, nk, nrounds),
                     "scaling": "weak", "modes": recs}
            del xs
            torch.cuda.empty_cache()
            # the same shape with the TOTAL fixed (strong scaling, SURVEY 8d: 2.5e8 rows = 128 GB over 1/2/4/8 GPUs): each
            # rank holds its RowGenerator slice.  (A total that does not fit this GPU's free memory falls back to
            # north_rows, which at one GPU is the record above.)
            total = args.strong_rows - args.strong_rows % world
            per = total // world
            free_b, _ = torch.cuda.mem_get_info(local)
            fits = torch.tensor([1 if free_b > per * nd * 8 + (8 << 30) else 0], dtype=torch.int64, device=dev)
            if world > 1:
                dist.all_reduce(fits, op=dist.ReduceOp.MIN)
            if not int(fits.item()):
                total = nn - nn % world
                per = total // world
            if world == 1 and per == nn:
                strong = {"workload": "%d x %d FP64 rows in all, k=%d (at one GPU: the north_star record)" % (nn, nd, nk),
                          "scaling": "strong", "rows_total": nn,
                          "modes": {m: {"value": v["value"], "unit": UNIT, "ms_per_round": v["ms_per_round"]} for m, v in recs.items()}}
            else:
                xs = generate(per, nd)
                srecs = {}
                for name, fl in (("f64", 0), ("tf32x3+f64-recheck", native.FLAG_FAST_TF32)):
                    if name != "f64" and args.no_fast:
                        continue
                    nr = timed_fits(fl, xs, nk, 2, 1, 1)
                    srecs[name] = {"value": nr["value"], "unit": UNIT, "ms_per_round": nr["ms_per_round"]}
                strong = {"workload": "%d x %d FP64 rows in all (%d per GPU), k=%d, 2 rounds, device-resident" % (
                              total, nd, per, nk), "scaling": "strong", "rows_total": total, "modes": srecs}
                del xs
                torch.cuda.empty_cache()

    base = None
    if rank == 0 and world == 1 and not args.no_cpu:
        base, _, _ = cpu_baseline(d, k)

    if rank == 0:
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "tf32x3+f64-recheck" if r["fast"] else "f64", "data": "synthetic",
            "config": {"workload": desc, "rows_per_gpu": n, "rows_total": total_rows, "d": d, "k": k,
                       "rounds_per_step": rounds, "sharding": "rows by subtask as RowGenerator (rank = subtask)",
                       "collective": "ncclAllReduce of k*d+k FP64 per round" if world > 1 else "none (1 GPU)",
                       "l2": "inputs (%.2f GB per GPU) are larger than L2, no flush needed" % (n * d * 8 / 1e9)
                       if n * d * 8 > 200e6 else "inputs fit in L2 (reference's own small config)",
                       "data_seed": DATA_SEED, "init": "reference sampler, KMeans default seed"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": prof["total_launches"],
            "roofline": roofline, "cpu_baseline": base, "parity": parity,
            "wall_ms_per_step": wall / args.steps * 1e3, "rechecked_rows": prof["rechecked_rows"],
            "fast_mode": fast_rec, "transform": transform_rec, "online": online_rec, "knn": knn_rec, "sgd": sgd_rec, "north_star": north, "strong_scaling": strong,
        }
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--rows", type=int, default=0, help="override rows per GPU")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the world > 1 parity leg (N > 1 only)")
    ap.add_argument("--fast", action="store_true",
                    help="headline line in the tcgen05 fast mode (FMLK_FLAG_FAST_TF32) instead of the FP64 mode")
    ap.add_argument("--no-fast", action="store_true", help="skip the fast-mode sub-record of the default run")
    ap.add_argument("--no-north-star", action="store_true", help="skip the config-4-share sub-record")
    ap.add_argument("--knn-rows", type=int, default=1 << 17, help="rows per GPU of the KnnModel.transform sub-record")
    ap.add_argument("--sgd-rows", type=int, default=10_000_000, help="rows per GPU of the SGD (LogisticRegression) sub-record")
    ap.add_argument("--no-extra", action="store_true", help="skip the transform (config 3) and online (config 5) sub-records")
    ap.add_argument("--transform-rows", type=int, default=4_000_000, help="rows per GPU of the transform sub-record")
    ap.add_argument("--online-batches", type=int, default=40)
    ap.add_argument("--north-rows", type=int, default=125_000_000, help="rows per GPU of the north-star sub-record")
    ap.add_argument("--strong-rows", type=int, default=250_000_000, help="total rows of the strong-scaling sub-record")
    args = ap.parse_args()
    if args.impl == "ours":
        args.warmup = max(args.warmup, 3)  # timing rule: at least 3 warm-up steps
    # exactly one line on stdout: libraries that talk on fd 1 (NCCL prints its version there on some boxes) are
    # sent to stderr for the duration of the run; the JSON line goes to the real stdout
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w", buffering=1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
    sys.stdout.flush()


if __name__ == "__main__":
    main()
