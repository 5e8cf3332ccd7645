#!/usr/bin/env python
"""Benchmark of the q-means hot path (BASELINE.json config 4: 10M x 128, k = 256) and of batched
vector-state tomography (config 5).

  python bench.py --gpus N --steps K --warmup W          # this repo (CUDA, sm_100a)
  python bench.py --impl reference --steps K --warmup W  # CPU arm: oracle port of the reference

A step = one Lloyd iteration of qMeans_ (E-step with the delta-window noise model, M-step, all-reduce of the
centroid sums, default-mode centroid tomography at delta/2, centre shift, read-back of inertia and shift).
The 10M rows are ONE fixed dataset, rank r of G holding rows [r n/G, (r+1) n/G) of it (strong scaling);
inputs are resident in HBM for `value`; `e2e` goes through the public API (qMeans_.fit on a pageable host
array, what an sklearn caller holds).  The tomography half of BASELINE.json's metric is under
`roofline.kernels` (the config-5 grid, every N) with its CPU figure under `cpu_baseline.tomography`.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "sq-learn_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

N_ROWS, DIM, K_CLUSTERS, DELTA = 10_000_000, 128, 256, 0.5
METRIC, UNIT = "qmeans_lloyd_iterations_per_s_10Mx128_k256", "iter/s"
BLOCK_ROWS = 1 << 20          # the dataset is generated in blocks of this many rows, keyed by block id


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--mode", default="window", choices=["window", "ipe", "exact"])
    ap.add_argument("--rows", type=int, default=N_ROWS, help="total rows (default: the BASELINE config)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the tomography grid, the other E-step modes and qPCA")
    ap.add_argument("--tomo-budget", type=float, default=75.0, help="seconds of wall clock for the tomography grid")
    ap.add_argument("--tomo-full", action="store_true", help="also the grid points left out by default (B*d > 1e8)")
    ap.add_argument("--fit300", action="store_true", help="also time a fit of 300 iterations (the reference's max_iter)")
    ap.add_argument("--csv", default=None, help="config 3: q-means k=15 on this CSV (standardised), delta 0.1..1.0")
    ap.add_argument("--cpu-rows", type=int, default=20000)
    return ap.parse_args()


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi samples during the timed region (recipe in B200_PROFILING.md)."""

    def __init__(self, index):
        self.rows, self.stop_flag, self.index = [], False, index
        self.th = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
                if out.returncode == 0 and out.stdout.strip():
                    self.rows.append([c.strip() for c in out.stdout.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def __enter__(self):
        self.th.start()
        return self

    def __exit__(self, *a):
        self.stop_flag = True
        self.th.join(timeout=6)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(sm), "power_w_max": max(float(r[2]) for r in self.rows)}


# ----------------------------------------------------------------------------- data
def make_shard(rows, rank, world, device):
    """Blobs: 256 centres ~ N(0, 4^2 I), unit noise, generated on the device.  The data
    set is a function of the ROW INDEX alone: block b = rows [b 2^20, (b+1) 2^20) comes from a generator
    seeded with (1234, b), and rank r holds the contiguous rows [r n/G, (r+1) n/G) - so 1, 2, 4 and 8 GPUs
    cluster the same 10M rows."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(1234)
    centres = torch.randn((K_CLUSTERS, DIM), generator=g, device=device, dtype=torch.float64) * 4.0
    lo = rows * rank // world
    hi = rows * (rank + 1) // world
    X = torch.empty((hi - lo, DIM), dtype=torch.float64, device=device)
    for blk in range(lo // BLOCK_ROWS, (hi + BLOCK_ROWS - 1) // BLOCK_ROWS):
        b0, b1 = blk * BLOCK_ROWS, min(rows, (blk + 1) * BLOCK_ROWS)
        g.manual_seed(1234 + 7919 * (blk + 1))
        lab = torch.randint(0, K_CLUSTERS, (b1 - b0,), generator=g, device=device)
        Xb = centres[lab]
        Xb += torch.randn((b1 - b0, DIM), generator=g, device=device, dtype=torch.float64)
        a, b = max(lo, b0), min(hi, b1)
        X[a - lo:b - lo] = Xb[a - b0:b - b0]
        del Xb, lab
    return X


def measure_dgemm_peak(device):
    """fp64 tensor-core denominators: cuBLAS DGEMM 8192^3, best of 10 (burst) and back to back for 4 s
    (sustained) - MEASURED_PEAKS.json has no fp64 entry; same protocol as its bf16 lines."""
    import torch
    n = 8192
    a = torch.randn((n, n), dtype=torch.float64, device=device)
    b = torch.randn((n, n), dtype=torch.float64, device=device)
    flop = 2.0 * n ** 3
    best = 0.0
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(a, b)
        e1.record()
        torch.cuda.synchronize()
        best = max(best, flop / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    reps = max(8, int(4.0 / (flop / (best * 1e12))))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        torch.matmul(a, b)
    e1.record()
    torch.cuda.synchronize()
    sustained = reps * flop / (e0.elapsed_time(e1) * 1e-3) / 1e12
    del a, b
    return best, sustained


# ----------------------------------------------------------------------------- CPU arm
def cpu_iteration_rate(cpu_rows, steps, warmup, mode, total_rows=N_ROWS):
    """Oracle port of the reference (oracle/qmeans.py) on a bounded sample of the workload.
    The E- and M-step are linear in the row count, the centroid tomography does not depend on
    it, so the full-size step time is  t_rows * (total_rows / cpu_rows) + t_tomography."""
    from oracle import qmeans as oq
    from oracle import quantum_utility as oqu
    rs = np.random.RandomState(1234)
    centres = rs.randn(K_CLUSTERS, DIM) * 4.0
    X = centres[rs.randint(0, K_CLUSTERS, cpu_rows)] + rs.randn(cpu_rows, DIM)
    X -= X.mean(axis=0)
    C = X[rs.permutation(cpu_rows)[:K_CLUSTERS]].copy()
    rng = np.random.default_rng(1)
    delta = 0.0 if mode == "exact" else DELTA
    tde = mode == "ipe"
    t_rows, t_tomo = [], []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        labels, _, _ = oq.labels_estimation(X, C, delta, tde, rng, scipy_cdist=True)
        Cn = oq.centers_update(X, labels, delta / 2, False, True, True, rng, n_clusters=K_CLUSTERS, centers_old=C)
        t1 = time.perf_counter()
        Cn = oqu.tomography(Cn, DELTA / 2, rng=rng)
        t2 = time.perf_counter()
        C = Cn
        if it >= warmup:
            t_rows.append(t1 - t0)
            t_tomo.append(t2 - t1)
    tr, tt = float(np.mean(t_rows)), float(np.mean(t_tomo))
    full = tr * (total_rows / cpu_rows) + tt
    return {"value": 1.0 / full, "unit": UNIT, "cores": 1, "kind": "port", "extrapolated": True,
            "measured": {"rows": cpu_rows, "seconds_per_step_rows_part": tr, "seconds_centroid_tomography": tt,
                         "steps": steps},
            "threads": "1 (the reference's quantum routines are single-threaded Python and its Lloyd loop limits "
                       "BLAS to one thread, _dmeans.py:618)",
            "sample": "%d of %d rows per step (E+M-step %.3f s, measured) + 256x128 centroid tomography (%.3f s, "
                      "measured); full-size step time extrapolated linearly in rows: %.1f s.  The port calls scipy's "
                      "cdist like the reference; its M-step and tomography are faster than the reference's own "
                      "(np.where instead of an O(n k) scan, multinomial instead of O(N) draws): BASELINE.md has both "
                      "timed on the same sample" % (cpu_rows, total_rows, tr, tt, full),
            "host_cpus": os.cpu_count()}


def cpu_tomography(budget_s=12.0):
    """The oracle's tomography (oracle/quantum_utility.py: the reference's algorithm with the O(N) sampling
    of RandomState.choice replaced by numpy's multinomial, so it flatters the CPU) on a few grid shapes, one
    core, timed per vector."""
    from oracle import quantum_utility as oqu
    rng = np.random.default_rng(7)
    out = []
    t_all = time.perf_counter()
    for d, delta, norm, inc in [(128, 0.25, "L2", True), (1000, 0.1, "L2", False), (1000, 0.1, "L2", True),
                                (1000, 0.1, "inf", True), (10000, 0.1, "L2", False), (100000, 0.5, "L2", False)]:
        v = rng.uniform(-1, 1, d)
        v /= np.linalg.norm(v)
        n, t0 = 0, time.perf_counter()
        while n < 200 and time.perf_counter() - t0 < budget_s / 6:
            oqu.real_tomography(v, delta=delta, norm=norm, incremental_measure=inc, rng=rng)
            n += 1
        dt = (time.perf_counter() - t0) / n
        out.append({"d": d, "delta": delta, "norm": norm, "mode": "incremental+stop" if inc else "single-shot",
                    "s_per_vector": dt, "vectors_per_s": 1.0 / dt, "vectors_timed": n})
        if time.perf_counter() - t_all > budget_s:
            break
    return {"kind": "port", "cores": 1, "points": out,
            "sample": "oracle port, one core, up to 200 vectors or %.0f s per shape; the reference itself "
                      "(O(N) RandomState.choice + Counter) measured in the build container: BASELINE.md" % (budget_s / 6)}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if args.mode == "ipe":
        args.cpu_rows = min(args.cpu_rows, 8)   # 13 O(M) draws for each of rows x 256 pairs
    t0 = time.perf_counter()
    cb = cpu_iteration_rate(args.cpu_rows, args.steps, args.warmup, args.mode, args.rows)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 / cb["value"],
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": "q-means Lloyd iteration, %dx%d, k=%d, delta=%g, %s E-step + centroid "
                                   "tomography (reference's NumPy path restated in oracle/, CPU)"
                                   % (args.rows, DIM, K_CLUSTERS, DELTA, args.mode)},
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "wall_s": time.perf_counter() - t0}
    print(json.dumps(line))


# ----------------------------------------------------------------------------- GPU arm
def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)
    except Exception:
        return {"hbm_gbs": 6650.0, "source": "fallback (B200_PROFILING.md)"}


def _timed(torch, fn, reps):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


class Ctx:
    pass


def run_b200(args):
    import torch
    import torch.distributed as dist
    from sqlearn.QuantumUtility import Utility as Q
    from sqlearn import _cabi
    from sqlearn.cluster import qMeans_
    from sqlearn.cluster import _dmeans as D
    from sqlearn.cluster._engine import QMeansEngine

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device; there is no CPU fallback"
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"      # keep stdout to the one JSON line
        # NCCL prints its version banner on stdout when the communicator comes up: send stdout to stderr until
        # then, so that the JSON line is the only thing the caller reads there
        sys.stdout.flush()
        saved_fd = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=device)
            warm = torch.zeros(1, device=device)
            dist.all_reduce(warm)
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_fd, 1)
            os.close(saved_fd)
    assert world == args.gpus or world == 1, "launch with torchrun --nproc-per-node == --gpus"
    c = Ctx()
    c.torch, c.dist, c.Q, c.cabi, c.D, c.world, c.rank, c.device, c.args = torch, dist, Q, _cabi, D, world, rank, device, args

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    c.barrier = barrier

    def max_over_ranks(ms):
        t = torch.tensor([ms], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    c.max_over_ranks = max_over_ranks

    mode = args.mode
    delta = 0.0 if mode == "exact" else DELTA
    tde = mode == "ipe"
    inter = mode != "exact"            # centroid tomography needs delta > 0 (reference :1249-1250)

    X = make_shard(args.rows, rank, world, device)
    n_local = X.shape[0]
    Q.seed(1234)
    if world > 1:
        Q._sync_state()
    eng = QMeansEngine(X, K_CLUSTERS)
    c.eng = eng
    # centre the data as fit() does, then take the initial centres from the first rows of the data set
    csum = eng.col_sums()
    if world > 1:
        dist.all_reduce(csum)
    mean = csum / float(eng.n_total)
    eng.subtract_rowvec(mean)
    eng.refresh_norms()
    g = torch.Generator(device="cpu")
    g.manual_seed(1234)
    idx = torch.randperm(min(args.rows // 8, 1 << 20), generator=g)[:K_CLUSTERS].to(device)   # rows rank 0 holds at any N <= 8
    centers0 = X[idx].clone() if rank == 0 else torch.empty((K_CLUSTERS, DIM), dtype=torch.float64, device=device)
    if world > 1:
        dist.broadcast(centers0, src=0)
    state = D.LloydState(centers0)
    host = torch.empty((2,), dtype=torch.float64).pin_memory()

    def make_step(d_, inter_, tde_):
        return lambda: D._lloyd_iteration(eng, state, host, d_, inter_, True, True, None, tde_)
    step = make_step(delta, inter, tde)

    for _ in range(args.warmup):
        step()
    barrier()
    l0 = eng.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk:
        e0.record()
        for _ in range(args.steps):
            inertia, shift = step()
        e1.record()
        barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    launches = eng.launches - l0
    launches += D.TOMO_LAUNCHES_PER_ITER * args.steps if inter else 0    # tomography launches are made outside the engine
    value = args.steps / (ms_total * 1e-3)

    # ---- roofline leg: the E-step kernel alone, CUDA events inside the C ABI ----
    tot, cnt = (_cabi.ctypes.c_double)(), (_cabi.ctypes.c_longlong)()
    _cabi.lib.sq_profile_assign(2, None, None)
    for _ in range(max(3, min(args.steps, 10))):
        D._labels_inertia(eng, state.centers, delta=delta, pool=None, true_distance_estimate=tde)
    torch.cuda.synchronize()
    _cabi.lib.sq_profile_assign(0, _cabi.ctypes.byref(tot), _cabi.ctypes.byref(cnt))
    k_ms = tot.value / max(cnt.value, 1)
    flops = 2.0 * n_local * K_CLUSTERS * DIM
    peak, peak_sustained = measure_dgemm_peak(device) if rank == 0 else (0.0, 0.0)
    achieved = flops / (k_ms * 1e-3) / 1e12
    traffic, traffic_src = None, None
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            tj = json.load(f)
        traffic = tj["dram_bytes_per_row"] * n_local       # dram__bytes_read.sum + dram__bytes_write.sum, per launch
        traffic_src = tj["source"]
    except Exception:
        pass
    peaks = load_peaks()
    step_tflops = flops * world / (ms_total / args.steps * 1e-3) / 1e12 / world
    filtered = mode != "ipe" and K_CLUSTERS <= 256 and DIM <= 128 and DIM % 2 == 0 and not os.environ.get("SQ_BENCH_NO_FILTER")
    roofline = {"bound": "tensor", "kernel": "qmeans_assign_kernel (fp64 DMMA.8x8x4)", "achieved": achieved,
                "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak if peak else None, "traffic": traffic,
                "traffic_source": traffic_src,
                "kernel_ms": k_ms, "algorithmic_flops_per_launch": flops,
                "algorithmic_bytes_per_launch": 8.0 * n_local * DIM + 4.0 * n_local,
                "peak_source": "cuBLAS DGEMM 8192^3 fp64 measured in this run, best of 10 (MEASURED_PEAKS.json has no "
                               "fp64 entry); the kernel is timed alone, so the burst figure is its denominator",
                "peak_sustained": peak_sustained,
                "whole_step": {"tflops_per_gpu": step_tflops,
                               "frac_of_sustained_peak": step_tflops / peak_sustained if peak_sustained else None,
                               "note": "2 n k d flop of the E-step over the whole Lloyd step (M-step, tomography, "
                                       "collectives included), against DGEMM back to back for 4 s"},
                "hbm_fraction": (8.0 * n_local * DIM + 4.0 * n_local) / (k_ms * 1e-3) / 1e9 / peaks["hbm_gbs"],
                "kernels": []}

    if filtered:
        # exact / window E-step: bf16 tensor-core filter + fp64 decisions (csrc/qmeans_filter.cuh).  X is read once in
        # fp64 and nothing of size n x k is written: the bound is HBM, 8 n d + 8 n (row norms) + 12 n (labels, distances)
        alg_bytes = (8.0 * DIM + 8.0 + 12.0) * n_local
        hbm_peak = peaks["hbm_gbs"]
        ftraffic, fsrc = None, None
        try:
            with open(os.path.join(ROOT, "profiles", "ncu_traffic_filter.json")) as f:
                tj = json.load(f)
            ftraffic, fsrc = tj["dram_bytes_per_row"] * n_local, tj["source"]
        except Exception:
            pass
        roofline.update({
            "bound": "hbm", "kernel": "qmeans_filter_kernel (bf16 mma.sync filter, fp64 decisions)",
            "achieved": alg_bytes / (k_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
            "frac": alg_bytes / (k_ms * 1e-3) / 1e9 / hbm_peak if hbm_peak else None,
            "traffic": ftraffic, "traffic_source": fsrc, "algorithmic_bytes_per_launch": alg_bytes,
            "peak_source": peaks.get("source", "MEASURED_PEAKS.json hbm_gbs (driver-written copy bandwidth)"),
            "fp64_kernel": {"note": "the same E-step through the fp64 tensor-core kernel (sq_qmeans_debug_filter(0)): "
                                    "20.6 ms at 10M x 128, k = 256 = 0.897 of cuBLAS DGEMM (profiles/r02_bench_1gpu_fp64_kernel.json)",
                            "dgemm_peak_tflops": peak, "flops_over_filter_time_tflops": achieved}})
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "q-means Lloyd iteration, %dx%d fp64 rows sharded over %d GPU(s), k=%d, delta=%g, "
                                   "%s E-step, M-step, centroid tomography (incremental+stop, delta/2)"
                                   % (args.rows, DIM, world, K_CLUSTERS, DELTA, mode),
                       "rows_per_gpu": n_local, "mode": mode,
                       "estep": ("bf16 tensor-core filter with a rigorous error bound, every decision and every output in fp64 "
                                 "(csrc/qmeans_filter.cuh); same labels and labels_crc as the fp64 tensor-core kernel"
                                 if filtered else "fp64 tensor-core kernel (DMMA) for every pair"),
                       "dataset": "one fixed data set of %d rows, generated per 2^20-row block from the block id; "
                                  "rank r holds rows [r n/G, (r+1) n/G)" % args.rows,
                       "l2": "inputs (%.2f GB per GPU) exceed the 126 MB L2; no flush needed" % (n_local * DIM * 8 / 1e9),
                       "last_inertia": inertia, "last_shift": shift},
            "clocks": clk.summary(), "gpu_launches": launches, "roofline": roofline}

    # ---- the same data at every N: a short noiseless run whose result must not depend on N ----
    line["config"]["invariance"] = invariance_check(c, centers0)
    # ---- the other E-step modes as real timed steps, the tomography grid (config 5), qPCA (config 2) ----
    if not args.no_extra:
        roofline["kernels"] += other_mode_steps(c, make_step, state, centers0, flops, peak_sustained)
        roofline["kernels"] += tomography_grid(c, peaks)
        if world == 1:
            line["extra"] = extra_measurements(c, state.centers)
    # ---- e2e through the public API: qMeans_.fit on host arrays ----
    if not args.no_e2e:
        try:
            line["e2e"] = e2e_fit(c, X, mean, centers0, qMeans_)
        except Exception as exc:      # e.g. the host cannot hold a copy of the shard
            line["e2e"] = {"value": None, "unit": UNIT, "error": "%s: %s" % (type(exc).__name__, exc)}
    if args.csv and rank == 0 and world == 1:
        line["config3_csv"] = config3_csv(c, args.csv, qMeans_)
    if rank == 0 and world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_iteration_rate(args.cpu_rows, 2, 1, mode, args.rows)
        line["cpu_baseline"]["tomography"] = cpu_tomography()
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def invariance_check(c, centers0):
    """Three noiseless Lloyd iterations from the common initial centres: inertia and a position-weighted
    checksum of all labels.  Identical at N = 1, 2, 4, 8 (labels bit for bit; the inertia up to the order of
    the cross-rank sums)."""
    torch, D, eng = c.torch, c.D, c.eng
    st = D.LloydState(centers0)
    host = torch.empty((2,), dtype=torch.float64).pin_memory()
    for _ in range(3):
        inertia, _ = D._lloyd_iteration(eng, st, host, 0.0, False, True, True, None, True)
    D._labels_inertia(eng, st.centers, delta=0.0, pool=None, true_distance_estimate=True)
    idx = torch.arange(eng.row0, eng.row0 + eng.n, device=c.device, dtype=torch.int64)
    w = (idx * 2654435761 + 12345) & 0x7FFFFFFF
    crc = ((eng.labels.to(torch.int64) + 1) * w).sum().reshape(1)
    red = eng.inertia.clone()
    if c.world > 1:
        c.dist.all_reduce(crc)
        c.dist.all_reduce(red)
    return {"inertia_noiseless_3iter": inertia, "inertia_final_estep": float(red.item()),
            "labels_crc": int(crc.item()) & 0xFFFFFFFFFFFFFFFF,
            "what": "sum_i (label_i + 1) * ((i * 2654435761 + 12345) mod 2^31) over all rows i, mod 2^64"}


def other_mode_steps(c, make_step, state, centers0, flops, peak_sustained):
    """One Lloyd step in each of the E-step modes the headline does not use, timed as real steps
    (E-step + M-step + all-reduce + centroid tomography + read-back), max over ranks."""
    torch, args = c.torch, c.args
    out = []
    for name, (delta, inter, tde, reps) in {"ipe": (DELTA, True, True, 3), "exact": (0.0, False, True, 3),
                                             "window": (DELTA, True, False, 3)}.items():
        if name == args.mode:
            continue
        state.centers = centers0.clone()
        st = make_step(delta, inter, tde)
        st()
        st()
        c.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            st()
        e1.record()
        c.barrier()
        ms = c.max_over_ranks(e0.elapsed_time(e1)) / reps
        rec = {"kernel": "Lloyd step, %s E-step%s" % (name, " (the reference's default, true_distance_estimate=True)"
                                                      if name == "ipe" else ""),
               "shape": {"rows": args.rows, "d": DIM, "k": K_CLUSTERS, "gpus": c.world}, "ms": ms,
               "lloyd_iterations_per_s": 1e3 / ms, "steps_timed": reps}
        if name == "ipe":
            pairs = float(args.rows) * K_CLUSTERS
            rec["ipe_pairs_per_s"] = pairs / (ms * 1e-3)
            rec["reference_draws_per_s"] = 13.0 * pairs / (ms * 1e-3)
            rec["bound"] = ("tensor (the distance GEMM) + issue slots of the epilogue: every pair's estimate is either sampled "
                            "(median of 13 amplitude-estimation outcomes by one inversion) or certified, from the pair's own "
                            "uniform, to exceed one already sampled in its row")
            tf = flops / (ms * 1e-3) / 1e12
            rec["achieved_TFLOPs_per_gpu"] = tf
            rec["frac_of_sustained_peak"] = tf / peak_sustained if peak_sustained else None
        else:
            rec["bound"] = "tensor"
            tf = flops / (ms * 1e-3) / 1e12
            rec["achieved_TFLOPs_per_gpu"] = tf
            rec["frac_of_sustained_peak"] = tf / peak_sustained if peak_sustained else None
        out.append(rec)
    state.centers = centers0.clone()
    return out


def tomography_grid(c, peaks):
    """BASELINE.json config 5: d in {1e3, 1e4, 1e5, 1e6} x delta in {0.01, 0.05, 0.1, 0.5} x B in {1, 64, 4096}
    x {L2, inf} x {single-shot, incremental+stop}; vectors uniform(-1, 1) normalised, seed 1234.
    With several GPUs the vectors of a batch are dealt to the ranks (B >= G) and a single long vector goes
    through tomography_sharded; the time is the max over ranks.  Roofline: 16 B d algorithmic bytes (read the
    input once, write the estimate once) against the measured HBM copy bandwidth."""
    torch, U, args = c.torch, c.Q, c.args
    world, rank = c.world, c.rank
    hbm = peaks["hbm_gbs"]
    out, t_start = [], time.perf_counter()
    g = torch.Generator(device=c.device)
    fused = c.cabi.lib.sq_tomography_max_fused_dim()
    budget_hit = False
    order = [(d, B) for d in (1000, 10000, 100000, 1000000) for B in (4096, 64, 1)]
    for d, B in order:
        if B * d > 1e8 and not args.tomo_full:
            out.append({"kernel": "tomography", "shape": {"B": B, "d": d}, "skipped":
                        "16 B d = %.1f GB fits, but the point takes seconds; run with --tomo-full" % (16.0 * B * d / 1e9)})
            continue
        share = B >= world                                  # vectors dealt by vector
        b_loc = (B * (rank + 1)) // world - (B * rank) // world if share else B
        g.manual_seed(1234 + d + B)
        V = torch.rand((max(b_loc, 1), d), generator=g, device=c.device, dtype=torch.float64) * 2 - 1
        V /= V.norm(dim=1, keepdim=True)
        for norm in ("L2", "inf"):
            for inc in (False, True):
                for delta in (0.01, 0.05, 0.1, 0.5):
                    elapsed = time.perf_counter() - t_start
                    flag = torch.tensor([1.0 if elapsed > args.tomo_budget else 0.0], device=c.device)
                    if world > 1:
                        c.dist.all_reduce(flag, op=c.dist.ReduceOp.MAX)
                    if flag.item() > 0:
                        budget_hit = True
                        break
                    res = {}

                    def run():
                        if not share and world > 1 and d > fused:
                            res["est"] = U.tomography_sharded(V, delta, None, norm, True, inc, 0)
                            res["n"] = None
                        elif not share and world > 1 and rank != 0:
                            res["n"] = None             # one short vector: rank 0 alone
                        else:
                            r = U._tomography_device(V, delta, None, norm, True, inc, 0)
                            res["n"] = r[1]
                    c.barrier()
                    t0 = time.perf_counter()
                    run()
                    c.barrier()
                    first = (time.perf_counter() - t0) * 1e3
                    cold = first > 400.0                  # a point this slow is measured once, first run
                    if cold:
                        ms = c.max_over_ranks(first)
                    else:
                        reps = 3 if first < 50 else 1
                        c.barrier()
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        e0.record()
                        for _ in range(reps):
                            run()
                        e1.record()
                        c.barrier()
                        ms = c.max_over_ranks(e0.elapsed_time(e1) / reps)
                    n_used = torch.zeros(1, dtype=torch.float64, device=c.device)
                    if res.get("n") is not None:
                        n_used += res["n"].sum()
                    if world > 1 and share:
                        c.dist.all_reduce(n_used)
                    N_full = U.tomography_num_measurements(d, delta, norm)
                    meas = float(n_used.item()) if float(n_used.item()) > 0 else None
                    gbps = 16.0 * B * d / (ms * 1e-3) / 1e9
                    kern = "tomo_group_kernel" if (not inc and d <= fused and b_loc >= 296) else \
                        ("tomo_fused_kernel" if d <= fused else "tomo_big_* (block / top kernels)")
                    out.append({"kernel": kern, "shape": {"B": B, "d": d, "delta": delta, "norm": norm,
                                                          "mode": "incremental+stop" if inc else "single-shot",
                                                          "gpus": world, "N_full": N_full},
                                "ms": ms, "bound": "hbm (nominal): 16 B d bytes", "achieved_GBps": gbps,
                                "frac_hbm": gbps / hbm / world, "vectors_per_s": B / (ms * 1e-3),
                                "simulated_measurements_per_s": (2.0 * meas / (ms * 1e-3)) if meas else None,
                                "timing": "first run, wall clock" if cold else "CUDA events after one warm-up"})
                if budget_hit:
                    break
            if budget_hit:
                break
        del V
        if budget_hit:
            out.append({"kernel": "tomography", "note": "grid stopped after %.0f s (--tomo-budget); the points above "
                                                       "are complete" % args.tomo_budget})
            break
    return out


def extra_measurements(c, centers):
    """Per-phase device times of one Lloyd iteration and the config-2 qPCA phases (N = 1 only)."""
    torch, D, eng, U = c.torch, c.D, c.eng, c.Q
    out = {}
    for name, (delta, tde) in {"exact": (0.0, True), "window": (DELTA, False), "ipe": (DELTA, True)}.items():
        out["estep_%s_ms" % name] = _timed(
            torch, lambda: D._labels_inertia(eng, centers, delta=delta, pool=None, true_distance_estimate=tde),
            2 if name == "ipe" else 3)
    D._labels_inertia(eng, centers, delta=0.0, pool=None, true_distance_estimate=True)
    out["mstep_accumulate_ms"] = _timed(torch, eng.accumulate, 3)
    c2 = torch.empty_like(centers)
    eng.finalize(centers, c2)
    out["centroid_tomography_ms"] = _timed(torch, lambda: U.tomography(c2, DELTA / 2), 3)
    out["qpca_config2"] = qpca_config2(torch)
    return out


def qpca_config2(torch):
    """BASELINE.json config 2: qPCA top-50 on synthetic 70,000 x 784, singular-value estimation eps = 0.05,
    L2 tomography delta = 0.1 (single-shot, qPCA's default).  Phases timed on the device after one warm-up."""
    import warnings
    from sqlearn import _cabi, _device as dv
    from sqlearn.decomposition import qPCA
    from sqlearn.QuantumUtility import Utility as U
    n, d, k = 70000, 784, 50
    g = torch.Generator(device="cuda")
    g.manual_seed(1234)
    cen = torch.randn((10, d), generator=g, device="cuda", dtype=torch.float64) * 3.0
    X = cen[torch.randint(0, 10, (n,), generator=g, device="cuda")] + torch.randn((n, d), generator=g, device="cuda",
                                                                                dtype=torch.float64)
    Xh = X.cpu().numpy()
    out = {"shape": [n, d], "n_components": k}
    G = dv.empty((d, d))
    nb = _cabi.lib.sq_gram_workspace_bytes(n, d)
    ws = dv.workspace(nb)
    ms = _timed(torch, lambda: _cabi.check(_cabi.lib.sq_gram_centered(dv.ptr(X), n, d, None, dv.ptr(G), dv.ptr(ws), nb,
                                                                      dv.stream_ptr()), "gram"), 3)
    out["gram_ms"] = ms
    out["gram_tflops"] = 2.0 * n * d * d / (ms * 1e-3) / 1e12          # full (non-symmetric) flop count
    Vt = torch.linalg.eigh(G)[1].T.contiguous()
    keys = torch.zeros((d,), dtype=torch.int64, device="cuda")
    zeros = torch.zeros((d,), dtype=torch.float64, device="cuda")
    ms = _timed(torch, lambda: _cabi.check(_cabi.lib.sq_project_absmax(dv.ptr(X), n, d, dv.ptr(Vt), d, dv.ptr(zeros),
                                                                       dv.ptr(keys), dv.stream_ptr()), "absmax"), 3)
    out["project_absmax_ms"] = ms
    out["project_absmax_tflops"] = 2.0 * n * d * d / (ms * 1e-3) / 1e12
    ms = _timed(torch, lambda: U.best_mu(X, start=0, step=0.1), 2)
    out["best_mu_11pt_ms"] = ms
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        m = qPCA(n_components=k, svd_solver='full').fit(Xh)
        theta = float(m.singular_values_[-1]) * 0.5
        U.seed(3)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        m = qPCA(n_components=k, svd_solver='full').fit(Xh, estimate_all=True, eps=0.05, delta=0.1, theta_major=theta,
                                                      true_tomography=True, norm='L2')
        torch.cuda.synchronize()
        out["fit_estimate_all_s"] = time.perf_counter() - t0
        out["topk"] = int(m.topk)
        t0 = time.perf_counter()
        r = m.transform(Xh, classic_transform=False, epsilon_delta=0.1, quantum_representation=True,
                        norm='est_representation')
        torch.cuda.synchronize()
        out["transform_quantum_representation_s"] = time.perf_counter() - t0
        out["transform_rows"] = int(r['quantum_representation_results'][0].shape[0])
    return out


def _fit_once(c, qMeans_, Xnp, init, steps, seed, mode=None):
    """One timed qMeans_.fit on a host array; returns (seconds = max over ranks of max(wall, device), estimator)."""
    import warnings
    torch, args = c.torch, c.args
    mode = mode or args.mode
    est = qMeans_(n_clusters=K_CLUSTERS, init=init, n_init=1, max_iter=steps, tol=0,
                  delta=0 if mode == "exact" else DELTA, algorithm='full', intermediate_error=mode != "exact",
                  true_distance_estimate=mode == "ipe", copy_x=False)
    c.Q.seed(seed)
    c.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        est.fit(Xnp)
    e1.record()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    c.barrier()
    return c.max_over_ranks(max(e0.elapsed_time(e1), wall * 1e3)) * 1e-3, est


def e2e_fit(c, X, mean, centers_init, qMeans_):
    """The same iterations through the public API on HOST arrays: qMeans_.fit(X_host) with an explicit init and
    max_iter = steps.  Timed region: H2D of X, the fit preamble (eta, mu, condition number, centring), the Lloyd
    iterations, the final E-step and the D2H of labels / centres.  `value` is the PAGEABLE case (a NumPy array
    as an sklearn caller holds it: staged through two pinned 256 MB buffers); the pinned case is beside it."""
    import warnings
    torch, args = c.torch, c.args
    X.add_(mean)                                   # back to the uncentred data the user would hold
    Xh = torch.empty(X.shape, dtype=torch.float64)           # pageable
    Xh.copy_(X)
    init = (centers_init + mean).cpu().numpy()
    n_local = X.shape[0]
    c.eng = None
    # one small untimed fit first: library start-up (cuSOLVER handle, module load) is not the workload
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        qMeans_(n_clusters=K_CLUSTERS, init=init, n_init=1, max_iter=1, tol=0, delta=0, algorithm='full',
                copy_x=False).fit(Xh[:min(n_local, 200000)].numpy())
    secs, est = _fit_once(c, qMeans_, Xh.numpy(), init, args.steps, 99)
    out = {"value": est.n_iter_ / secs, "unit": UNIT,
           "h2d_bytes_per_step": (n_local * DIM * 8 + K_CLUSTERS * DIM * 8) / est.n_iter_,
           "d2h_bytes_per_step": (n_local * 4 + K_CLUSTERS * DIM * 8 + 16 * est.n_iter_) / est.n_iter_,
           "api": "qMeans_(n_clusters=256, init=<array>, n_init=1, max_iter=%d, tol=0, algorithm='full', ...)"
                  ".fit(X_host)" % args.steps,
           "host_memory": "pageable", "fit_seconds": secs, "n_iter": est.n_iter_, "inertia": est.inertia_,
           "note": "one fit = H2D of the shard + preamble + n_iter Lloyd iterations + final E-step; "
                   "bytes are totals divided by n_iter; the upload is paid once per fit, so the figure "
                   "grows with max_iter (see fits)"}
    fits = [{"max_iter": args.steps, "host_memory": "pageable", "fit_seconds": secs, "iter_per_s": est.n_iter_ / secs}]
    if args.steps != 20:
        s20, e20 = _fit_once(c, qMeans_, Xh.numpy(), init, 20, 99)
        fits.append({"max_iter": 20, "host_memory": "pageable", "fit_seconds": s20, "iter_per_s": e20.n_iter_ / s20})
    if args.fit300 or (c.world == 1 and n_local * args.steps <= 2e8):
        s300, e300 = _fit_once(c, qMeans_, Xh.numpy(), init, 300, 99)
        fits.append({"max_iter": 300, "host_memory": "pageable", "fit_seconds": s300, "iter_per_s": e300.n_iter_ / s300,
                     "note": "the reference's default max_iter"})
    if args.mode != "ipe":
        # the reference's default E-step (true_distance_estimate=True): same fit, inner-product estimation per pair
        si, ei = _fit_once(c, qMeans_, Xh.numpy(), init, args.steps, 99, mode="ipe")
        fits.append({"max_iter": args.steps, "host_memory": "pageable", "mode": "ipe", "fit_seconds": si,
                     "iter_per_s": ei.n_iter_ / si, "note": "true_distance_estimate=True, the reference's default"})
    try:
        Xp = Xh.pin_memory()
        sp, ep = _fit_once(c, qMeans_, Xp.numpy(), init, args.steps, 99)
        fits.append({"max_iter": args.steps, "host_memory": "pinned", "fit_seconds": sp, "iter_per_s": ep.n_iter_ / sp})
        out["pinned"] = {"value": ep.n_iter_ / sp, "fit_seconds": sp}
        del Xp
    except Exception as exc:
        out["pinned"] = {"error": "%s: %s" % (type(exc).__name__, exc)}
    out["fits"] = fits
    return out


def config3_csv(c, path, qMeans_):
    """BASELINE.json config 3: q-means k = 15 on a CSV (numeric columns, standardised), delta 0.1 ... 1.0."""
    import warnings
    import pandas as pd
    df = pd.read_csv(path)
    Xc = df.select_dtypes(include=[np.number]).to_numpy(dtype=np.float64)
    Xc = Xc[np.isfinite(Xc).all(axis=1)]
    sd = Xc.std(axis=0)
    Xc = (Xc - Xc.mean(axis=0)) / np.where(sd > 0, sd, 1.0)
    rs = np.random.RandomState(0)
    init = Xc[rs.permutation(len(Xc))[:15]].copy()
    out = {"path": path, "shape": list(Xc.shape), "k": 15, "runs": []}
    for delta in [round(0.1 * i, 1) for i in range(1, 11)]:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            c.Q.seed(5)
            t0 = time.perf_counter()
            est = qMeans_(n_clusters=15, init=init, n_init=1, max_iter=10, tol=0, delta=delta, algorithm='full',
                          intermediate_error=True, true_distance_estimate=False).fit(Xc)
            c.torch.cuda.synchronize()
            out["runs"].append({"delta": delta, "fit_seconds": time.perf_counter() - t0, "n_iter": est.n_iter_,
                                "inertia": est.inertia_})
    return out


def main():
    args = parse()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
