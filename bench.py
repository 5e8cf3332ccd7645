#!/usr/bin/env python
"""Benchmark of the q-means hot path (BASELINE.json config 4: 10M x 128, k = 256).

  python bench.py --gpus N --steps K --warmup W          # this repo (CUDA, sm_100a)
  python bench.py --impl reference --steps K --warmup W  # CPU arm: oracle port of the reference

A step = one Lloyd iteration of qMeans_ (E-step with the delta-window noise
model, M-step, all-reduce of centroid sums, default-mode centroid tomography
at delta/2, centre shift, read-back of inertia and shift).  The 10M rows are
sharded over the N ranks (strong scaling); inputs are resident in HBM for
`value`; `e2e` goes through the public API (qMeans_.fit on host arrays).
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
    ap.add_argument("--no-extra", action="store_true")
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
    """Blobs: 256 centres ~ N(0, 4^2 I), unit noise, seed 1234.  Generated
    on the device in chunks; rank r holds the contiguous rows [r*n/G, (r+1)*n/G)."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(1234)
    centres = torch.randn((K_CLUSTERS, DIM), generator=g, device=device, dtype=torch.float64) * 4.0
    lo = rows * rank // world
    hi = rows * (rank + 1) // world
    n = hi - lo
    X = torch.empty((n, DIM), dtype=torch.float64, device=device)
    g.manual_seed(1234 + 7919 * (rank + 1))
    step = 1 << 20
    for a in range(0, n, step):
        b = min(n, a + step)
        lab = torch.randint(0, K_CLUSTERS, (b - a,), generator=g, device=device)
        X[a:b] = centres[lab]
        X[a:b] += torch.randn((b - a, DIM), generator=g, device=device, dtype=torch.float64)
    return X


def measure_dgemm_peak(device):
    """fp64 tensor-core denominator: cuBLAS DGEMM 8192^3, best of 5 (MEASURED_PEAKS.json has no
    fp64 entry; same protocol as its bf16 line)."""
    import torch
    n = 8192
    a = torch.randn((n, n), dtype=torch.float64, device=device)
    b = torch.randn((n, n), dtype=torch.float64, device=device)
    best = 0.0
    for _ in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(a, b)
        e1.record()
        torch.cuda.synchronize()
        best = max(best, 2.0 * n ** 3 / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    del a, b
    return best


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
        labels, _, _ = oq.labels_estimation(X, C, delta, tde, rng)
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
    return {"value": 1.0 / full, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": "%d of %d rows per step (E+M-step %.3f s, measured) + 256x128 centroid tomography (%.3f s, "
                      "measured); full-size step time extrapolated linearly in rows: %.1f s"
                      % (cpu_rows, total_rows, tr, tt, full),
            "host_cpus": os.cpu_count()}


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
        dist.init_process_group("nccl", device_id=device)
    assert world == args.gpus or world == 1, "launch with torchrun --nproc-per-node == --gpus"

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

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
    # centre the data as fit() does, then take the initial centres from rank 0's rows
    csum = eng.col_sums()
    if world > 1:
        dist.all_reduce(csum)
    mean = csum / float(eng.n_total)
    eng.subtract_rowvec(mean)
    eng.refresh_norms()
    g = torch.Generator(device="cpu")
    g.manual_seed(1234)
    idx = torch.randperm(min(n_local, 1 << 22), generator=g)[:K_CLUSTERS].to(device)
    centers = X[idx].clone()
    if world > 1:
        dist.broadcast(centers, src=0)
    c_old, c_new = torch.empty_like(centers), torch.empty_like(centers)
    host = torch.empty((2,), dtype=torch.float64).pin_memory()

    def step():
        return D._lloyd_iteration(eng, centers, c_old, c_new, host, delta, inter, True, True, None, tde)

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
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    launches = eng.launches - l0
    # tomography launches are made outside the engine
    launches += args.steps if inter else 0
    value = args.steps / (ms_total * 1e-3)

    # ---- roofline leg: the E-step kernel alone, CUDA events inside the C ABI ----
    tot, cnt = (_cabi.ctypes.c_double)(), (_cabi.ctypes.c_longlong)()
    _cabi.lib.sq_profile_assign(2, None, None)
    for _ in range(max(3, min(args.steps, 10))):
        D._labels_inertia(eng, centers, delta=delta, pool=None, true_distance_estimate=tde)
    torch.cuda.synchronize()
    _cabi.lib.sq_profile_assign(0, _cabi.ctypes.byref(tot), _cabi.ctypes.byref(cnt))
    k_ms = tot.value / max(cnt.value, 1)
    flops = 2.0 * n_local * K_CLUSTERS * DIM
    peak = measure_dgemm_peak(device) if rank == 0 else 0.0
    achieved = flops / (k_ms * 1e-3) / 1e12
    traffic, traffic_src = None, None
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            tj = json.load(f)
        traffic = tj["dram_bytes_per_row"] * n_local       # dram__bytes_read.sum + dram__bytes_write.sum, per launch
        traffic_src = tj["source"]
    except Exception:
        pass
    roofline = {"bound": "tensor", "kernel": "qmeans_assign_kernel (fp64 DMMA.8x8x4)", "achieved": achieved,
                "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak if peak else None, "traffic": traffic,
                "traffic_source": traffic_src,
                "kernel_ms": k_ms, "algorithmic_flops_per_launch": flops,
                "algorithmic_bytes_per_launch": 8.0 * n_local * DIM + 4.0 * n_local,
                "peak_source": "cuBLAS DGEMM 8192^3 fp64 measured in this run (MEASURED_PEAKS.json has no fp64 entry)",
                "hbm_fraction": (8.0 * n_local * DIM + 4.0 * n_local) / (k_ms * 1e-3) / 1e9 / load_peaks()["hbm_gbs"]}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "q-means Lloyd iteration, %dx%d fp64 rows sharded over %d GPU(s), k=%d, delta=%g, "
                                   "%s E-step, M-step, centroid tomography (incremental+stop, delta/2)"
                                   % (args.rows, DIM, world, K_CLUSTERS, DELTA, mode),
                       "rows_per_gpu": n_local, "mode": mode,
                       "l2": "inputs (%.2f GB per GPU) exceed the 126 MB L2; no flush needed" % (n_local * DIM * 8 / 1e9),
                       "last_inertia": inertia, "last_shift": shift},
            "clocks": clk.summary(), "gpu_launches": launches, "roofline": roofline}

    # ---- extra, rank 0 at N=1: other E-step modes and the tomography sweep ----
    if not args.no_extra and world == 1:
        line["extra"] = extra_measurements(eng, centers, D, Q, _cabi, torch)
    # ---- e2e through the public API: qMeans_.fit on pinned host arrays ----
    if not args.no_e2e:
        try:
            line["e2e"] = e2e_fit(args, X, mean, centers, eng, qMeans_, Q, torch, dist, world, device)
        except Exception as exc:      # e.g. the host cannot pin a copy of the shard
            line["e2e"] = {"value": None, "unit": UNIT, "error": "%s: %s" % (type(exc).__name__, exc)}
    if rank == 0 and world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_iteration_rate(args.cpu_rows, 2, 1, mode, args.rows)
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)
    except Exception:
        return {"hbm_gbs": 6650.0, "source": "fallback"}


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


def extra_measurements(eng, centers, D, Q, _cabi, torch):
    """Per-phase device times of one Lloyd iteration and the batched-tomography sweep
    (BASELINE.json config 5); explanatory numbers, not the headline."""
    from sqlearn.QuantumUtility import Utility as U
    out = {}
    for name, (delta, tde) in {"exact": (0.0, True), "window": (DELTA, False), "ipe": (DELTA, True)}.items():
        out["estep_%s_ms" % name] = _timed(
            torch, lambda: D._labels_inertia(eng, centers, delta=delta, pool=None, true_distance_estimate=tde),
            2 if name == "ipe" else 3)
    # IPE mode samples, for every (row, centroid) pair, the median of 13 amplitude-estimation outcomes
    pairs = float(eng.n) * float(eng.k)
    out["ipe_pairs_per_s"] = pairs / (out["estep_ipe_ms"] * 1e-3)
    out["ipe_reference_draws_per_s"] = 13.0 * pairs / (out["estep_ipe_ms"] * 1e-3)
    D._labels_inertia(eng, centers, delta=0.0, pool=None, true_distance_estimate=True)
    out["mstep_accumulate_ms"] = _timed(torch, eng.accumulate, 3)
    c2 = torch.empty_like(centers)
    eng.finalize(centers, c2)
    out["centroid_tomography_ms"] = _timed(torch, lambda: U.tomography(c2, DELTA / 2), 3)
    # the reference's default E-step (true_distance_estimate=True) in the same iteration, from the phases above
    out["lloyd_iterations_per_s_ipe_estep_from_phases"] = 1e3 / (
        out["estep_ipe_ms"] + out["mstep_accumulate_ms"] + out["centroid_tomography_ms"])
    sweep = []
    g = torch.Generator(device=centers.device)
    g.manual_seed(1234)
    for B, d, delta, incremental in [(4096, 1000, 0.1, False), (4096, 1000, 0.1, True), (65536, 128, 0.25, True),
                                     (64, 4096, 0.05, False), (70000, 61, 0.1, True), (64, 100000, 0.1, False),
                                     (1, 1000000, 0.01, False), (1, 1000000, 0.5, True)]:
        V = torch.rand((B, d), generator=g, device=centers.device, dtype=torch.float64) * 2 - 1
        V /= V.norm(dim=1, keepdim=True)
        res = {}

        def run():
            res["r"] = U._tomography_device(V, delta, None, 'L2', True, incremental, 0)
        ms = _timed(torch, run, 2)
        n_used = float(res["r"][1].sum().item())
        sweep.append({"B": B, "d": d, "delta": delta, "mode": "incremental+stop" if incremental else "single-shot",
                      "ms": ms, "vectors_per_s": B / (ms * 1e-3), "simulated_measurements_per_s": 2 * n_used / (ms * 1e-3),
                      "algorithmic_GBps": 16.0 * B * d / (ms * 1e-3) / 1e9})
    out["tomography_sweep"] = sweep
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
    # kernels alone
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


def e2e_fit(args, X, mean, centers_init, eng, qMeans_, Q, torch, dist, world, device):
    """The same iterations through the public API on HOST arrays: qMeans_.fit(X_host) with an
    explicit init and max_iter = steps.  Timed region: H2D of X from pinned memory, the fit
    preamble (eta, mu, condition number, centring), the Lloyd iterations, the final E-step and
    the D2H of labels / centres."""
    import warnings
    mode = args.mode
    X.add_(mean)                                   # back to the uncentred data the user would hold
    Xh = torch.empty(X.shape, dtype=torch.float64, pin_memory=True)
    Xh.copy_(X)
    init = (centers_init + mean).cpu().numpy()
    n_local = X.shape[0]
    del eng
    est = qMeans_(n_clusters=K_CLUSTERS, init=init, n_init=1, max_iter=args.steps, tol=0,
                  delta=0 if mode == "exact" else DELTA, algorithm='full', intermediate_error=mode != "exact",
                  true_distance_estimate=mode == "ipe", copy_x=False)
    # one small untimed fit first: library start-up (cuSOLVER handle, module load) is not the workload
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        qMeans_(n_clusters=K_CLUSTERS, init=init, n_init=1, max_iter=1, tol=0, delta=0, algorithm='full',
                copy_x=False).fit(Xh[:min(n_local, 200000)].numpy())
    Q.seed(99)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        est.fit(Xh.numpy())
    e1.record()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    if world > 1:
        dist.barrier()
    ms = torch.tensor([max(e0.elapsed_time(e1), wall * 1e3)], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    secs = float(ms.item()) * 1e-3
    return {"value": est.n_iter_ / secs, "unit": UNIT,
            "h2d_bytes_per_step": (n_local * DIM * 8 + K_CLUSTERS * DIM * 8) / est.n_iter_,
            "d2h_bytes_per_step": (n_local * 4 + K_CLUSTERS * DIM * 8 + 16 * est.n_iter_) / est.n_iter_,
            "api": "qMeans_(n_clusters=256, init=<array>, n_init=1, max_iter=%d, tol=0, algorithm='full', ...)"
                   ".fit(X_host)" % args.steps,
            "fit_seconds": secs, "n_iter": est.n_iter_, "inertia": est.inertia_,
            "note": "one fit = H2D of the shard + preamble + n_iter Lloyd iterations + final E-step; "
                    "bytes are totals divided by n_iter"}


def main():
    args = parse()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
