#!/usr/bin/env python
"""bench.py -- PFML date-steps/sec (r_tilde + denom, N=500, p=1024) on N B200s.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--dates B]

A "step" is one batch of B date-steps per GPU (one iteration each of the loop PFML_Input_Data.py:318-491: Sigma,
Lambda, m_func, standardise, chains, solves, r_tilde / risk / tc / denom) WITH the expanding-window sums of
PFML_Search_Coef.py:77-121 fused into the a9 finish kernel and, at N > 1, the FP64 NCCL all-reduce of those sums
(pfml_comm_allreduce_sums, the one exchange step of the path) inside the timed region.  Dates are independent, so every
rank runs its own B dates per step (weak scaling).

Printed JSON line (rank 0): value = whole-job date-steps/s with inputs resident in HBM; e2e = the same through
pfml_date_steps_host_packed (pinned HOST buffers in and out, H2D/D2H inside the timed region; the symmetric outputs
travel as packed upper triangles, lossless); roofline = the dominant kernel (the FP64-equivalent batched GEMM on the
INT8 tensor cores) with SURVEY 8(d)'s algorithmic flops against the INT8 tensor-pipe ceiling MEASURED in this run
(pfml_int8_mma_probe), the cuBLAS DGEMM peak measured beside it; hp_grid = the whole C3 + C4 pipeline (360 months x 3 g
date-steps -> window sums -> all-reduce -> ridge grid dealt over ranks -> beta all-gather) wall-time at N GPUs next to the
CPU arm; hbm_rows = the HBM-bound kernels (C2 covariance estimator, RFF) in GB/s; cpu_baseline = the numpy oracle (a
port of the reference's algorithm) on the host cores.  --impl reference times only that CPU arm.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "oracle")):
    if _p not in sys.path:
        sys.path.insert(0, _p)
PKG = "jkmp22-machine-learning-and-the-implementable-efficient-frontier-replication_b200"

N_STOCKS, P_FEAT = 500, 1024
METRIC = "PFML date-steps/sec (r_tilde+denom, N=500, p=1024)"
UNIT = "date-steps/s"
WORKLOAD = "PFML date-step N=500 p=1024 (P=1025), lb=11, 10 m_func iterations (BASELINE configs[2] at the metric's p)"


def official_flops_per_date(n=N_STOCKS, P=P_FEAT + 1):
    """SURVEY.md section 8(d), 'Date-step total (reduced, official)'."""
    a9 = 2 * n * n * P + 2 * n * P * P + 2 * n * P
    chain = 22 * n ** 3 + 24 * n * n * P
    lu = 2 * (2 / 3 * n ** 3 + 2 * n * n * P)
    mf = 23 * n ** 3
    return dict(total=a9 + chain + lu + mf, denom=a9, chain=chain, solve=lu, m_func=mf)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for k, nm in enumerate(names):
                if len(r) > 4 + k and r[4 + k].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max((float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "").isdigit()), default=None),
                "samples": len(sm), "reasons": sorted(reasons)}


def use_all_host_threads():
    """torchrun exports OMP_NUM_THREADS=1; the CPU arm must use every host core (BLAS threads = os.cpu_count())."""
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(limits=os.cpu_count())
    except Exception:
        pass
    return os.cpu_count()


def make_bench_panel(syn, B, seed):
    return syn.make_panel(N=N_STOCKS, T=syn.LB + 2 + B, p=P_FEAT, seed=seed, g=float(np.exp(-3.0)))


def cpu_dates_per_sec(syn, budget_s=20.0, max_dates=8, seed=101):
    """Oracle (numpy port of the reference algorithm, all host BLAS threads) on a bounded sample of the workload."""
    import pfml_oracle as O
    use_all_host_threads()
    panel = make_bench_panel(syn, max_dates, seed)
    T, N, F = panel.feats.shape
    rff_x = O.rff(panel.feats.reshape(T * N, F), panel.W).reshape(T, N, -1)      # a1, excluded like on the GPU side
    vs = O.vol_scale_panel(panel)                                               # a2
    hot = list(panel.hot_months())
    t0 = time.perf_counter()
    done = 0
    for t in hot:
        O.date_step(panel, int(t), rff_x, vs)
        done += 1
        if time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    return done / dt, done, dt


def run_reference(args, rank, world):
    if rank != 0:
        return
    syn = importlib.import_module(PKG + ".synthetic")
    cores = use_all_host_threads()
    per_step = 2
    import pfml_oracle as O
    panel = make_bench_panel(syn, per_step, 101)
    T, N, F = panel.feats.shape
    rff_x = O.rff(panel.feats.reshape(T * N, F), panel.W).reshape(T, N, -1)
    vs = O.vol_scale_panel(panel)
    hot = [int(t) for t in panel.hot_months()]

    def step():
        for t in hot:
            O.date_step(panel, t, rff_x, vs)

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    val = per_step * args.steps / dt
    sample = f"{per_step} date-steps per step of the N=500, p=1024 workload (panel-level RFF / vol_scale excluded, as on the GPU arm)"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD}, "details": {"dates_per_step": per_step},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference algorithm restated in numpy/scipy (oracle/pfml_oracle.py), OpenBLAS threads = all host cores; "
                "the reference's pandas per-column standardisation is replaced by vectorised numpy, so this CPU arm is "
                "faster than the reference script itself"}))


def algorithmic_flops_emu_per_date(n=N_STOCKS, P=P_FEAT + 1):
    """SURVEY 8(d) work of the launches that run on the INT8 tensor cores: 11 chain products 2n^3, 14 n x n x P products
    2n^2 P (12 sums + 2 solves) and the two symmetric a9 contractions n P^2 each (only the upper triangle is formed)."""
    return 11 * 2 * n ** 3 + 14 * 2 * n * n * P + 2 * n * P * P


C3_MONTHS, C3_G = 360, 3
C4_P_VEC = [64, 128, 256, 512, 1024]
C4_L_VEC = [0.0] + list(np.exp(np.linspace(-10, 10, 14)))


def hp_grid_pipeline_walltime(torch, dist, dev, comm, rank, world, batch):
    """hp-grid wall-time (BASELINE metric, second half; configs[2] + configs[3]): N = 500, 360 months, 3 bandwidths g,
    p_max = 1024, 15 lambda x p in {64..1024} x every complete yearly refit -- host panel in, beta on the host out:
    date-steps sharded over the ranks with the window sums fused in, NCCL all-reduce of the sums, ridge problems dealt over
    the ranks, beta all-gathered (strong scaling: the work is fixed, ranks split it)."""
    syn = importlib.import_module(PKG + ".synthetic")
    pl = importlib.import_module(PKG + ".pfml_pipeline")
    panel = syn.make_panel(N=N_STOCKS, T=syn.LB + 2 + C3_MONTHS, p=P_FEAT, seed=303, g=float(np.exp(-3.0)), start="1990-01")
    months = [int(t) for t in panel.hot_months()]
    me = panel.month_ends[months]
    y0 = int(str(me[0])[:4])
    years = list(range(y0 + 2, int(str(me[-1])[:4]) + 1))                # burn-in = the first calendar year (11 months)
    rng = np.random.default_rng(304)
    Ws = [np.sqrt(g) * rng.standard_normal(panel.W.shape) for g in np.exp([-3.0, -2.0, -1.0])]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    tm = {}
    pl.hp_grid_pipeline(panel, Ws[:1], months[:min(len(months), 40)], years[:1], C4_P_VEC, C4_L_VEC, comm=comm, device=dev, batch=batch)  # warm-up
    barrier()
    t0 = time.perf_counter()
    coef = pl.hp_grid_pipeline(panel, Ws, months, years, C4_P_VEC, C4_L_VEC, comm=comm, device=dev, batch=batch, timing=tm)
    barrier()
    wall = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(wall, op=dist.ReduceOp.MAX)
    out = {"workload": f"C3 + C4: N={N_STOCKS}, {len(months)} months x {C3_G} g date-steps (p={P_FEAT}) -> {len(years)} yearly refits x "
                       f"{len(C4_L_VEC)} lambda x p in {C4_P_VEC}",
           "n_gpus": world, "scaling": "strong", "wall_s": float(wall.item()), "date_steps": len(months) * C3_G,
           "ridge_solves": len(years) * len(C4_L_VEC) * len(C4_P_VEC) * C3_G,
           "phases_rank0_s": {k: round(v, 4) for k, v in tm.items() if k.endswith("_s")},
           "allreduce_bytes": tm.get("allreduce_bytes", 0),
           "includes": "H2D of this rank's panel slice, RFF + vol_scale, date-steps with fused window sums, NCCL all-reduce, running "
                       "sums, gather of the pfml_feat_fun(p) sub-matrices, SPD inverses + refinement, beta all-gather, D2H of beta"}
    if world > 1 and rank == 0:
        # sharded vs unsharded on the first bandwidth: FP64 summation order only
        one = pl.hp_grid_pipeline(panel, Ws[:1], months, years, C4_P_VEC, C4_L_VEC, comm=None, device=dev, batch=batch)
        worst = 0.0
        for y in years:
            for p in C4_P_VEC:
                for j in range(1, len(C4_L_VEC)):
                    worst = max(worst, float(np.max(np.abs(coef[0][y][p][j] - one[0][y][p][j])) / np.max(np.abs(one[0][y][p][j]))))
        out["beta_max_rel_diff_vs_1gpu_lambda_gt_0"] = worst
    out["_sizes"] = (len(months), len(years))
    return out


def hp_grid_cpu_arm(out):
    """CPU arm of hp_grid on a bounded sample, extrapolated linearly (date-steps and refits are independent): numpy oracle, all
    BLAS threads.  Rank 0 only, AFTER every collective of the run (the other ranks sleep meanwhile)."""
    syn = importlib.import_module(PKG + ".synthetic")
    n_months, n_years = out.pop("_sizes")
    months, years = range(n_months), range(n_years)
    rng = np.random.default_rng(306)
    import pfml_oracle as O
    use_all_host_threads()
    small = make_bench_panel(syn, 10, 305)
    T, N, F = small.feats.shape
    rff_x = O.rff(small.feats.reshape(T * N, F), small.W).reshape(T, N, -1)
    vs = O.vol_scale_panel(small)
    t0 = time.perf_counter()
    for t in small.hot_months():
        O.date_step(small, int(t), rff_x, vs)
    cpu_date_s = (time.perf_counter() - t0) / len(small.hot_months())
    Pn = P_FEAT + 1
    Tr = 36
    r_t = rng.standard_normal((Tr, Pn)) * 1e-3
    A = rng.standard_normal((Tr, Pn, 4)) * 0.2
    dn = np.einsum("tik,tjk->tij", A, A)
    dn[:, np.arange(Pn), np.arange(Pn)] += 1e-2
    me_s = syn.month_end_range("1990-01", Tr)
    t0 = time.perf_counter()
    O.search_coef(r_t, dn, me_s, [1993], C4_P_VEC, C4_L_VEC)
    cpu_year_s = time.perf_counter() - t0
    cpu_total = cpu_date_s * len(months) * C3_G + cpu_year_s * len(years) * C3_G
    out["cpu_port"] = {"cores": os.cpu_count(), "s_per_date_step": cpu_date_s, "s_per_yearly_refit": cpu_year_s,
                       "wall_s_extrapolated": cpu_total,
                       "sample": f"{len(small.hot_months())} date-steps and 1 yearly refit (all p, all lambda) timed, extrapolated to "
                                 f"{len(months) * C3_G} date-steps and {len(years) * C3_G} refits"}
    out["speedup_vs_cpu_port"] = cpu_total / out["wall_s"]
    out["north_star_target"] = ">= 50x the reference's host-CPU time at 8 GPUs"
    return out


def c5_universe(torch, dist, dev, lib, rank, world, dates=8):
    """BASELINE configs[4] (stress universe): N = 3000 stocks (padded 3008), p = 1024 -- `dates` date-steps per rank, timed on
    the device (max over ranks), with the INT8 result of one date checked against the FP64 DMMA path of the same library.
    The oracle comparison at this universe size is a test (tests/test_baseline_configs_gpu.py, p = 64: one oracle date-step
    takes a minute of CPU)."""
    syn = importlib.import_module(PKG + ".synthetic")
    dm = importlib.import_module(PKG + ".device")
    panel = syn.make_panel(N=3000, T=syn.LB + 2 + dates, p=P_FEAT, seed=500 + rank, ragged=0.01)
    months = [int(t) for t in panel.hot_months()]
    dp = dm.DevicePanel(panel, dev)
    sel, n_sel, n_max, NP = dm.universe(panel, months)
    r = dm.DateStepRunner(dp, dates, n_max, want_signal=False)
    r.set_batch(months, sel, n_sel)
    r.run(); r.run()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); r.run(); r.run(); e1.record()
    torch.cuda.synchronize(dev)
    ms = torch.tensor([e0.elapsed_time(e1) / 2], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ok = int(r.status.abs().sum().item()) == 0
    d1 = r.denom[0].clone()
    lib.pfml_set_emulation(0)
    r.run()
    torch.cuda.synchronize(dev)
    lib.pfml_set_emulation(1)
    diff = float(((d1 - r.denom[0]).abs().max() / r.denom[0].abs().max()).item())
    n_eff = int(n_sel.max())
    P = P_FEAT + 1
    flops = (23 + 22 + 4 / 3) * n_eff ** 3 + (24 + 4 + 2) * n_eff ** 2 * P + 2 * n_eff * P * P     # SURVEY 8(d) formulas at this n
    out = {"workload": f"N=3000 (n = {n_eff}, padded {NP}), p={P_FEAT}, {dates} date-steps per GPU", "n_gpus": world,
           "ms_per_date": float(ms.item()) / dates, "value": world * dates / (float(ms.item()) * 1e-3), "unit": UNIT,
           "workspace_gb_per_gpu": r.ws_bytes / 1e9, "status_ok": ok, "int8_vs_dmma_denom_rel_diff": diff,
           "algorithmic_tflops_per_gpu": flops * dates / (float(ms.item()) * 1e-3) * 1e-12}
    del r, dp
    torch.cuda.empty_cache()
    return out


def hbm_rows(torch, dev, lib):
    """The HBM-bound rows of the path (north_star item 5): achieved GB/s = algorithmic bytes / CUDA-event time.
    C2: 500 stocks x 5 040 trading days, K = 25 factors (a15 daily OLS, a16 factor covariance, a17 EWMA vol); a1: RFF of the
    C3 panel (186 500 rows x 115 characteristics -> 1 024 features)."""
    import ctypes as C
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = float(peaks.get("hbm_gbs", 6500.0))
    ptr = lambda t: C.c_void_p(t.data_ptr())
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def timed(fn, reps=int(os.environ.get("PFML_HBM_REPS", "5"))):
        fn(); fn()
        best = 1e30
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize(dev)
            best = min(best, e0.elapsed_time(e1))
        return best

    out = {"hbm_peak_gbs": hbm, "hbm_peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6500 GB/s"}
    g = torch.Generator(device=dev).manual_seed(1)
    D, n, K = 5040, 500, 25
    rows = D * n
    X = torch.randn(rows, K, dtype=torch.float64, device=dev, generator=g)
    X[:, :12] = 0.0
    X[torch.arange(rows, device=dev), torch.randint(0, 12, (rows,), device=dev, generator=g)] = 1.0      # one-hot industries
    y = 0.01 * torch.randn(rows, dtype=torch.float64, device=dev, generator=g)
    day_ptr = torch.arange(0, rows + 1, n, dtype=torch.int64, device=dev)
    coef = torch.empty(D, K, dtype=torch.float64, device=dev)
    resid = torch.empty(rows, dtype=torch.float64, device=dev)
    status = torch.zeros(D, dtype=torch.int32, device=dev)
    ms = timed(lambda: lib.pfml_xs_ols_daily(ptr(X), ptr(y), ptr(day_ptr), D, K, ptr(coef), ptr(resid), ptr(status), st))
    by = rows * (K + 2) * 8 + D * K * 8                     # X and y read once, residuals written, coefficients written
    out["a15_xs_ols_daily"] = {"ms": ms, "algorithmic_bytes": by, "gbs": by / ms * 1e-6, "frac_of_hbm_peak": by / ms * 1e-6 / hbm,
                               "days": D, "stocks_per_day": n, "K": K}
    n_calc, obs = 240, 2520
    ends = torch.linspace(obs - 1, D - 1, n_calc, device=dev).to(torch.int32)
    wc = torch.as_tensor((0.5 ** (1 / 378)) ** np.arange(obs, 0, -1)).to(dev)
    wv = torch.as_tensor((0.5 ** (1 / 126)) ** np.arange(obs, 0, -1)).to(dev)
    cov = torch.empty(n_calc, K, K, dtype=torch.float64, device=dev)
    ms = timed(lambda: lib.pfml_wcov_wcor(ptr(coef), ptr(ends), n_calc, K, obs, ptr(wc), ptr(wv), 21.0, ptr(cov), st))
    by = n_calc * (obs * K * 8 + 2 * obs * 8)
    out["a16_wcov_wcor"] = {"ms": ms, "algorithmic_bytes": by, "gbs": by / ms * 1e-6, "calc_dates": n_calc, "window": obs,
                            "note": "the D x K factor-return matrix (1 MB) stays in L2: bound by L2 / the K x K accumulation, not HBM"}
    series_ptr = torch.arange(0, rows + 1, D, dtype=torch.int64, device=dev)
    vol = torch.empty(rows, dtype=torch.float64, device=dev)
    lam = 0.5 ** (1 / 126)
    ms = timed(lambda: lib.pfml_ewma_vol(ptr(resid), ptr(series_ptr), n, C.c_double(lam), 63, ptr(vol), st))
    by = 2 * rows * 8
    out["a17_ewma_vol"] = {"ms": ms, "algorithmic_bytes": by, "gbs": by / ms * 1e-6, "series": n, "length": D,
                           "note": "20 MB in, 20 MB out; each series is a 5 040-step dependent recurrence (un-fused multiply and add "
                                   "like the reference's numba kernel): the floor is the dependency chain, about 16 clk per step"}
    del X, y, resid, vol
    rows_c3, F, ph = N_STOCKS * (C3_MONTHS + 13), 115, P_FEAT // 2
    Xf = torch.rand(rows_c3, F, dtype=torch.float64, device=dev, generator=g)
    W = 0.22 * torch.randn(F, ph, dtype=torch.float64, device=dev, generator=g)
    o = torch.empty(rows_c3, 2 * ph, dtype=torch.float64, device=dev)
    ms = timed(lambda: lib.pfml_rff_sincos(ptr(Xf), ptr(W), rows_c3, F, ph, ptr(o), st))
    by = rows_c3 * (F + 2 * ph) * 8
    out["a1_rff_sincos"] = {"ms": ms, "algorithmic_bytes": by, "gbs": by / ms * 1e-6, "frac_of_hbm_peak": by / ms * 1e-6 / hbm,
                            "rows": rows_c3, "flops": 2.0 * rows_c3 * F * ph, "tflops": 2.0 * rows_c3 * F * ph / ms * 1e-9,
                            "note": "FP64 GEMM (K = 115) with a sincos epilogue: 1.53 GB written; the FP64 sin/cos of the epilogue, not "
                                    "HBM, paces it"}
    return out


def measure_fp64_peak(torch, dev):
    n = 8192
    a = torch.randn(n, n, dtype=torch.float64, device=dev)
    b = torch.randn(n, n, dtype=torch.float64, device=dev)
    for _ in range(2):
        a @ b
    torch.cuda.synchronize(dev)
    best = 1e30
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); a @ b; e1.record()
        torch.cuda.synchronize(dev)
        best = min(best, e0.elapsed_time(e1))
    del a, b
    return 2 * n ** 3 / best * 1e-9        # TFLOP/s


def run_ours(args, rank, world, local_rank):
    import ctypes as C
    import torch
    import torch.distributed as dist
    syn = importlib.import_module(PKG + ".synthetic")
    device_mod = importlib.import_module(PKG + ".device")
    host_mod = importlib.import_module(PKG + ".host")
    lib = importlib.import_module(PKG + "._lib").load()
    dev = torch.device(f"cuda:{local_rank}")
    torch.cuda.set_device(dev)
    B = args.dates

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    panel = make_bench_panel(syn, B, seed=7 + rank)
    months = [int(t) for t in panel.hot_months()]
    assert len(months) == B
    dpanel = device_mod.DevicePanel(panel, dev)
    sel, n_sel, n_max, NP = device_mod.universe(panel, months)
    runner = device_mod.DateStepRunner(dpanel, B, n_max)
    runner.set_batch(months, sel, n_sel)

    if args.gemm_only:
        # the dominant launch shape in isolation (Horner step: m~ [n x n] @ Z [n x P], batch = dates), for ncu --set full
        A = torch.randn(B, NP, NP, dtype=torch.float64, device=dev)
        Z = torch.randn(B, NP, 1032, dtype=torch.float64, device=dev)
        D = torch.empty(B, NP, 1032, dtype=torch.float64, device=dev)
        P = P_FEAT + 1
        evs = []
        for i in range(args.warmup + args.steps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rc = lib.pfml_dgemm_batched(0, 0, NP, P, NP, 1.0, C.c_void_p(A.data_ptr()), NP, NP * NP, C.c_void_p(Z.data_ptr()), 1032,
                                        NP * 1032, 0.0, C.c_void_p(D.data_ptr()), 1032, NP * 1032, B,
                                        C.c_void_p(torch.cuda.current_stream().cuda_stream))
            e1.record()
            assert rc == 0
            evs.append((e0, e1))
        torch.cuda.synchronize(dev)
        ms = min(a.elapsed_time(b) for a, b in evs[args.warmup:])
        # the same product on the INT8 tensor cores (slicing included / excluded)
        wsb = lib.pfml_dgemm_emulated_workspace_bytes(NP, P, NP, B)
        ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
        D2 = torch.empty_like(D)
        evs2 = []
        for i in range(args.warmup + args.steps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rc = lib.pfml_dgemm_emulated(NP, P, NP, 1.0, C.c_void_p(A.data_ptr()), NP, NP * NP, C.c_void_p(Z.data_ptr()), 1032, NP * 1032,
                                         0.0, C.c_void_p(D2.data_ptr()), 1032, NP * 1032, B, None, C.c_void_p(ws.data_ptr()), wsb,
                                         C.c_void_p(torch.cuda.current_stream().cuda_stream))
            e1.record()
            assert rc == 0
            evs2.append((e0, e1))
        torch.cuda.synchronize(dev)
        ms2 = min(a.elapsed_time(b) for a, b in evs2[args.warmup:])
        err2 = float((D2[0, :, :P] - D[0, :, :P]).abs().max() / D[0, :, :P].abs().max())
        print(json.dumps({"emulated_int8": True, "ms_incl_slicing": ms2, "fp64_equiv_tflops": 2.0 * B * NP * P * NP / ms2 * 1e-9,
                          "rel_diff_vs_dmma": err2}))
        ref = torch.matmul(A[0], Z[0, :, :P])
        err = float((D[0, :, :P] - ref).abs().max() / ref.abs().max())
        print(json.dumps({"gemm_only": True, "shape": [B, NP, P, NP], "ms": ms,
                          "executed_tflops": 2.0 * B * NP * P * NP / ms * 1e-9, "rel_err_vs_cublas": err}))
        return

    # ---- the step: B date-steps with the window sums fused in, then (N > 1) the all-reduce of the sums ----------
    sc = importlib.import_module(PKG + ".pfml_search_coef")
    comm_mod = importlib.import_module(PKG + ".comm")
    comm = comm_mod.Communicator.from_torch_distributed(device=dev) if world > 1 else None
    me = panel.month_ends[months]
    y0 = int(str(me[0])[:4])
    hp_years = list(range(y0 + 2, int(str(me[-1])[:4]) + 2))          # burn-in = first calendar year; every date lies in a window
    bucket_np = sc.rank_buckets(me, hp_years)
    assert (bucket_np >= 0).all()
    nbk = len(hp_years) + 1
    bucket = torch.as_tensor(bucket_np.astype(np.int32)).to(dev)
    Pn = P_FEAT + 1
    r_sum = torch.zeros((nbk, Pn), dtype=torch.float64, device=dev)
    d_sum = torch.zeros((nbk, Pn, Pn), dtype=torch.float64, device=dev)
    cnt = torch.as_tensor(np.bincount(bucket_np, minlength=nbk).astype(np.float64)).to(dev)
    cnt_step = cnt.clone()
    allreduce_bytes = int(r_sum.numel() + d_sum.numel() + cnt.numel()) * 8 if world > 1 else 0

    def step():
        r_sum.zero_(); d_sum.zero_(); cnt_step.copy_(cnt)
        runner.run_bucketed(bucket, nbk, r_sum, d_sum)
        if comm is not None:
            comm.allreduce_sums(r_sum, d_sum, cnt_step)

    for _ in range(args.warmup):
        step()
    if args.profile_only:
        for _ in range(args.steps):
            step()
        torch.cuda.synchronize(dev)
        print(json.dumps({"profile_only": True, "dates": B, "steps": args.steps, "launches_per_step":
                          lib.pfml_launch_count() // (args.warmup + args.steps)}))
        return
    barrier()
    sampler = ClockSampler(local_rank).start() if rank == 0 else None
    l0 = lib.pfml_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    launches = lib.pfml_launch_count() - l0
    clocks = sampler.stop() if sampler else None
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    value = world * B * args.steps / (ms_total * 1e-3)
    assert int(runner.status.abs().sum().item()) == 0, "numerical failure flag set"
    # the fused sums against a separate pass over the outputs (bit-exact), and the all-reduced count
    r_chk, d_chk = sc.bucket_sums(runner.r_tilde, runner.denom, bucket_np, nbk)
    if world == 1:
        assert torch.equal(d_chk, d_sum) and torch.equal(r_chk, r_sum), "fused window sums differ from the separate pass"
    assert float(cnt_step.sum().item()) == world * B
    # the all-reduce alone (same buffers), for the record
    ar_ms = None
    if comm is not None:
        barrier()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        for _ in range(5):
            comm.allreduce_sums(r_sum, d_sum, cnt_step)
        a1.record()
        barrier()
        t = torch.tensor([a0.elapsed_time(a1) / 5], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ar_ms = float(t.item())

    # ---- the opt-in 6-digit mode of the INT8 path (pfml_set_emulation(2)), same steps: reported beside the default ----
    lib.pfml_set_emulation(2)
    step()
    barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    for _ in range(args.steps):
        step()
    f1.record()
    barrier()
    lib.pfml_set_emulation(1)
    ms6 = torch.tensor([f0.elapsed_time(f1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms6, op=dist.ReduceOp.MAX)
    value_six = world * B * args.steps / (float(ms6.item()) * 1e-3)

    # ---- three bandwidths g of the same dates in one call (BASELINE configs[2] has 3 g; quirk Q10: m, the chains and the
    # two inverses are shared over g): (date, g) pairs per second ------------------------------------------------------
    rng3 = np.random.default_rng(9)
    dp3 = device_mod.DevicePanel(panel, dev, extra_W=[np.sqrt(g_) * rng3.standard_normal(panel.W.shape) for g_ in np.exp([-2.0, -1.0])])
    mg = device_mod.MultiGRunner(dp3, B, n_max)
    mg.set_batch(months, sel, n_sel)
    r3 = torch.zeros((3, nbk, Pn), dtype=torch.float64, device=dev)
    d3 = torch.zeros((3, nbk, Pn, Pn), dtype=torch.float64, device=dev)
    mg.run(bucket, nbk, r3, d3)
    barrier()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record()
    for _ in range(args.steps):
        mg.run(bucket, nbk, r3, d3)
    g1.record()
    barrier()
    ms3 = torch.tensor([g0.elapsed_time(g1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms3, op=dist.ReduceOp.MAX)
    value_3g = world * 3 * B * args.steps / (float(ms3.item()) * 1e-3)
    ms_3g = float(ms3.item()) / args.steps
    assert int(mg.status.abs().sum().item()) == 0
    step()                                                 # the single-g step in the default mode again (the last one ran in mode 2)
    torch.cuda.synchronize(dev)
    same_g0 = bool(torch.equal(mg.denom[0], runner.denom)) and bool(torch.equal(mg.r_tilde[0], runner.r_tilde))
    del mg, dp3, r3, d3
    torch.cuda.empty_cache()

    # ---- live per-kernel accounting (CUDA events around every launch group, same steps) ------------
    nt = lib.pfml_prof_ntags()
    pms, pfl, pln = (np.zeros(nt) for _ in range(3))
    lib.pfml_prof_enable(1)
    for _ in range(args.steps):
        runner.run_bucketed(bucket, nbk, r_sum, d_sum)
    lib.pfml_prof_collect(C.c_void_p(pms.ctypes.data), C.c_void_p(pfl.ctypes.data), C.c_void_p(pln.ctypes.data))
    lib.pfml_prof_enable(0)
    names = [lib.pfml_prof_tag_name(i).decode() for i in range(nt)]
    resident_ref = {k: getattr(runner, k)[0].cpu().numpy() for k in ("risk", "tc", "denom")}
    ws_gb = runner.ws_bytes / 1e9

    # ---- e2e: host buffers in / out through pfml_date_steps_host_packed ---------------------------------
    # pinned host buffers are placed on the NUMA node of the allocating thread: bind it to the GPU's node while they are
    # allocated (with 8 ranks the downloads otherwise cross the socket link), then give the affinity back
    numa = bind_to_gpu_numa(torch, dev)
    blk = host_mod.HostBlock(panel, months, pin=True)
    if numa is not None:
        os.sched_setaffinity(0, numa["before"])

    def e2e_rate(fn):
        for _ in range(max(1, min(args.warmup, 2))):
            fn()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            fn()
        torch.cuda.synchronize(dev)
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        assert int(blk.status.abs().sum()) == 0
        return world * B * args.steps / float(dt.item())

    e2e_val = e2e_rate(lambda: blk.run_packed(chunk=args.chunk, device=local_rank))
    risk0, tc0, den0 = blk.unpack(0)
    chk = max(float(np.max(np.abs(risk0 - resident_ref["risk"]))), float(np.max(np.abs(tc0 - resident_ref["tc"]))),
              float(np.max(np.abs(den0 - resident_ref["denom"]))))
    e2e_full = e2e_rate(lambda: blk.run(chunk=args.chunk, device=local_rank))
    chk_full = float(np.max(np.abs(blk.out["denom"][0].numpy() - resident_ref["denom"])))
    e2e_two = e2e_rate(lambda: blk.run(chunk=args.chunk, device=local_rank, full_outputs=False))
    h2d_bytes, d2h_packed, d2h_full = int(blk.h2d_bytes), int(blk.d2h_bytes), int(blk.d2h_bytes_full_squares)
    del blk, runner, dpanel
    torch.cuda.empty_cache()

    # ---- hp-grid wall-time: the whole C3 + C4 pipeline at N GPUs (all ranks take part) ---------------------
    hp = hp_grid_pipeline_walltime(torch, dist, dev, comm, rank, world, B) if not args.no_hp_grid else None     # GPU part (all ranks)
    c5 = c5_universe(torch, dist, dev, lib, rank, world) if not args.no_hp_grid else None
    if comm is not None:
        comm.close()
    if rank != 0:
        return
    # ---- roofline of the dominant kernel ---------------------------------------------------------------
    # pfml::emu_gemm_kernel: the n x n x P products of the sums / solves (14 launches per step), the n x n x n chain
    # products (11) and the two symmetric a9 contractions as FP64-equivalent GEMMs on the INT8 tensor cores.  Duration:
    # CUDA events around every launch (library profiling slot `emu_i8_gemm`) over the timed steps.  Work: SURVEY 8(d)'s
    # algorithmic flops (n = 500, not the padded 512; n P^2 per symmetric launch).  Ceiling: the INT8 tensor pipe MEASURED
    # here with back-to-back tcgen05.mma.kind::i8 (pfml_int8_mma_probe), divided by the 28 digit products per FP64 product.
    peak_fp64 = measure_fp64_peak(torch, dev)
    probe = {}
    for nm, n_, ts in (("pipe_peak_M128_N256_A_in_TMEM", 256, 1), ("N256_SS", 256, 0), ("kernel_shape_N64_A_in_TMEM", 64, 1),
                       ("kernel_shape_N64_SS", 64, 0)):
        tops, clk = C.c_double(), C.c_double()
        rc = lib.pfml_int8_mma_probe(n_, 8192, ts, C.byref(tops), C.byref(clk))
        probe[nm] = {"tops": tops.value, "nominal_clk_per_mma": clk.value} if rc == 0 else {"error": rc}
    int8_peak = probe["pipe_peak_M128_N256_A_in_TMEM"].get("tops", 0.0)
    n_pairs = 28
    # the kernel's own instruction stream per 32-byte k slab (three tcgen05.cp of A digits, 18 products with A in TMEM, 10 with
    # both operands in shared memory), issued back to back on resident operands: the ceiling of THIS decomposition
    mix = None
    clk_s, tops_s = C.c_double(), C.c_double()
    if lib.pfml_int8_pipe_probe(4096, 0, C.byref(clk_s), C.byref(tops_s)) == 0:
        mix = tops_s.value
        probe["kernel_k_slab_stream"] = {"tops": tops_s.value, "nominal_clk_per_k_slab": clk_s.value}
        c1, t1 = C.c_double(), C.c_double()
        if lib.pfml_int8_pipe_probe(4096, 1, C.byref(c1), C.byref(t1)) == 0:
            probe["k_slab_stream_with_tcgen05_st_stagers"] = {"tops": t1.value, "nominal_clk_per_k_slab": c1.value,
                                                              "note": "A digits 0..3 staged to TMEM by four extra warps instead of tcgen05.cp: no faster"}
    off = official_flops_per_date()
    k_emu = names.index("emu_i8_gemm")
    emu_ms = pms[k_emu] / args.steps
    emu_launches = pln[k_emu] / args.steps
    emu_alg = algorithmic_flops_emu_per_date() * B
    peak = int8_peak / n_pairs if int8_peak > 0 else None
    achieved = emu_alg / (emu_ms * 1e-3) * 1e-12 if emu_ms > 0 else 0.0
    gemm_ms = pms[0] / args.steps                      # all DMMA GEMM launches of one step
    total_stage_ms = sum(pms[1:7]) / args.steps
    traffic = None
    ncu_json = os.path.join(ROOT, "profiles", "ncu_summary.json")
    if os.path.exists(ncu_json):
        try:
            per_date = json.load(open(ncu_json)).get("emu_gemm_nxP_dram_bytes_per_launch_per_date")
            traffic = per_date * B if per_date else None
        except Exception:
            traffic = None
    n, P = N_STOCKS, P_FEAT + 1
    stage_map = {"m_func": off["m_func"], "chain_nxn": 22 * n ** 3, "horner_nxP": 24 * n * n * P, "solve": off["solve"], "denom": off["denom"]}
    stages = {}
    for i in range(1, 7):
        st_ms = pms[i] / args.steps
        stages[names[i]] = {"ms_per_step": round(st_ms, 3), "share": round(st_ms / total_stage_ms, 4)}
        if names[i] in stage_map:
            tf = stage_map[names[i]] * B / (st_ms * 1e-3) * 1e-12
            stages[names[i]]["algorithmic_tflops"] = round(tf, 2)
            stages[names[i]]["vs_cublas_dgemm_peak"] = round(tf / peak_fp64, 3)
    if peak:
        stages["denom"]["roofline"] = {"bound": "tensor", "achieved": stages["denom"]["algorithmic_tflops"], "peak": peak,
                                       "unit": "TFLOP/s", "frac": stages["denom"]["algorithmic_tflops"] / peak,
                                       "note": "whole a9 stage (graded contraction): stacked factor GEMMs on the DMMA pipe, two symmetric "
                                               "INT8 contractions, their operand slicing and the finish kernel with the fused window sums"}
    if hp is not None:
        hp_grid_cpu_arm(hp)
    hb = hbm_rows(torch, dev, lib) if world == 1 else None
    cpu_v, cpu_n, cpu_dt = cpu_dates_per_sec(syn)
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD},
        "details": {"dates_per_step_per_gpu": B,
                    "step": "B date-steps with the expanding-window sums fused into the a9 finish kernel"
                            + (f", then one FP64 NCCL all-reduce of the [{nbk}] x (P + P^2 + 1) sums over {world} ranks "
                               f"({allreduce_bytes} bytes per rank per step)" if world > 1 else " (one rank: no exchange)"),
                    "parallelism": f"dates sharded over {world} GPU(s); collective in the timed region: "
                                   + ("pfml_comm_allreduce_sums (NCCL, FP64 sum)" if world > 1 else "none at N = 1"),
                    "allreduce_ms_alone": ar_ms, "allreduce_bytes_per_step": allreduce_bytes,
                    "l2": "per-step working set ~%.1f GB >> 126 MB L2 (inputs larger than L2)" % ws_gb},
        "clocks": clocks,
        "value_six_digit_mode": {"value": value_six, "unit": UNIT,
                                 "note": "pfml_set_emulation(2): 6 of 7 digits for the non-compounding products; 7e-12 of the "
                                         "DMMA result instead of 1e-13 (tolerance 1e-8); not the default"},
        "value_three_g_shared": {"value": value_3g, "unit": "(date, g) steps/s", "ms_per_step": ms_3g, "g_per_date": 3,
                                 "g0_bit_identical_to_single_g_step": same_g0,
                                 "note": "pfml_date_steps_multi_g: the 3 bandwidths of BASELINE configs[2] for the same 148 dates in one call; "
                                         "Sigma, m_func, the gtm chains and the two inverses do not depend on g (the reference recomputes "
                                         "them per g, PFML_Input_Data.py:225,318) and are computed once; every g's outputs are bit-identical "
                                         "to the single-g step; this is what hp_grid runs"},
        "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_packed,
                "entry_point": "pfml_date_steps_host_packed",
                "outputs": "r_tilde, signal_t and the upper triangles of risk and tc (denom = risk + tc exactly; all three symmetric): "
                           "everything PFML_Input_Data.py stores per date, without the redundant two thirds of the squares",
                "chunk": args.chunk, "max_abs_diff_vs_resident_path": chk,
                "value_full_squares": e2e_full, "d2h_bytes_per_step_full_squares": d2h_full,
                "max_abs_diff_vs_resident_path_full_squares": chk_full,
                "value_r_tilde_denom_only": e2e_two,
                "host_buffers": "pinned, allocated on the GPU's NUMA node" if numa else "pinned"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "tensor",
                     "kernel": "pfml::emu_gemm_kernel (tcgen05.mma kind::i8, TMEM accumulators, bulk-TMA operands): FP64-equivalent "
                               "batched products [500 x 1025 x 500] (sums, solves), [500 x 500 x 500] (chain) and the symmetric "
                               "[1025 x 1025 x 500] a9 contractions (upper triangle)",
                     "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": (achieved / peak) if peak else None, "traffic": traffic,
                     "traffic_note": "DRAM bytes of one n x P launch (ncu --set full, profiles/ncu_summary.json)",
                     "flop_kind": "FP64-equivalent, SURVEY 8(d) algorithmic flops: (22 n^3 + 28 n^2 P + 2 n P^2) per date at n = 500, "
                                  "P = 1025 = 10.98 GFLOP; executed INT8 rate = 28 x achieved x padding",
                     "peak_source": "measured in this run: pfml_int8_mma_probe, back-to-back tcgen05.mma.kind::i8 M=128 N=256 K=32 with A "
                                    "in TMEM on all SMs (%.0f TOPS) / 28 digit products per FP64 product" % int8_peak,
                     "int8_probe": probe,
                     "frac_of_instruction_stream_ceiling": (achieved * n_pairs / mix) if mix else None,
                     "executed_frac_of_instruction_stream_ceiling": (pfl[k_emu] / args.steps / (emu_ms * 1e-3) * 1e-12 * n_pairs / mix) if mix else None,
                     "instruction_stream_ceiling_note": "the kernel issues M=128 N=64 K=32 (7 int32 accumulators x 64 columns fill TMEM); its "
                                                        "per-k-slab stream (3 tcgen05.cp + 18 products with A in TMEM + 10 from shared memory) "
                                                        "issued back to back on resident operands measures %s TOPS (pfml_int8_pipe_probe); "
                                                        "`executed_...` counts the padded flops actually issued" % (("%.0f" % mix) if mix else "n/a"),
                     "vs_cublas_dgemm_peak": achieved / peak_fp64, "cublas_dgemm_peak_tflops": peak_fp64,
                     "algorithmic_flops_per_step": emu_alg, "executed_padded_flops_per_step": pfl[k_emu] / args.steps,
                     "launch_ms_avg": emu_ms / max(emu_launches, 1.0),
                     "launches_per_step": emu_launches, "share_of_step": emu_ms / (ms_total / args.steps),
                     "dmma_gemm_launches": {"executed_tflops": pfl[0] / args.steps / (gemm_ms * 1e-3) * 1e-12,
                                            "ms_per_step": gemm_ms, "launches_per_step": pln[0] / args.steps,
                                            "share_of_step": gemm_ms / (ms_total / args.steps),
                                            "peak": peak_fp64, "peak_source": "cuBLAS DGEMM 8192^3 via torch.matmul, best of 5, this run"},
                     "algorithmic_flops_per_date": off["total"],
                     "whole_step_algorithmic_tflops": off["total"] * B / (ms_total / args.steps * 1e-3) * 1e-12},
        "hp_grid": hp,
        "c5_universe": c5,
        "hbm_rows": hb,
        "stages": stages,
        "cpu_baseline": {"value": cpu_v, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
                         "sample": f"{cpu_n} date-steps of the same N=500, p=1024 workload in {cpu_dt:.1f} s (numpy oracle, all host BLAS threads)"},
    }
    print(json.dumps(out))


def bind_to_gpu_numa(torch, dev):
    """Set this thread's CPU affinity to the cores NVML reports as local to `dev`; returns the previous affinity."""
    try:
        import pynvml
        before = os.sched_getaffinity(0)
        pr = torch.cuda.get_device_properties(dev)
        bus_id = "%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByPciBusId(bus_id.encode())
        pynvml.nvmlDeviceSetCpuAffinity(h)
        after = os.sched_getaffinity(0)
        return {"before": before, "cores": len(after)}
    except Exception:
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--dates", type=int, default=148, help="date-steps per step per GPU")
    ap.add_argument("--chunk", type=int, default=0, help="dates per device chunk in the host entry point (0 = library schedule)")
    ap.add_argument("--profile-only", action="store_true", help="resident-path steps only (for ncu launch lists)")
    ap.add_argument("--gemm-only", action="store_true", help="the dominant GEMM shape alone (for ncu --set full)")
    ap.add_argument("--no-hp-grid", action="store_true", help="skip the C3 + C4 pipeline wall-time")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world > 1:
        import torch
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))
    # Ranks != 0 are done once the collectives are: rank 0 then times the CPU arms alone.  They must not spin meanwhile (a
    # NCCL barrier or the process-group shutdown busy-waits on a host core each, and an over-subscribed OpenBLAS collapses),
    # so they sleep on a flag file until rank 0 is through, and everybody tears the process group down together.
    flag = os.path.join("/tmp", "pfml_bench_done_%s_%s" % (os.environ.get("MASTER_PORT", "0"), os.environ.get("TORCHELASTIC_RUN_ID", "x")))
    if world > 1 and rank == 0 and os.path.exists(flag):
        os.remove(flag)
    try:
        run_ours(args, rank, world, local_rank)
        if world > 1:
            if rank == 0:
                open(flag, "w").close()
            else:
                t_wait = time.time()
                while not os.path.exists(flag) and time.time() - t_wait < 900:
                    time.sleep(0.25)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
