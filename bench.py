#!/usr/bin/env python
"""bench.py -- beta coefficient-draws/sec (and ESS/sec) of the Gibbs sweep, BASELINE.json's metric.

    python bench.py --gpus N --steps K --warmup W [--workload c3] [--no-sub] [--impl reference]

The headline line is north_star's target shape, **c3** (bridge.reg.tri, n=10000, p=1000, alpha=0.5, 512 chains per GPU =
4096 on 8 GPUs, weak scaling).  A "step" is `iters` Gibbs iterations of every chain of this rank; every beta draw is
written to the device draw buffer inside the timed region, as the reference stores every draw.  The other BASELINE.json
configurations ride along in the same JSON line under "workloads" (c1; c2; c4 as 1024 chains TOTAL split over the
ranks = strong scaling; c5 = {tri, stb} x alpha in {0.125, 0.5, 0.9} on the row-sharded Gram + all-reduce path), each
with value, e2e, roofline and ESS/sec.  One JSON line on rank 0.

  value        chains * P * iterations / device time, inputs resident in HBM (CUDA events on the launching stream,
               reported by the C ABI; max over ranks)
  e2e          the same metric end to end with HOST buffers: N = 1 -> the C-ABI call bb_bridge_batched (upload of X, y +
               constructor + sweep + download of every draw); N > 1 -> the same steps through bayesbridge_b200.dist (rank 0
               uploads, ONE NCCL broadcast of [X|y] -- or per-rank row shards and ONE all-reduce of the Gram for c5 --
               then constructor, sweep, download on every rank)
  comm_ms      device time of the collective(s) of one such call (CUDA events around the NCCL calls), max over ranks
  roofline     algorithmic FP64 work of the dominant kernel / its measured duration / the measured peak of the pipe that
               bounds it (DFMA vector pipe for the triangle sweep, cuBLAS DGEMM for the Cholesky paths)
  cpu_baseline the reference's own C++ (oracle/_ref) on one host core, bounded sample, with its ESS/sec and -- for the
               triangle sampler -- the O(P^2) "fairness" sweep (the reference's as-written sweep is O(P^3))
  --impl reference   the reference's CPU sampler on all host cores (one chain per core) on the headline workload
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
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "beta coefficient-draws/sec"
UNIT = "coefficient-draws/s"

# data = key of bayesbridge_b200.synth.CONFIGS; chains = per GPU (weak) unless chains_total is given (strong scaling)
WORKLOADS = {
    "c1": dict(desc="bridge.reg.tri n=442 p=64 alpha=0.5, 1 chain (per GPU: independent replicas)", data="c1", method="triangle",
               chains=1, iters=2000, scaling="replicas"),
    "c2": dict(desc="bridge.reg.stb n=442 p=64 alpha=0.5, 512 chains/GPU", data="c2", method="stable", chains=512, iters=200,
               scaling="weak"),
    "c3": dict(desc="bridge.reg.tri n=10000 p=1000 alpha=0.5, 512 chains/GPU (4096 on 8)", data="c3", method="triangle",
               chains=512, iters=100, scaling="weak"),
    "c4": dict(desc="bridge.reg.stb n=500 p=2000 alpha=0.25 (p>n), 1024 chains in total split over the GPUs", data="c4",
               method="stable", chains_total=1024, iters=2, scaling="strong", burn=6),
    "c4w": dict(desc="c4 with the opt-in N x N Gaussian draw (same law, not the reference's arithmetic; SURVEY row f4), 1024 chains in total",
                data="c4", method="stable", chains_total=1024, iters=4, scaling="strong", burn=8, wide=True),
    "c5": dict(desc="bridge.reg.tri n=100000 p=256, row-sharded Gram + all-reduce, 256 chains/GPU", data="c5", method="triangle",
               chains=256, iters=100, scaling="weak", sharded=True),
    "c5s": dict(desc="bridge.reg.stb n=100000 p=256, row-sharded Gram + all-reduce, 256 chains/GPU", data="c5", method="stable",
                chains=256, iters=20, scaling="weak", sharded=True),
}
C5_ALPHAS = (0.125, 0.5, 0.9)


def flops_per_chain_iter(method, P):
    """DESIGN.md section 4.  Triangle: 6 P^2 FP64 operations as the kernel evaluates the sweep (1 mul, 2 FMA, 2 min, 1 axpy FMA
    per (i,j)); SURVEY.md section 8(d) counts 10 P^2 for the as-written form and both fractions are reported.
    Stable: P^3/3 + 5 P^2 flops (Cholesky + three triangular solves + the quadratic form of rss)."""
    return 6.0 * P * P if method == "triangle" else P ** 3 / 3.0 + 5.0 * P * P


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._read, daemon=True)
            self.thr.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); smax.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, idx in (("hw_slowdown", 5), ("hw_thermal_slowdown", 6), ("sw_thermal_slowdown", 7), ("sw_power_cap", 8)):
                if len(r) > idx and r[idx].lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


_DATA = {}


def make_data(key):
    from bayesbridge_b200 import synth
    if key not in _DATA:
        _DATA[key] = synth.make(key)
    return _DATA[key]


def workload_config(name, alpha=None):
    from bayesbridge_b200 import synth
    wl = WORKLOADS[name]
    c = synth.CONFIGS[wl["data"]]
    return {"workload": f"{name}: {wl['desc']}", "method": wl["method"], "N": c["N"], "P": c["P"],
            "alpha": float(alpha if alpha else c["alpha"])}


# ------------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the reference's own C++ compiled unmodified (oracle/_ref), else the port
# ------------------------------------------------------------------------------------------------
def _cpu_worker(args):
    """one chain on one core; returns (iterations, seconds of the post-burn-in loop as the reference reports it, wall, beta draws)."""
    kind, method, X, y, alpha, nsamp, seed, keep = args
    os.environ["OPENBLAS_NUM_THREADS"] = "1"
    import bb_oracle
    t0 = time.perf_counter()
    if kind == "reference":
        ref = bb_oracle.Ref()
        fn = ref.bridge_regression if method == "triangle" else ref.bridge_reg_stable
        out = fn(y, X, nsamp, alpha=alpha, burn=0, seed=seed)
    else:   # "port" (as written) or "port-o-p2" (the incremental O(P^2) sweep, same law: the fairness row)
        pm = bb_oracle.Port().model(X, y, _lapack_factors(X, y) if X.shape[0] >= X.shape[1] else None)
        fn = pm.bridge_regression if method == "triangle" else pm.bridge_reg_stable
        out = fn(nsamp, true_alpha=alpha, burn=0, seed=seed, as_written=0 if kind == "port-o-p2" else 1)
    wall = time.perf_counter() - t0
    return nsamp - 1, out["runtime"], wall, (out["beta"][1:] if keep else None)


def _lapack_factors(X, y):
    """tV, d, a, bhat as the reference's constructor forms them (BridgeRegression.cpp:41-54), by LAPACK through numpy (the
    port's own Jacobi SVD takes minutes at the c3 / c5 sizes; the timed loop does not depend on where the factors come from)"""
    U, d, Vt = np.linalg.svd(X, full_matrices=False)
    a = d * (U.T @ y)
    bhat = Vt.T @ (a / (d * d)) if d.min() > 1e-13 * d.max() else np.zeros(X.shape[1])
    return np.asfortranarray(Vt), d, a, bhat


def cpu_kind():
    import bb_oracle
    try:
        if bb_oracle.ref_available():
            bb_oracle.Ref()
            return "reference"
    except OSError:
        pass
    return "port"


# seconds per iteration of one chain on one core (measured in the authoring container; only used to size the bounded sample --
# the reported throughput always comes from the run's own `runtime`).  The .C entry points rebuild the constructor (SVD) on
# every call -- 10 s at c3, 20 s at c4 -- so the sample is sized without a probe call.
CPU_ITER_S = {("triangle", 64): 3e-4, ("stable", 64): 2.5e-4, ("triangle", 1000): 0.4, ("stable", 2000): 0.15,
              ("triangle", 256): 0.025, ("stable", 256): 0.025}
CPU_ITER_S_OP2 = {("triangle", 64): 6e-5, ("triangle", 1000): 0.02, ("triangle", 256): 0.012}


def cpu_iters_for(method, P, target_s, table=CPU_ITER_S):
    per = table.get((method, P)) or (2e-9 * P ** 3 + 1e-6 * P)
    return int(max(2, min(200000, target_s / per)))


def _ess_block(draws, runtime, P):
    """ESS/sec (coda definition, the host restatement ess.py) and posterior summaries with MCSE for a few coefficients"""
    from bayesbridge_b200 import ess as bess
    n = draws.shape[0]
    if n < 50:
        return {"n": int(n), "note": "sample too short for an ESS estimate"}
    e = np.array([bess.ess_ar(draws[:, j]) for j in range(P)])
    cols = [int(c) for c in np.unique(np.linspace(0, P - 1, min(P, 4)).astype(int))]
    post = {}
    for j in cols:
        sd = draws[:, j].std(ddof=1)
        post[f"beta[{j}]"] = {"mean": float(draws[:, j].mean()), "Q.1": float(np.quantile(draws[:, j], 0.1)),
                              "Q.9": float(np.quantile(draws[:, j], 0.9)), "mcse_mean": float(sd / np.sqrt(max(e[j], 1.0)))}
    return {"n": int(n), "ess_per_sec_min": float(e.min() / runtime), "ess_per_sec_median": float(np.median(e) / runtime),
            "estimator": "coda AR spectral (bayesbridge_b200/ess.py), one chain", "posterior": post}


def cpu_baseline(method, X, y, alpha, target_s=12.0):
    """one core, one chain: what the reference does (its own `runtime`, BridgeWrapper.cpp:165,197-202), plus ESS/sec of
    that chain and, for the triangle sampler, the fairness row: the same law with the O(P^2) incremental sweep."""
    kind = cpu_kind()
    P = X.shape[1]
    iters = cpu_iters_for(method, P, target_s)
    n, rt, wall, draws = _cpu_worker((kind, method, X, y, alpha, iters + 1, 7, True))
    out = {"value": n * P / rt, "unit": UNIT, "cores": 1, "kind": kind,
           "sample": f"1 chain, {n} post-burn-in iterations of the reference driver ({rt:.2f} s CPU, its own `runtime`; constructor excluded)",
           "ess": _ess_block(draws, rt, P)}
    if method == "triangle":
        it2 = cpu_iters_for(method, P, 0.5 * target_s, CPU_ITER_S_OP2)
        n2, rt2, _, d2 = _cpu_worker(("port-o-p2", method, X, y, alpha, it2 + 1, 7, True))
        out["fairness_o_p2"] = {"value": n2 * P / rt2, "unit": UNIT, "cores": 1, "kind": "port, O(P^2) incremental sweep (same law)",
                                "sample": f"1 chain, {n2} iterations ({rt2:.2f} s CPU)", "ess": _ess_block(d2, rt2, P),
                                "note": "the reference's sweep as written is O(P^3) per iteration (BridgeRegression.cpp:250-265); "
                                        "this row removes that algorithmic factor from the GPU/CPU ratio"}
    return out


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    wl = WORKLOADS[args.workload]
    X, y, _ = make_data(wl["data"])
    cfg = workload_config(args.workload, args.alpha)
    method, alpha, P = cfg["method"], cfg["alpha"], X.shape[1]
    kind = cpu_kind()
    cores = os.cpu_count() or 1
    # Bounded sample: about 60 s of timed CPU work per core.  Every call of the reference's .C entry point rebuilds its
    # constructor (two SVD-sized LAPACK calls: 10-20 s at c3, outside its own `runtime`), so the W warm-up steps are ONE call
    # of 2 W iterations per core and the K timed steps are ONE call of K x iters iterations per core; a step's time is the
    # slowest core's `runtime` / K (the reference's clock around its sampling loop, BridgeWrapper.cpp:165,197-202).
    per_step_s = max(1.5, min(20.0, 60.0 / max(1, args.steps)))
    iters = cpu_iters_for(method, P, per_step_s)
    ctx = mp.get_context("fork")
    with ctx.Pool(cores) as pool:
        if args.warmup > 0:
            pool.map(_cpu_worker, [(kind, method, X, y, alpha, 2 * args.warmup + 1, 500 + c, False) for c in range(cores)])
        res = pool.map(_cpu_worker, [(kind, method, X, y, alpha, iters * args.steps + 1, 1000 + c, False) for c in range(cores)])
    times = [max(r[1] for r in res)]
    T = float(np.sum(times))
    value = cores * iters * P * args.steps / T
    cfg["cpu_sample"] = f"{cores} independent chains (one per host core) x {iters} iterations per step, the {args.steps} steps in one driver call per core"
    line = {"metric": METRIC, "value": value, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * T / args.steps, "higher_is_better": True, "scaling": wl["scaling"],
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                             "sample": f"{cores} chains x {iters} post-burn-in iterations per step, reference driver's own runtime"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def measure_peaks(torch, L, dev):
    """FP64 roofline denominators, measured in this run (MEASURED_PEAKS.json carries no FP64 figure): cuBLAS DGEMM 8192^3
    (best of 5, burst) for the tensor (DMMA) rows and a DFMA loop (bb_measure_fp64_fma) for the vector-pipe row."""
    import ctypes as C
    n = 8192
    a = torch.randn(n, n, dtype=torch.float64, device=dev)
    b = torch.randn(n, n, dtype=torch.float64, device=dev)
    torch.matmul(a, b)
    torch.cuda.synchronize(dev)
    best = 1e30
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); torch.matmul(a, b); e1.record(); torch.cuda.synchronize(dev)
        best = min(best, e0.elapsed_time(e1))
    del a, b
    torch.cuda.empty_cache()
    fma = C.c_double(0.0)
    if L.bb_measure_fp64_fma(dev.index, C.byref(fma)):
        fma.value = float("nan")
    return {"dgemm_tflops": 2.0 * n ** 3 / (best * 1e-3) / 1e12, "dfma_tflops": fma.value}


def measured_hbm():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return json.load(open(p))["hbm_gbs"], "MEASURED_PEAKS.json"
    except (OSError, KeyError, ValueError):
        return 6650.0, "B200_PROFILING.md fallback"


def kernel_name(method, P):
    if method == "triangle":
        return "tri_sweep_kernel"
    if P <= 64:
        return "stb64_kernel"
    return "stb_small_kernel" if P <= 128 else "stbl_colblock_kernel (+ potrf / solve / scalar / lambda)"


def roofline_of(name, method, P, nch, ips, steps, T_local, peaks, N=None):
    per_chain = flops_per_chain_iter(method, P) if N is None else (float(N) * N * P + N ** 3 / 3.0 + 4.0 * N * P + 2.0 * N * N)
    flops_step = per_chain * nch * ips
    achieved = flops_step * steps / T_local / 1e12
    traffic, ncu_info = None, None
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(tpath):
        ncu_info = json.load(open(tpath)).get(name) or None
        traffic = (ncu_info or {}).get("bytes_per_launch")
    hbm, hbm_src = measured_hbm()
    out = {"kernel": kernel_name(method, P), "achieved": achieved, "unit": "TFLOP/s", "traffic": traffic,
           "algorithmic_flops_per_chain_iter": per_chain,
           "hbm_store_GBs": 8.0 * P * nch * ips * steps / T_local / 1e9, "hbm_peak_GBs": hbm, "hbm_peak_source": hbm_src, "ncu": ncu_info}
    if method == "triangle":
        out.update(bound="fp64-vector", peak=peaks["dfma_tflops"], frac=achieved / peaks["dfma_tflops"],
                   peak_source="DFMA loop measured in this run (bb_measure_fp64_fma); the sweep has no contraction, so no tensor pipe",
                   frac_survey_10P2=achieved * (10.0 / 6.0) / peaks["dfma_tflops"],
                   note="6 P^2 FP64 operations per chain-iteration as evaluated (SURVEY's as-written count is 10 P^2: frac_survey_10P2); "
                        "latency-bound: P sequential coordinate steps per iteration, see profiles/ for fp64-pipe utilisation and stalls")
    elif P <= 128:
        ls = (ncu_info or {}).get("lambda_stage")
        if ls:   # trials/s and mean trials per draw of the tilted-stable stage (counters of a dev build, profiles/)
            draws_per_s = P * nch * ips * steps / T_local
            out["lambda_stage"] = dict(ls, sequential_rounds_per_s=ls["sequential_rounds_per_draw"] * draws_per_s,
                                       rounds_evaluated_per_s=ls["rounds_evaluated_per_draw"] * draws_per_s)
        out.update(bound="fp64-latency", peak=peaks["dgemm_tflops"], frac=achieved / peaks["dgemm_tflops"],
                   peak_source="cuBLAS DGEMM 8192^3 measured in this run",
                   note="P^3/3 + 5 P^2 flops are tiny at this P; the iteration is dependent FP64 chains (half of it the tilted-stable "
                        "rejection stage, which has no contraction): see ncu.fp64_pipe_pct / issue_pct and lambda_stage")
    else:
        out.update(bound="tensor", peak=peaks["dgemm_tflops"], frac=achieved / peaks["dgemm_tflops"],
                   peak_source="cuBLAS DGEMM 8192^3 measured in this run (DMMA pipe)")
    return out


def measure_sweep(torch, bb, bdist, L, dev, name, method, alpha, model, P, nch, chain0, ips, steps, warmup, burn, peaks,
                  want_ess=True, sample_clocks=False, world=1, wide=False):
    """device-timed sweep of `nch` chains of this rank: W warm-up + K timed steps of `ips` iterations, L2 flushed between steps."""
    from bayesbridge_b200._lib import ST_ALL
    hyp = bb.make_hyper(alpha=alpha)
    chains = model.chains(method, nch, seed=20121027, chain0=chain0)
    if wide:
        chains.wide_gaussian()
    chains.reserve(steps * ips, ("beta",))
    chains.run(hyp, burn, 0, ST_ALL, pre_tau=1)        # burn-in, untimed: the timed iterations run in the stationary regime
    it0 = burn
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2
    clocks = ClockSampler(dev.index) if sample_clocks else None
    step_ms, launches0 = [], 0
    for s in range(warmup + steps):
        timed = s >= warmup
        flush.fill_(s & 0xFF)
        torch.cuda.synchronize(dev)
        if s == warmup:
            bdist.barrier(); torch.cuda.synchronize(dev)
            if clocks:
                clocks.start()
            launches0 = L.bb_launch_count()
        ms = chains.run(hyp, ips, it0, ST_ALL, thin=1 if timed else 0, slot0=(s - warmup) * ips if timed else 0)
        it0 += ips
        if timed:
            step_ms.append(ms)
    torch.cuda.synchronize(dev); bdist.barrier()
    launches = L.bb_launch_count() - launches0
    clk = clocks.stop() if clocks else None
    T_local = float(np.sum(step_ms)) * 1e-3
    T = bdist.max_over_ranks(T_local, dev)
    chains_total = int(bdist.sum_over_ranks(np.array([float(nch)]), dev)[0])
    res = {"value": chains_total * P * ips * steps / T, "unit": UNIT, "ms_per_step": 1e3 * T / steps, "chains_total": chains_total,
           "chains_this_rank": nch, "iters_per_step": ips, "burn_in_untimed": burn, "gpu_launches": int(launches),
           "chain_status_nonzero": int((chains.get()["status"] != 0).sum()),
           "roofline": roofline_of(name, method, P, nch, ips, steps, T_local, peaks, N=model.N if wide else None)}
    if clk:
        res["clocks"] = clk
    if want_ess:
        nsl = min(steps * ips, 16384)
        t_s = time.perf_counter()
        sm = chains.summary("beta", slot0=0, nslots=nsl)
        summary_s = time.perf_counter() - t_s
        e = bdist.sum_over_ranks(sm["ESS"], dev) * (steps * ips / nsl)
        res["ess_per_sec"] = {"min": float(e.min() / T), "median": float(np.median(e) / T), "unit": "ESS/s summed over all chains",
                              "estimator": "coda AR spectral (bb_chains_summary, on the device), all chains x all coefficients",
                              "n_per_chain": int(nsl), "summary_s": summary_s,
                              "posterior_beta0": {"mean": float(sm["mean"][0]), "Q.1": float(sm["Q.1"][0]), "Q.9": float(sm["Q.9"][0])}}
    chains.close()
    return res


def e2e_capi(torch, bb, L, dev, method, Xp, yp, N, P, alpha, nch, chain0, ips, steps):
    """K calls of bb_bridge_batched: pinned host X, y in, every beta draw (+ sig2, tau) out to pinned host memory."""
    import ctypes as C
    from bayesbridge_b200._lib import BB_STB, BB_TRI, dp, up
    nsamp = ips + 1
    beta = torch.empty((nch, nsamp, P), dtype=torch.float64).pin_memory()
    sig2 = torch.empty((nch, nsamp), dtype=torch.float64).pin_memory()
    tau = torch.empty((nch, nsamp), dtype=torch.float64).pin_memory()
    status = np.zeros(nch, dtype=np.uint32)
    hyp = bb.make_hyper(alpha=alpha)
    rt = C.c_double(0)
    cast = lambda t: C.cast(t.data_ptr(), dp)
    walls = []
    for s in range(1 + steps):
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        rc = L.bb_bridge_batched(BB_TRI if method == "triangle" else BB_STB, 0, cast(Xp), cast(yp), N, P, C.byref(hyp), nch,
                                 20121027 + s, chain0, 0, nsamp, 1, dev.index, cast(beta), None, None, None, cast(sig2), cast(tau),
                                 None, status.ctypes.data_as(up), C.byref(rt))
        torch.cuda.synchronize(dev)
        wall = time.perf_counter() - t0
        if rc:
            raise RuntimeError("bb_bridge_batched: " + bb._lib.last_error())
        if s > 0:
            walls.append(wall)
    return float(np.sum(walls)), {"h2d_bytes_per_step": int(8 * (N * P + N)), "d2h_bytes_per_step": int(8 * (nch * nsamp * (P + 2)) + 4 * nch),
                                  "call": "bb_bridge_batched(host X,y -> host draws): upload + Gram/SVD constructor + sweep + download",
                                  "posterior_mean_beta0": float(beta[:, 1:, 0].mean()), "comm_ms": 0.0}


def e2e_dist(torch, bb, bdist, dev, method, host_pack, N, P, alpha, nch, chain0, ips, steps, sharded, rank, wide=False):
    """The same call at N > 1 through bayesbridge_b200.dist: host -> device on the source rank(s), ONE collective, then
    constructor + sweep + download on every rank.  host_pack: pinned packed [X | y] (rank 0) or this rank's row shard."""
    from bayesbridge_b200._lib import ST_ALL
    hyp = bb.make_hyper(alpha=alpha)
    nsamp = ips + 1
    walls, comm, h2d = [], [], 0
    mean0 = 0.0
    hb = torch.empty((nch, nsamp, P), dtype=torch.float64).pin_memory()      # pinned host buffers for the draws, as in e2e_capi
    hs = torch.empty((nch, nsamp), dtype=torch.float64).pin_memory()
    ht = torch.empty((nch, nsamp), dtype=torch.float64).pin_memory()
    for s in range(1 + steps):
        bdist.barrier(); torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        tm = {}
        if sharded:
            model = bdist.sharded_model(host_pack, None, N, P, dev.index, timing=tm)
            h2d = 8 * host_pack.numel()
        else:
            model = bdist.replicated_model(host_pack, None, N, P, dev.index, timing=tm)
            h2d = 8 * host_pack.numel() if rank == 0 else 0
        ch = model.chains(method, nch, seed=20121027 + s, chain0=chain0)
        if wide:
            ch.wide_gaussian()
        ch.reserve(nsamp, ("beta", "sig2", "tau"))
        nburn = 0 if method == "triangle" else 1      # run_driver's schedule at burn = 0 (BridgeWrapper.cpp:142-145, 262-266)
        ch.run(hyp, nburn, 0, ST_ALL, pre_tau=1)
        ch.store_state(0)
        ch.run(hyp, ips, nburn, ST_ALL, thin=1, slot0=1)
        ch.fetch("beta", out=hb); ch.fetch("sig2", out=hs); ch.fetch("tau", out=ht)
        torch.cuda.synchronize(dev); bdist.barrier()
        wall = time.perf_counter() - t0
        mean0 = float(hb[:, 1:, 0].mean())
        ch.close(); model.close()
        if s > 0:
            walls.append(wall); comm.append(tm.get("comm_ms", 0.0))
        coll = tm.get("collective")
    return float(np.sum(walls)), {"h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(8 * nch * nsamp * (P + 2)),
                                  "call": "dist.%s_model(host rows) -> Model.chains -> run -> fetch (host draws); " % ("sharded" if sharded else "replicated")
                                          + (coll or "no collective"),
                                  "posterior_mean_beta0": mean0, "comm_ms": float(np.mean(comm)) if comm else 0.0}


def run_workload(ctx, name, alpha=None, model=None, steps=None, warmup=None, main=False, keep_model=False, cpu=True):
    """one BASELINE configuration on this job: device-timed sweep + end-to-end leg (+ CPU baseline at N = 1)."""
    torch, bb, bdist, L, dev, rank, world, args, peaks = (ctx[k] for k in ("torch", "bb", "bdist", "L", "dev", "rank", "world", "args", "peaks"))
    wl = WORKLOADS[name]
    cfg = workload_config(name, alpha)
    method, alpha, N, P = cfg["method"], cfg["alpha"], cfg["N"], cfg["P"]
    steps = steps or args.steps
    warmup = max(3, warmup if warmup is not None else args.warmup)
    ips = (args.iters if (main and args.iters) else 0) or wl["iters"]
    if "chains_total" in wl:                             # strong scaling: a fixed job split over the ranks
        lo, hi = bdist.chain_range(wl["chains_total"], rank, world)
        nch, chain0 = hi - lo, lo
    else:
        nch = (args.chains if (main and args.chains) else 0) or wl["chains"]
        chain0 = rank * nch
    sharded = bool(wl.get("sharded")) and world > 1
    X, y, _ = make_data(wl["data"])
    own_model = model is None
    timing = {}
    t0 = time.perf_counter()
    if own_model:
        if sharded:
            lo, hi = bdist.row_range(N, rank, world)
            model = bdist.sharded_model(X[lo:hi], y[lo:hi], N, P, dev.index, timing=timing)
        else:
            model = bdist.replicated_model(X if rank == 0 else None, y if rank == 0 else None, N, P, dev.index, timing=timing)
    precompute_s = time.perf_counter() - t0
    burn = args.burn if (main and args.burn is not None) else wl.get("burn", max(ips, 50 if P >= 1000 else 200))
    res = measure_sweep(torch, bb, bdist, L, dev, name, method, alpha, model, P, nch, chain0, ips, steps, warmup, burn, peaks,
                        want_ess=args.ess, sample_clocks=main, world=world, wide=bool(wl.get("wide")))
    res["scaling"] = wl["scaling"]
    res["config"] = dict(cfg, chains_per_gpu=nch if "chains_total" not in wl else None, chains_total=res["chains_total"],
                         iters_per_step=ips, multi_gpu=("row-sharded Gram + all-reduce" if sharded else
                                                        ("X replicated by 1 NCCL broadcast" if world > 1 else "single GPU")))
    if own_model:
        res["precompute_s"] = precompute_s
        res["constructor_comm_ms"] = bdist.max_over_ranks(timing.get("comm_ms", 0.0), dev)

    # ---- end to end with host buffers ----
    e2e_steps = max(1, min(steps, args.e2e_steps))
    L.bb_set_wide_gaussian(1 if wl.get("wide") else 0)
    if world == 1:
        Xp = torch.from_numpy(np.asfortranarray(X).ravel(order="F").copy()).pin_memory()
        yp = torch.from_numpy(np.ascontiguousarray(y).copy()).pin_memory()
        T_e, info = e2e_capi(torch, bb, L, dev, method, Xp, yp, N, P, alpha, nch, chain0, ips, e2e_steps)
    else:
        if sharded:
            lo, hi = bdist.row_range(N, rank, world)
            pack = torch.from_numpy(np.concatenate([np.asfortranarray(X[lo:hi]).ravel(order="F"), y[lo:hi]])).pin_memory()
        else:
            pack = torch.from_numpy(np.concatenate([np.asfortranarray(X).ravel(order="F"), y])).pin_memory() if rank == 0 else None
        T_e, info = e2e_dist(torch, bb, bdist, dev, method, pack, N, P, alpha, nch, chain0, ips, e2e_steps, sharded, rank, wide=bool(wl.get("wide")))
    L.bb_set_wide_gaussian(0)
    T_e = bdist.max_over_ranks(T_e, dev)
    info["comm_ms"] = bdist.max_over_ranks(info["comm_ms"], dev)
    res["e2e"] = dict(value=res["chains_total"] * P * ips * e2e_steps / T_e, unit=UNIT, steps=e2e_steps, **info)
    res["comm_ms"] = info["comm_ms"]
    if own_model and not keep_model:
        model.close()
        model = None
    if world == 1 and rank == 0 and not args.no_cpu and cpu:
        res["cpu_baseline"] = cpu_baseline(method, X, y, alpha, args.cpu_seconds if main else args.cpu_seconds_sub)
        res["gpu_over_cpu_1core"] = res["value"] / res["cpu_baseline"]["value"]
    return res, model


def run_ours(args):
    import torch
    import bayesbridge_b200 as bb
    from bayesbridge_b200 import dist as bdist

    rank, world, local = bdist.init()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the sampler has no CPU fallback (use --impl reference for the CPU arm)")
    dev = torch.device("cuda", local)
    L = bb._lib.load()
    if os.environ.get("BB_VERBOSE"):
        L.bb_set_verbose(int(os.environ["BB_VERBOSE"]))   # 2: per-phase timings of the batched driver on stdout
    peaks = measure_peaks(torch, L, dev)
    ctx = dict(torch=torch, bb=bb, bdist=bdist, L=L, dev=dev, rank=rank, world=world, args=args, peaks=peaks)

    main, _ = run_workload(ctx, args.workload, alpha=args.alpha or None, main=True)
    wl = WORKLOADS[args.workload]
    line = {"metric": METRIC, "value": main["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": main["ms_per_step"], "higher_is_better": True, "scaling": wl["scaling"], "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": dict(main["config"], burn_in_untimed=main["burn_in_untimed"],
                           stored="every beta draw written to HBM inside the timed region",
                           l2="256 MiB buffer written between steps (L2 flush); chain state lives on chip"),
            "e2e": main["e2e"], "comm_ms": main["comm_ms"], "gpu_launches": main["gpu_launches"], "clocks": main.get("clocks"),
            "roofline": main["roofline"], "precompute_s": main.get("precompute_s"), "constructor_comm_ms": main.get("constructor_comm_ms"),
            "chain_status_nonzero": main["chain_status_nonzero"], "peaks_measured": peaks}
    for k in ("ess_per_sec", "cpu_baseline", "gpu_over_cpu_1core"):
        if k in main:
            line[k] = main[k]

    # ---- the other BASELINE configurations, same contract, fewer steps ----
    if args.sub:
        subs = {}
        t_sub = time.perf_counter()
        for name in [s for s in args.sub.split(",") if s and s != args.workload]:
            try:
                if name == "c5":
                    shared = None
                    for meth_name in ("c5", "c5s"):
                        for a in C5_ALPHAS:
                            r, shared = run_workload(ctx, meth_name, alpha=a, model=shared, steps=args.sub_steps, warmup=3,
                                                     keep_model=True, cpu=(a == 0.5))
                            subs[f"{meth_name} alpha={a}"] = r
                    if shared is not None:
                        shared.close()
                else:
                    r, _ = run_workload(ctx, name, steps=args.sub_steps, warmup=3)
                    subs[name] = r
            except Exception as e:   # a sub-result must never take the headline down with it
                subs[name] = {"error": f"{type(e).__name__}: {e}"}
        line["workloads"] = subs
        line["workloads_wall_s"] = time.perf_counter() - t_sub
    if rank == 0:
        print(json.dumps(line), flush=True)
    bdist.barrier()
    bdist.shutdown()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--sub", default=None, help="comma list of the other configurations carried under `workloads` "
                                                "(default: c1,c2,c4,c4w,c5 when --workload is the default, none otherwise)")
    ap.add_argument("--no-sub", action="store_true")
    ap.add_argument("--sub-steps", type=int, default=3)
    ap.add_argument("--chains", type=int, default=0, help="chains per GPU of the headline workload (default: the workload's)")
    ap.add_argument("--iters", type=int, default=0, help="Gibbs iterations per step of the headline workload")
    ap.add_argument("--burn", type=int, default=None)
    ap.add_argument("--alpha", type=float, default=0.0, help="override the headline workload's alpha")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--cpu-seconds-sub", type=float, default=4.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-ess", dest="ess", action="store_false")
    args = ap.parse_args()
    if args.sub is None:
        args.sub = "c1,c2,c4,c4w,c5" if ("--workload" not in " ".join(sys.argv)) else ""
    if args.no_sub:
        args.sub = ""
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
