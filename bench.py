#!/usr/bin/env python
"""bench.py -- beta coefficient-draws/sec (and ESS/sec) of the Gibbs sweep, BASELINE.json's metric.

    python bench.py --gpus N --steps K --warmup W [--workload c2] [--impl reference]

A "step" is `--iters` Gibbs iterations of every chain on this rank (default workload c2: the
stable-mixture sampler, n=442, p=64, 512 chains per GPU -- weak scaling: rank r owns global
chains [512 r, 512 (r+1))).  Every draw of beta is written to the device draw buffer inside the
timed region, as the reference stores every draw.  One JSON line on rank 0.

  value      : chains * P * iterations / device time, inputs resident in HBM (CUDA events on the
               launching stream, reported by the C ABI; max over ranks)
  e2e        : the same metric through the public call `bb_bridge_batched` with HOST buffers
               (upload of X,y + precompute + sweep + download of every draw inside the timed region)
  roofline   : algorithmic FP64 work of the dominant kernel / its measured duration / measured FP64 peak
  cpu_baseline: the reference's own C++ (oracle/_ref) on one host core, bounded sample
  --impl reference : the reference's CPU sampler on all host cores (one chain per core)
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

# per-GPU chain counts (weak scaling) and iterations per step for each BASELINE.json configuration
WORKLOADS = {
    "c1": dict(desc="bridge.reg.tri n=442 p=64 alpha=0.5, 1 chain", chains=1, iters=2000),
    "c2": dict(desc="bridge.reg.stb n=442 p=64 alpha=0.5, 512 chains/GPU", chains=512, iters=200),
    "c3": dict(desc="bridge.reg.tri n=10000 p=1000 alpha=0.5, 512 chains/GPU (4096 on 8)", chains=512, iters=40),
    "c4": dict(desc="bridge.reg.stb n=500 p=2000 alpha=0.25 (p>n), 128 chains/GPU (1024 on 8)", chains=128, iters=4),
    "c5": dict(desc="bridge.reg.tri n=100000 p=256 alpha=0.5 row-sharded Gram, 256 chains/GPU", chains=256, iters=100),
    "c5s": dict(desc="bridge.reg.stb n=100000 p=256 alpha=0.5 row-sharded Gram, 256 chains/GPU", chains=256, iters=20),
}


def algorithmic_flops_per_chain_iter(method, P):
    """triangle 6 P^2 FP64 operations (1 mul, 2 FMA, 2 min, 1 axpy FMA per (i,j); SURVEY.md
    section 8(d) counted 10 for the as-written form); stable P^3/3 + 5 P^2 flops (Cholesky + solves + rss form)."""
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


def make_data(name):
    from bayesbridge_b200 import synth
    wl = "c5" if name == "c5s" else name
    cfg = dict(synth.CONFIGS[wl])
    if name == "c5s":
        cfg["method"] = "stable"
    return cfg, synth.make(wl)


# ------------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the reference's own C++ compiled unmodified (oracle/_ref), else the port
# ------------------------------------------------------------------------------------------------
def _cpu_worker(args):
    """one chain on one core; returns (iterations, seconds of the post-burn-in loop as the reference reports it)."""
    kind, method, X, y, alpha, nsamp, seed = args
    os.environ["OPENBLAS_NUM_THREADS"] = "1"
    import bb_oracle
    if kind == "reference":
        ref = bb_oracle.Ref()
        fn = ref.bridge_regression if method == "triangle" else ref.bridge_reg_stable
        t0 = time.perf_counter()
        out = fn(y, X, nsamp, alpha=alpha, burn=0, seed=seed)
        wall = time.perf_counter() - t0
        return nsamp - 1, out["runtime"], wall
    pm = bb_oracle.Port().model(X, y)
    fn = pm.bridge_regression if method == "triangle" else pm.bridge_reg_stable
    t0 = time.perf_counter()
    out = fn(nsamp, true_alpha=alpha, burn=0, seed=seed, as_written=1)
    wall = time.perf_counter() - t0
    return nsamp - 1, out["runtime"], wall


def cpu_kind():
    import bb_oracle
    try:
        if bb_oracle.ref_available():
            bb_oracle.Ref()
            return "reference"
    except OSError:
        pass
    return "port"


def cpu_iters_for(method, X, y, alpha, kind, target_s):
    """choose a sample size (iterations) worth about target_s seconds on one core"""
    probe = 3 if X.shape[1] >= 1000 else 20
    n, rt, wall = _cpu_worker((kind, method, X, y, alpha, probe + 1, 1))
    per = max(rt, 1e-6) / n
    return int(max(2, min(200000, target_s / per))), per


def cpu_baseline(method, X, y, alpha, target_s=12.0):
    """one core, one chain: what the reference does (BridgeWrapper.cpp:165,197-202 timing)."""
    kind = cpu_kind()
    iters, _ = cpu_iters_for(method, X, y, alpha, kind, target_s)
    n, rt, wall = _cpu_worker((kind, method, X, y, alpha, iters + 1, 7))
    P = X.shape[1]
    return {"value": n * P / rt, "unit": UNIT, "cores": 1, "kind": kind,
            "sample": f"1 chain, {n} post-burn-in iterations of the reference driver ({rt:.2f} s CPU, its own `runtime`)"}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    cfg, (X, y, _) = make_data(args.workload)
    method, alpha, P = cfg["method"], (args.alpha if args.alpha > 0 else cfg["alpha"]), X.shape[1]
    kind = cpu_kind()
    cores = os.cpu_count() or 1
    per_step_s = max(2.0, min(20.0, 150.0 / max(1, args.steps + args.warmup)))
    iters, _ = cpu_iters_for(method, X, y, alpha, kind, per_step_s)
    ctx = mp.get_context("fork")
    times = []
    with ctx.Pool(cores) as pool:
        for s in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            res = pool.map(_cpu_worker, [(kind, method, X, y, alpha, iters + 1, 1000 * s + c) for c in range(cores)])
            wall = time.perf_counter() - t0
            if s >= args.warmup:
                times.append(max(r[1] for r in res))   # the reference's own runtime, slowest core
    T = float(np.sum(times))
    value = cores * iters * P * args.steps / T
    line = {"metric": METRIC, "value": value, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * T / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {WORKLOADS[args.workload]['desc']}",
                       "cpu_sample": f"{cores} independent chains (one per host core) x {iters} iterations per step"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                             "sample": f"{cores} chains x {iters} post-burn-in iterations per step, reference driver's own runtime"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def measure_fp64_peak(torch, dev):
    """cuBLAS DGEMM 8192^3, best of 5 (burst): the FP64 roofline denominator (MEASURED_PEAKS.json has none)."""
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
    return 2.0 * n ** 3 / (best * 1e-3) / 1e12


def run_ours(args):
    import torch
    import bayesbridge_b200 as bb
    from bayesbridge_b200 import dist as bdist, ess as bess
    from bayesbridge_b200._lib import ST_ALL

    rank, world, local = bdist.init()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the sampler has no CPU fallback (use --impl reference for the CPU arm)")
    dev = torch.device("cuda", local)
    L = bb._lib.load()
    if os.environ.get("BB_VERBOSE"):
        L.bb_set_verbose(int(os.environ["BB_VERBOSE"]))   # 2: per-phase timings of the batched driver on stdout
    wl = WORKLOADS[args.workload]
    cfg, (X, y, _) = make_data(args.workload)
    method, alpha = cfg["method"], (args.alpha if args.alpha > 0 else cfg["alpha"])
    N, P = X.shape
    nch = args.chains or wl["chains"]
    ips = args.iters or wl["iters"]
    chain0 = rank * nch
    seed = 20121027

    # ---- one-time: replicate X (one NCCL broadcast) or row-shard the Gram build (one all-reduce) ----
    t0 = time.perf_counter()
    if args.workload in ("c5", "c5s") and world > 1:
        lo, hi = bdist.row_range(N, rank, world)
        model = bdist.sharded_model(X[lo:hi], y[lo:hi], N, P, local)
        gram_mode = "row-sharded Gram + 1 NCCL all-reduce"
    else:
        model = bdist.replicated_model(X if rank == 0 else None, y if rank == 0 else None, N, P, local)
        gram_mode = "X replicated by 1 NCCL broadcast" if world > 1 else "single GPU"
    precompute_s = time.perf_counter() - t0
    hyp = bb.make_hyper(alpha=alpha)
    chains = model.chains(method, nch, seed=seed, chain0=chain0)
    total_steps = args.warmup + args.steps
    out_name = "beta"
    chains.reserve(args.steps * ips, (out_name,))
    # burn-in so that the timed iterations run in the stationary regime (not timed)
    burn = args.burn if args.burn is not None else max(ips, 50 if P >= 1000 else 200)
    chains.run(hyp, burn, 0, ST_ALL, pre_tau=1)
    it0 = burn
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    fp64_peak = measure_fp64_peak(torch, dev)
    clocks = ClockSampler(local)
    step_ms = []
    launches0 = None
    for s in range(total_steps):
        timed = s >= args.warmup
        flush.fill_(s & 0xFF)
        torch.cuda.synchronize(dev)
        if s == args.warmup:
            bdist.barrier(); torch.cuda.synchronize(dev)
            clocks.start()
            launches0 = L.bb_launch_count()
        slot0 = (s - args.warmup) * ips if timed else 0
        ms = chains.run(hyp, ips, it0, ST_ALL, thin=1 if timed else 0, slot0=slot0)   # device time, CUDA events on the launching stream
        it0 += ips
        if timed:
            step_ms.append(ms)
    torch.cuda.synchronize(dev); bdist.barrier()
    launches = L.bb_launch_count() - launches0
    clk = clocks.stop()
    T_local = float(np.sum(step_ms)) * 1e-3
    T = bdist.max_over_ranks(T_local, dev)
    total_chains = nch * world
    value = total_chains * P * ips * args.steps / T
    status = chains.get()["status"]

    # ---- roofline of the dominant kernel (the sweep kernel; one launch per step except the large-P stable path) ----
    flops_step = algorithmic_flops_per_chain_iter(method, P) * nch * ips
    achieved = flops_step * args.steps / T_local / 1e12
    traffic, ncu_info = None, None
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(tpath):
        ncu_info = json.load(open(tpath)).get(args.workload) or None
        traffic = (ncu_info or {}).get("bytes_per_launch")
    store_bytes = 8.0 * P * nch * ips
    roofline = {"bound": "tensor", "achieved": achieved, "peak": fp64_peak, "unit": "TFLOP/s", "frac": achieved / fp64_peak,
                "traffic": traffic, "kernel": kernel_name(method, P),
                "peak_source": "cuBLAS DGEMM 8192^3 measured in this run (FP64; MEASURED_PEAKS.json carries no FP64 figure)",
                "algorithmic_flops_per_chain_iter": algorithmic_flops_per_chain_iter(method, P),
                "hbm_store_GBs": store_bytes * args.steps / T_local / 1e9, "hbm_peak_GBs": measured_hbm(),
                "ncu": ncu_info,   # committed capture of this kernel (profiles/): FP64-pipe and issue-slot utilisation, DRAM bytes per launch
                "note": "FP64 path: no bf16/HBM roofline applies; the sweep kernels are issue/latency-bound (sequential Gibbs steps, "
                        "3.5 chains per SM), see profiles/*.md for fp64-pipe utilisation and stall breakdowns"}

    # ---- ESS/sec: sum.stat on the device over EVERY chain and coefficient of the timed region (bb_chains_summary:
    # coda's effectiveSize per series, summed over chains and ranks); the host estimator cross-checks one chain ----
    ess_info = None
    if args.ess:
        nsl = min(args.steps * ips, 16384)
        t_s = time.perf_counter()
        sm = chains.summary(out_name, slot0=0, nslots=nsl)
        summary_s = time.perf_counter() - t_s
        e = bdist.sum_over_ranks(sm["ESS"], dev) * (args.steps * ips / nsl)
        if rank == 0:
            cols = np.unique(np.linspace(0, P - 1, min(P, 8)).astype(int))
            d0 = chains.fetch(out_name, 0, nsl)[0] if nch * nsl * P <= (1 << 27) else None
            check = None
            if d0 is not None:
                pc = chains.summary(out_name, slot0=0, nslots=nsl, per_chain=True)["chains"]
                check = float(max(abs(pc[0, 2, j] / bess.ess_ar(d0[:, j]) - 1.0) for j in cols))
            ess_info = {"min": float(e.min() / T), "median": float(np.median(e) / T), "unit": "ESS/s summed over all chains",
                        "estimator": "coda AR spectral (bb_chains_summary, on the device), all chains x all coefficients",
                        "n_per_chain": int(nsl), "summary_s": summary_s, "host_check_max_rel_diff": check,
                        "posterior_mean_beta0": float(sm["mean"][0]), "q10_q90_beta0": [float(sm["Q.1"][0]), float(sm["Q.9"][0])]}

    # ---- end to end through the public C ABI call with host buffers ----
    e2e = run_e2e(L, torch, bb, method, X, y, alpha, nch, chain0, ips, args, dev, world)

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": 1e3 * T / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": {"workload": f"{args.workload}: {wl['desc']}", "method": method, "N": N, "P": P, "alpha": alpha,
                           "chains_per_gpu": nch, "chains_total": total_chains, "iters_per_step": ips, "burn_in_untimed": burn,
                           "stored": "every beta draw written to HBM inside the timed region", "multi_gpu": gram_mode,
                           "l2": "256 MiB buffer written between steps (L2 flush); chain state lives on chip"},
                "e2e": e2e, "gpu_launches": int(launches), "clocks": clk, "roofline": roofline,
                "precompute_s": precompute_s, "chain_status_nonzero": int((status != 0).sum())}
        if ess_info:
            line["ess_per_sec"] = ess_info
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline(method, X, y, alpha, args.cpu_seconds)
        print(json.dumps(line), flush=True)
    bdist.barrier()
    bdist.shutdown()


def kernel_name(method, P):
    if method == "triangle":
        return "tri_sweep_kernel"
    if P <= 64:
        return "stb64_kernel"
    return "stb_small_kernel" if P <= 128 else "stbl_update_kernel (+potrf/trsm/solve)"


def measured_hbm():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return json.load(open(p))["hbm_gbs"]
    except (OSError, KeyError, ValueError):
        return 6650.0   # B200_PROFILING.md fallback


def run_e2e(L, torch, bb, method, X, y, alpha, nch, chain0, ips, args, dev, world):
    """K calls of bb_bridge_batched: host X,y in (pinned), every beta draw out to pinned host memory."""
    import ctypes as C
    from bayesbridge_b200 import dist as bdist
    from bayesbridge_b200._lib import BB_STB, BB_TRI, dp, up
    N, P = X.shape
    nsamp = ips + 1
    Xp = torch.from_numpy(np.asfortranarray(X).ravel(order="F").copy()).pin_memory()
    yp = torch.from_numpy(np.ascontiguousarray(y).copy()).pin_memory()
    beta = torch.empty((nch, nsamp, P), dtype=torch.float64).pin_memory()
    sig2 = torch.empty((nch, nsamp), dtype=torch.float64).pin_memory()
    tau = torch.empty((nch, nsamp), dtype=torch.float64).pin_memory()
    status = np.zeros(nch, dtype=np.uint32)
    hyp = bb.make_hyper(alpha=alpha)
    rt = C.c_double(0)
    cast = lambda t: C.cast(t.data_ptr(), dp)
    steps = max(1, min(args.steps, args.e2e_steps))
    walls = []
    for s in range(1 + steps):
        bdist.barrier(); torch.cuda.synchronize(dev)
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
    T = bdist.max_over_ranks(float(np.sum(walls)), dev)
    return {"value": nch * world * P * ips * steps / T, "unit": UNIT,
            "h2d_bytes_per_step": int(8 * (N * P + N)), "d2h_bytes_per_step": int(8 * (nch * nsamp * (P + 2)) + 4 * nch),
            "steps": steps, "call": "bb_bridge_batched(host X,y -> host draws): upload + Gram/SVD precompute + sweep + download",
            "posterior_mean_beta0": float(beta[:, 1:, 0].mean())}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--chains", type=int, default=0, help="chains per GPU (default: the workload's)")
    ap.add_argument("--iters", type=int, default=0, help="Gibbs iterations per step (default: the workload's)")
    ap.add_argument("--burn", type=int, default=None)
    ap.add_argument("--alpha", type=float, default=0.0, help="override the workload's alpha (the c5 sweep: 0.125, 0.5, 0.9)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-ess", dest="ess", action="store_false")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3   # timing rule: at least 3 warm-up steps
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
