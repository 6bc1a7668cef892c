"""Generates tests/golden/posterior_c2.npz: end-to-end posterior summaries of the reference's own C++ drivers
(oracle/_ref = /root/reference/Code/C compiled unmodified by oracle/Makefile) on the diabetes-shaped 442 x 64 design of
BASELINE configs c1 / c2, for north_star's end-to-end criterion ("posterior means and quantiles of beta, sig2 and tau
agree with the reference within 4 Monte-Carlo standard errors").

NREP independent chains per sampler (different seeds of the reference's sequential RNG shim); per chain the mean, Q.1 and
Q.9 (the quantiles of sum.stat, Code/R/PublicBenchmark.R:120-128; numpy's default = R type 7) of every beta_j, of sig2 and
of tau.  The replicates give the Monte-Carlo standard error of each statistic without any autocorrelation model.
Run in the authoring container:   python tests/golden/make_posterior_golden.py
The .npz travels to the GPU box; /root/reference does not."""
import os
import sys
import time
from multiprocessing import get_context

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "oracle"), ROOT]

NREP, BURN, NSAMP, ALPHA = 128, 1000, 10001, 0.5


def stats(out):
    """[3, P + 2]: mean, Q.1, Q.9 of beta_1..P, sig2, tau over the post-burn-in samples (sample 0 = end of burn-in, dropped)"""
    M = np.column_stack([out["beta"][1:], out["sig2"][1:], out["tau"][1:]])
    return np.stack([M.mean(0), np.quantile(M, 0.1, axis=0), np.quantile(M, 0.9, axis=0)])


def one(args):
    method, seed = args
    os.environ["OPENBLAS_NUM_THREADS"] = "1"
    import bb_oracle as O
    from bayesbridge_b200 import synth
    X, y, _ = synth.make("c2")
    ref = O.Ref()
    fn = ref.bridge_regression if method == "tri" else ref.bridge_reg_stable
    return stats(fn(y, X, NSAMP, alpha=ALPHA, burn=BURN, seed=seed))


if __name__ == "__main__":
    t0 = time.time()
    with get_context("fork").Pool(min(8, os.cpu_count() or 1)) as pool:
        tri = np.stack(pool.map(one, [("tri", 1000 + r) for r in range(NREP)]))
        stb = np.stack(pool.map(one, [("stb", 2000 + r) for r in range(NREP)]))
    np.savez_compressed(os.path.join(HERE, "posterior_c2.npz"), tri=tri, stb=stb, nrep=NREP, burn=BURN, nsamp=NSAMP, alpha=ALPHA)
    se = lambda a: a.std(0, ddof=1) / np.sqrt(a.shape[0])
    z = np.abs(tri.mean(0) - stb.mean(0)) / np.hypot(se(tri), se(stb))
    print(f"wrote posterior_c2.npz in {time.time() - t0:.0f} s; tri vs stb (the reference's own criterion, Trial.R:1): max z = {z.max():.2f}")
