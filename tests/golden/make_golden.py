"""Generates tests/golden/conditionals.npz from the reference's own C++ compiled unmodified
(oracle/_ref, built from /root/reference by oracle/Makefile).  Run in the authoring container:
    python tests/golden/make_golden.py
The .npz travels to the GPU box; /root/reference does not."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import bb_oracle as O  # noqa: E402
from helpers import small_problem  # noqa: E402

X, y, _ = small_problem(150, 24, seed=42)
ref = O.Ref()
rm = ref.model(X, y)
f = rm.factors()
seed, ch, it, alpha = 20121027, 3, 17, 0.5
rng = np.random.default_rng(7)
beta = f["bhat"] * (1 + 0.1 * rng.standard_normal(24))
u = rng.uniform(0, 0.9, 24)
tau = rm.sample_tau_marg(beta, alpha, 2.0, 2.0, seed, ch, it)
sig2 = rm.sample_sig2(beta, 0.0, 0.0, seed, ch, it)
omega, shape = rm.sample_omega(beta, u, tau, alpha, seed, ch, it)
u_new = rm.sample_u(beta, omega, tau, alpha, seed, ch, it)
beta_tri = rm.sample_beta(beta, u_new, omega, sig2, tau, alpha, seed, ch, it)
lam = rm.sample_lambda(beta, alpha, tau, seed, ch, it)
beta_stb = rm.sample_beta_stable(lam, sig2, tau, seed, ch, it)
# end-to-end drivers, the .C entry points themselves (sequential stream)
tri = ref.bridge_regression(y, X, 6000, alpha=alpha, burn=1000, seed=11)
stb = ref.bridge_reg_stable(y, X, 6000, alpha=alpha, burn=1000, seed=12)
np.savez_compressed(os.path.join(HERE, "conditionals.npz"), X=X, y=y, tV=f["tV"], d=f["d"], a=f["a"], bhat=f["bhat"],
                    seed=seed, chain=ch, iter=it, alpha=alpha, beta=beta, u=u, tau=tau, sig2=sig2, omega=omega, shape=shape,
                    u_new=u_new, beta_tri=beta_tri, beta_stb=beta_stb, **{"lambda": lam},
                    drv_tri_mean=tri["beta"].mean(0), drv_tri_sd=tri["beta"].std(0), drv_stb_mean=stb["beta"].mean(0),
                    drv_stb_sd=stb["beta"].std(0), drv_tri_sig2=tri["sig2"][1:].mean(), drv_stb_sig2=stb["sig2"][1:].mean(),
                    drv_tri_tau=tri["tau"][1:].mean(), drv_stb_tau=stb["tau"][1:].mean())
print("wrote conditionals.npz; tri vs stb posterior mean max |diff|/sd:",
      np.abs(tri["beta"].mean(0) - stb["beta"].mean(0)).max() / stb["beta"].std(0).max())
