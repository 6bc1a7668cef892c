import os, sys, numpy as np
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import bayesbridge_b200 as bb
from bayesbridge_b200 import synth
g = np.load("tests/golden/posterior_c2.npz")
X, y, _ = synth.make("c2")
P = X.shape[1]
for method in ("stable",):
    ref = g["tri" if method == "triangle" else "stb"]
    for seed in range(10, 26):
        out = bb.bridge_batched(y, X, int(g['nsamp']), method=method, nchains=128, seed=seed, burn=int(g["burn"]), alpha=float(g["alpha"]), want=("beta", "sig2", "tau"))
        M = np.concatenate([out["beta"][:, 1:], out["sig2"][:, 1:, None], out["tau"][:, 1:, None]], axis=2)
        ours = np.stack([M.mean(1), np.quantile(M, 0.1, axis=1), np.quantile(M, 0.9, axis=1)], axis=1)
        se = lambda a: a.std(0, ddof=1) / np.sqrt(a.shape[0])
        z = np.abs(ours.mean(0) - ref.mean(0)) / np.hypot(se(ours), se(ref))
        k, j = np.unravel_index(z.argmax(), z.shape)
        print(method, seed, "max z %.2f at stat %d coord %d; rms z %.3f; count z>3: %d" % (z.max(), k, j, np.sqrt((z**2).mean()), (z > 3).sum()), flush=True)
