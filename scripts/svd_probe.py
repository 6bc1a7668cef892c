"""Constructor factors on the device: time and accuracy of the SVD paths (block vs pairwise Jacobi, Cholesky-
preconditioned vs direct), against numpy's LAPACK SVD.  usage: python scripts/svd_probe.py [shape ...]   shape = NxP"""
import os, sys, time
import numpy as np
sys.path.insert(0, '.')
import bayesbridge_b200 as bb

shapes = [tuple(int(v) for v in a.split("x")) for a in sys.argv[1:]] or [(442, 64), (300, 130), (90, 17), (40, 60), (10000, 1000), (100000, 256), (500, 2000), (2500, 1000)]
rng = np.random.default_rng(7)
for N, P in shapes:
    X = rng.standard_normal((N, P)); X -= X.mean(0); X /= np.linalg.norm(X, axis=0)
    y = X[:, :3].sum(1) + 0.1 * rng.standard_normal(N)
    dref = np.linalg.svd(X, compute_uv=False)
    dref = np.concatenate([dref, np.zeros(P - len(dref))])
    XX = X.T @ X
    for env in ({}, {"BB_JACOBI_PAIRWISE": "1"}, {"BB_SVD_DIRECT": "1"}, {"BB_SVD_DIRECT": "1", "BB_JACOBI_PAIRWISE": "1"}):
        if (N < 4 * P) and "BB_SVD_DIRECT" in env:
            continue
        for k in ("BB_JACOBI_PAIRWISE", "BB_SVD_DIRECT"):
            os.environ.pop(k, None)
        os.environ.update(env)
        os.environ["BB_JACOBI_VERBOSE"] = "1"
        bb.Model(X[:50, :8], y[:50]).close()   # warm the context
        t0 = time.perf_counter()
        m = bb.Model(X, y)
        t1 = time.perf_counter()
        f = m.factors()
        t2 = time.perf_counter()
        V = f["tV"].T
        d = f["d"]
        e_d = np.abs(d - dref).max() / dref.max()
        e_rec = np.abs(V @ np.diag(d ** 2) @ V.T - XX).max() / np.abs(XX).max()
        e_orth = np.abs(V.T @ V - np.eye(P)).max()
        print(f"{N}x{P} {env}: create {1e3 * (t1 - t0):.1f} ms (+factors {1e3 * (t2 - t1):.1f} ms)  err d {e_d:.1e} recon {e_rec:.1e} orth {e_orth:.1e}", flush=True)
        m.close()
