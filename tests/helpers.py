"""Shared test helpers: small synthetic problems and state generators."""
import os

import numpy as np

# Achieved parity errors of this pytest session: (row, config, against) -> max error seen.  tests/conftest.py writes them to
# gpurun_out/parity_errors.json at the end of the session; scripts/make_parity_md.py turns that file into PARITY.md.
PARITY = {}


def check(row, config, err, tol, against="port"):
    """Record the achieved error of one SURVEY section-8 row and assert it against `tol`.
    BB_PARITY_RECORD_ONLY=1 records without asserting (used once to calibrate the tolerances)."""
    err = float(err)
    key = (row, str(config), against)
    old = PARITY.get(key)
    PARITY[key] = {"err": max(err, old["err"]) if old else err, "tol": tol, "n": (old["n"] + 1) if old else 1}
    if not os.environ.get("BB_PARITY_RECORD_ONLY"):
        assert err < tol, f"{row} [{config}] vs {against}: {err:.3e} >= {tol:.1e}"


def small_problem(N, P, seed=0, rho=0.4, nnz=3, noise=0.3):
    rng = np.random.default_rng(seed)
    E = rng.standard_normal((N, P))
    X = E.copy()
    for j in range(1, P):
        X[:, j] = rho * X[:, j - 1] + np.sqrt(1 - rho * rho) * E[:, j]
    X -= X.mean(0)
    X /= np.linalg.norm(X, axis=0)
    bt = np.zeros(P)
    bt[rng.choice(P, min(nnz, P), replace=False)] = rng.normal(0, 3.0, min(nnz, P))
    y = X @ bt + noise * rng.standard_normal(N)
    y -= y.mean()
    return np.asfortranarray(X), y, bt


def rel_err(a, b):
    a, b = np.asarray(a, dtype=float), np.asarray(b, dtype=float)
    scale = max(np.abs(b).max(), 1e-300)
    return np.abs(a - b).max() / scale


def teacher_forced_sweep_error(f, beta0, b, sig2, seed, chain, it, beta_dev):
    """Per-coordinate check of one rotated sweep (BridgeRegression.cpp:250-283): coordinate i of the
    oracle is drawn from the DEVICE's state before step i (its z_0..z_{i-1}), on the same Philox block,
    and compared with the device's z_i.  Returns max_i |z_i(oracle) - z_i(device)| / scale.
    This isolates every single truncated-normal draw; rounding differences cannot compound."""
    import bb_oracle as O
    tV, d, a = np.asarray(f["tV"]), np.asarray(f["d"]), np.asarray(f["a"])
    P = beta0.size
    z = tV @ beta0
    zdev = tV @ beta_dev
    cur = beta0.astype(float).copy()
    sig = np.sqrt(sig2)
    worst, scale = 0.0, max(np.abs(zdev).max(), 1e-300)
    for i in range(P):
        v = tV[i, :]                       # V(j,i) over j
        r = cur - v * z[i]
        with np.errstate(divide="ignore", invalid="ignore"):
            left = np.where(v > 0, -(b + r), -(b - r)) / np.abs(v)
            right = np.where(v > 0, b - r, b + r) / np.abs(v)
        lmax, rmin = np.nanmax(left), np.nanmin(right)
        if d[i] > 1e-16:
            want = O.tnorm_at(seed, chain, it, O.ST_BETA, i, lmax, rmin, a[i] / (d[i] * d[i]), sig / d[i])
        else:
            want = O.flat_at(seed, chain, it, O.ST_BETA, i, lmax, rmin)
        worst = max(worst, abs(want - zdev[i]) / scale)
        cur += v * (zdev[i] - z[i])
        z[i] = zdev[i]
    return worst


def engines(port, X, y):
    """(reference model | None, port model, factors): the reference's LAPACK factors when oracle/_ref is built, otherwise
    the port's Jacobi factors; the port model always runs on the same factors as the device."""
    import bb_oracle
    if bb_oracle.ref_available():
        rm = bb_oracle.Ref().model(X, y)
        f = rm.factors()
    else:
        rm, f = None, port.model(X, y).factors()
    pm = port.model(X, y, (f["tV"], f["d"], f["a"], f["bhat"]))
    return rm, pm, f
