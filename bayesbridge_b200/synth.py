"""Synthetic inputs of the BASELINE.json configurations.

The reference's own data sets are not in its tree (diabetes.RData is git-ignored), so every
benchmark and test uses data of the named SHAPES generated here from fixed seeds.
"""
import numpy as np

BASE_SEED = 20121027


def _standardise(X):
    X = X - X.mean(axis=0)
    nrm = np.linalg.norm(X, axis=0)
    nrm[nrm == 0] = 1.0
    return X / nrm


def diabetes_shaped(seed=BASE_SEED + 1, N=442):
    """442 x 64: 10 base columns (AR(0.5) correlated, column 2 binary) + 45 pairwise products + 9 squares,
    centred, unit norm (the x2 construction, reference man/diabetes.Rd:16-23)."""
    rng = np.random.default_rng(seed)
    idx = np.arange(10)
    Sig = 0.5 ** np.abs(idx[:, None] - idx[None, :])
    B = rng.multivariate_normal(np.zeros(10), Sig, size=N)
    B[:, 1] = (B[:, 1] > 0).astype(float)
    cols = [B[:, i] for i in range(10)]
    for i in range(10):
        for j in range(i + 1, 10):
            cols.append(B[:, i] * B[:, j])
    for i in range(10):
        if i != 1:
            cols.append(B[:, i] ** 2)
    X = _standardise(np.stack(cols, axis=1))
    beta = np.zeros(X.shape[1])
    nz = rng.choice(X.shape[1], 8, replace=False)
    beta[nz] = rng.normal(0, 300.0, 8)
    y = X @ beta + rng.normal(0, 54.0, N)
    y = y - y.mean()
    return np.asfortranarray(X), y, beta


def gaussian_design(N, P, nnz, rho=0.3, r2=0.8, seed=BASE_SEED):
    """iid N(0,1) rows with AR(1) column correlation rho, centred, unit-norm columns; nnz non-zero
    coefficients; noise sd chosen for the requested R^2."""
    rng = np.random.default_rng(seed)
    E = rng.standard_normal((N, P))
    if rho:
        X = np.empty_like(E)
        X[:, 0] = E[:, 0]
        c = np.sqrt(1 - rho * rho)
        for j in range(1, P):
            X[:, j] = rho * X[:, j - 1] + c * E[:, j]
    else:
        X = E
    X = _standardise(X)
    beta = np.zeros(P)
    nz = rng.choice(P, nnz, replace=False)
    beta[nz] = rng.normal(0, 10.0, nnz)
    mu = X @ beta
    sd = np.sqrt(mu.var() * (1 - r2) / r2)
    y = mu + rng.normal(0, sd, N)
    y = y - y.mean()
    return np.asfortranarray(X), y, beta


CONFIGS = {
    # name: (method, N, P, alpha, chains (total at 8 GPUs or as named), generator)
    "c1": dict(method="triangle", N=442, P=64, alpha=0.5, chains=1),
    "c2": dict(method="stable", N=442, P=64, alpha=0.5, chains=512),
    "c3": dict(method="triangle", N=10000, P=1000, alpha=0.5, chains=4096),
    "c4": dict(method="stable", N=500, P=2000, alpha=0.25, chains=1024),
    "c5": dict(method="triangle", N=100000, P=256, alpha=0.5, chains=2048),
}


def make(name, seed=None):
    """(X, y, beta_true) for a named configuration."""
    c = CONFIGS[name]
    if name in ("c1", "c2"):
        return diabetes_shaped(BASE_SEED + 1 if seed is None else seed)
    if name == "c3":
        return gaussian_design(c["N"], c["P"], 50, 0.3, 0.8, BASE_SEED + 3 if seed is None else seed)
    if name == "c4":
        return gaussian_design(c["N"], c["P"], 20, 0.0, 0.9, BASE_SEED + 4 if seed is None else seed)
    if name == "c5":
        return gaussian_design(c["N"], c["P"], 30, 0.3, 0.8, BASE_SEED + 5 if seed is None else seed)
    raise KeyError(name)
