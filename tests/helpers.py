"""Shared test helpers: small synthetic problems and state generators."""
import numpy as np


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
