"""Effective sample size, the metric of the reference's benchmarks.

The reference uses coda::effectiveSize (Code/R/PublicBenchmark.R:122), which is not in its tree:
ESS = n * var(x) / S(0) with S(0) the spectral density at zero of an AR(p) fit, p chosen by AIC
(Yule-Walker).  `ess_ar` restates that definition; `ess_batch_means` is the cross-check.
Multi-chain ESS is the sum over chains; MCSE = sd / sqrt(ESS).
"""
import numpy as np


def _yule_walker(x, order_max):
    n = x.size
    xc = x - x.mean()
    acov = np.array([xc[: n - k] @ xc[k:] / n for k in range(order_max + 1)])
    if acov[0] <= 0:
        return 0, np.zeros(0), 0.0
    best = (np.inf, 0, np.zeros(0), acov[0])
    phi = np.zeros(0)
    v = acov[0]
    # Levinson-Durbin; AIC as in R's ar(): n*log(var) + 2*order
    aic0 = n * np.log(v)
    best = (aic0, 0, phi.copy(), v)
    for k in range(1, order_max + 1):
        num = acov[k] - (phi @ acov[k - 1:0:-1] if k > 1 else 0.0)
        ref = num / v
        phi = np.concatenate([phi - ref * phi[::-1], [ref]])
        v = v * (1 - ref * ref)
        if v <= 0:
            break
        aic = n * np.log(v) + 2 * k
        if aic < best[0]:
            best = (aic, k, phi.copy(), v)
    _, order, phi, v = best
    # R's ar.yw rescales the innovations variance by n/(n-(order+1))
    v = v * n / (n - (order + 1))
    return order, phi, v


def ess_ar(x):
    """coda::effectiveSize for one series."""
    x = np.asarray(x, dtype=np.float64)
    n = x.size
    var = x.var(ddof=1)
    if var == 0 or n < 4:
        return 0.0
    order_max = min(n - 1, int(np.floor(10 * np.log10(n))))
    order, phi, v = _yule_walker(x, order_max)
    s0 = v / (1 - phi.sum()) ** 2
    return float(n * var / s0)


def ess_batch_means(x, nbatch=None):
    x = np.asarray(x, dtype=np.float64)
    n = x.size
    nb = int(np.floor(np.sqrt(n))) if nbatch is None else nbatch
    b = n // nb
    if b < 1 or nb < 2:
        return 0.0
    m = x[: nb * b].reshape(nb, b).mean(axis=1)
    var = x.var(ddof=1)
    vbm = b * m.var(ddof=1)
    return float(n * var / vbm) if vbm > 0 else 0.0


def summary(draws, runtime=None, est=ess_ar):
    """sum.stat (Code/R/PublicBenchmark.R:112-134) for draws [chain, sample, P] or [sample, P]:
    mean, sd, ESS (summed over chains), ESS.sec, t, Q.1, Q.9 per coefficient."""
    d = np.asarray(draws)
    if d.ndim == 2:
        d = d[None]
    C_, M, P = d.shape
    flat = d.reshape(C_ * M, P)
    mean, sd = flat.mean(0), flat.std(0, ddof=1)
    ess = np.array([sum(est(d[c, :, j]) for c in range(C_)) for j in range(P)])
    out = {"mean": mean, "sd": sd, "ESS": ess, "t": mean / np.where(sd > 0, sd, np.nan),
           "Q.1": np.quantile(flat, 0.1, axis=0), "Q.9": np.quantile(flat, 0.9, axis=0)}
    if runtime:
        out["ESS.sec"] = ess / runtime
    return out
