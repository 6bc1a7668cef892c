"""-m gpu: the opt-in N x N form of the Gaussian draw of beta for wide designs (bb_stb_wide.cuh).

It is NOT the reference's arithmetic (the reference factors the P x P matrix, BridgeRegression.cpp:552-575), so it is checked
  (1) against a numpy restatement of its own formulas fed with the same Philox normals (the kernels: product on the tensor
      cores, blocked Cholesky of M, both substitutions, X'w) to 1e-10;
  (2) against the closed-form law N(VInv^-1 X'y, sig2 VInv^-1) (Notes/bbnotes.tex:369-384) by Monte Carlo; and
  (3) end to end against the default (reference-arithmetic) sampler: posterior means within Monte-Carlo error.
"""
import numpy as np
import pytest

import bb_oracle as O
from helpers import check, rel_err

pytestmark = pytest.mark.gpu
SEED = 0x5EED0F4


def _wide_problem(N, P, seed, nnz=5):
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((N, P))
    X -= X.mean(0); X /= np.linalg.norm(X, axis=0)
    bt = np.zeros(P); bt[rng.choice(P, nnz, replace=False)] = rng.normal(0, 4.0, nnz)
    y = X @ bt + 0.2 * rng.standard_normal(N)
    y -= y.mean()
    return np.asfortranarray(X), y


@pytest.mark.parametrize("N,P", [(60, 150), (200, 333), (130, 520)])
def test_wide_draw_matches_its_host_restatement(N, P):
    import bayesbridge_b200 as bb
    from bayesbridge_b200._lib import ST_BETA
    X, y = _wide_problem(N, P, P)
    model = bb.Model(X, y)
    nch, it, chain0 = 3, 12, 5
    rng = np.random.default_rng(1)
    lam = rng.gamma(1.0, 2.0, (nch, P)) + 1e-2
    sig2 = rng.uniform(0.02, 0.3, nch); tau = rng.uniform(0.05, 0.5, nch)
    ch = model.chains("stable", nch, seed=SEED, chain0=chain0).wide_gaussian()
    ch.set(beta=np.zeros((nch, P)), aux=lam, sig2=sig2, tau=tau)
    ch.run(bb.make_hyper(alpha=0.5), 1, it, ST_BETA)
    st = ch.get()
    assert not st["status"].any()
    for c in range(nch):
        eps = np.array([O.norm_at(SEED, chain0 + c, it, O.ST_BETA_STB, j) for j in range(P)])
        dlt = np.array([O.norm_at(SEED, chain0 + c, it, O.ST_BETA_STB, P + n) for n in range(N)])
        dinv = tau[c] ** 2 / (lam[c] * sig2[c])
        u = eps * np.sqrt(dinv)
        M = (X * dinv) @ X.T + np.eye(N)
        w = np.linalg.solve(M, y / np.sqrt(sig2[c]) - X @ u - dlt)
        want = np.sqrt(sig2[c]) * (u + dinv * (X.T @ w))
        check("f4 wide Gaussian draw (N x N system)", f"stb {N}x{P}", rel_err(st["beta"][c], want), 1e-10, "numpy restatement")


def test_wide_draw_has_the_reference_law():
    """beta | lambda, sig2, tau ~ N(VInv^-1 X'y, sig2 VInv^-1): 64 chains x 400 conditional draws at fixed (lambda, sig2, tau)
    against the closed-form mean and covariance, along 40 random directions and every coordinate axis."""
    import bayesbridge_b200 as bb
    from bayesbridge_b200._lib import ST_BETA
    N, P = 50, 140
    X, y = _wide_problem(N, P, 7)
    model = bb.Model(X, y)
    rng = np.random.default_rng(2)
    lam = rng.gamma(1.0, 2.0, P) + 1e-2
    sig2, tau = 0.07, 0.3
    nch, ndraw = 64, 400
    ch = model.chains("stable", nch, seed=SEED + 1).wide_gaussian()
    ch.set(beta=np.zeros((nch, P)), aux=np.tile(lam, (nch, 1)), sig2=np.full(nch, sig2), tau=np.full(nch, tau))
    ch.reserve(ndraw, ("beta",))
    ch.run(bb.make_hyper(alpha=0.5, sig2_true=sig2, tau_true=tau), ndraw, 0, ST_BETA, thin=1, slot0=0)
    draws = ch.fetch("beta").reshape(-1, P)                        # independent draws of the same Gaussian
    VInv = X.T @ X + np.diag(lam * sig2 / tau ** 2)
    Sigma = sig2 * np.linalg.inv(VInv)
    mu = np.linalg.solve(VInv, X.T @ y)
    n = draws.shape[0]
    dirs = np.vstack([np.eye(P), rng.standard_normal((40, P))])
    proj = draws @ dirs.T
    m_true, v_true = dirs @ mu, np.einsum("ip,pq,iq->i", dirs, Sigma, dirs)
    z_mean = (proj.mean(0) - m_true) / np.sqrt(v_true / n)
    z_var = (proj.var(0, ddof=1) / v_true - 1.0) / np.sqrt(2.0 / (n - 1))
    assert np.abs(z_mean).max() < 4.5, np.abs(z_mean).max()
    assert np.abs(z_var).max() < 4.5, np.abs(z_var).max()
    # and a full-covariance check: the sample covariance against Sigma in relative Frobenius norm (Monte-Carlo error ~ sqrt(P/n))
    S = np.cov(draws.T)
    assert np.linalg.norm(S - Sigma) / np.linalg.norm(Sigma) < 4.0 * np.sqrt(P / n)


def test_wide_sampler_posterior_agrees_with_the_default_sampler():
    """whole runs on a P > N problem: the N x N form against the P x P form (the reference's arithmetic): posterior means of
    beta, sig2 and tau within 4.5 Monte-Carlo standard errors, errors from the spread over independent chains."""
    import bayesbridge_b200 as bb
    from bayesbridge_b200._lib import ST_ALL
    N, P = 60, 160
    X, y = _wide_problem(N, P, 11)
    model = bb.Model(X, y)
    # a proper prior on sig2: with P > N the fit can be exact and the improper prior (shape = scale = 0, the reference's
    # default) lets sig2 collapse towards 0 on this small problem, for either form of the draw
    hyp = bb.make_hyper(alpha=0.5, sig2_shape=3.0, sig2_scale=0.2)
    stats = []
    for wide in (False, True):
        ch = model.chains("stable", 96, seed=SEED + 2 + wide)
        if wide:
            ch.wide_gaussian()
        ch.run(hyp, 300, 0, ST_ALL, pre_tau=1)
        ch.reserve(300, ("beta", "sig2", "tau"))
        ch.run(hyp, 900, 300, ST_ALL, thin=3, slot0=0)
        assert not ch.get()["status"].any()
        M = np.concatenate([ch.fetch("beta"), ch.fetch("sig2")[:, :, None], ch.fetch("tau")[:, :, None]], axis=2).mean(axis=1)
        stats.append((M.mean(0), M.std(0, ddof=1) / np.sqrt(M.shape[0])))
    z = np.abs(stats[0][0] - stats[1][0]) / np.hypot(stats[0][1], stats[1][1])
    assert z.max() < 4.5, (int(z.argmax()), z.max())


def test_wide_option_is_refused_where_it_does_not_apply():
    import bayesbridge_b200 as bb
    X, y = _wide_problem(60, 20, 3)
    model = bb.Model(X, y)
    with pytest.raises(RuntimeError, match="P > 128"):
        model.chains("stable", 2).wide_gaussian()
    with pytest.raises(RuntimeError, match="stable sampler"):
        model.chains("triangle", 2).wide_gaussian()
