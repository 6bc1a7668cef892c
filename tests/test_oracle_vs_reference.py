"""CPU: the C restatement (oracle/port) against the reference's own C++ compiled unmodified
(oracle/_ref), conditional by conditional on identical Philox blocks, and against the committed
golden vectors generated from that compiled reference (tests/golden/make_golden.py)."""
import os

import numpy as np
import pytest

import bb_oracle as O
from helpers import rel_err, small_problem

GOLD = os.path.join(os.path.dirname(__file__), "golden", "conditionals.npz")
SEED = 987654321


def _case(port, ref, N, P, seed):
    X, y, _ = small_problem(N, P, seed=seed)
    rm = ref.model(X, y)
    f = rm.factors()
    pm = port.model(X, y, (f["tV"], f["d"], f["a"], f["bhat"]))
    return X, y, rm, pm, f


@pytest.mark.parametrize("N,P,alpha", [(50, 6, 0.5), (442, 64, 0.5), (90, 30, 0.125), (200, 17, 0.9), (40, 40, 0.5)])
def test_conditionals_match(port, ref, N, P, alpha):
    X, y, rm, pm, f = _case(port, ref, N, P, P)
    rng = np.random.default_rng(P)
    for trial in range(3):
        beta = f["bhat"] * (1 + 0.1 * rng.standard_normal(P)) + 0.01 * rng.standard_normal(P)
        if trial == 2:
            beta[0] = 0.0   # h = 0 in retstable, |beta|^alpha = 0 in tau
        u = rng.uniform(0, 0.95, P)
        ch, it = trial, 10 + trial
        tau = rm.sample_tau_marg(beta, alpha, 2.0, 2.0, SEED, ch, it)
        assert rel_err(pm.sample_tau_marg(beta, alpha, 2.0, 2.0, SEED, ch, it), tau) < 1e-13
        sig2 = rm.sample_sig2(beta, 0.0, 0.0, SEED, ch, it)
        assert rel_err(pm.sample_sig2(beta, 0.0, 0.0, SEED, ch, it), sig2) < 1e-12
        if trial < 2:
            o_r, s_r = rm.sample_omega(beta, u, tau, alpha, SEED, ch, it)
            o_p, s_p = pm.sample_omega(beta, u, tau, alpha, SEED, ch, it)
            assert rel_err(o_p, o_r) < 1e-13 and np.array_equal(s_p, s_r)
            u_r = rm.sample_u(beta, o_r, tau, alpha, SEED, ch, it)
            assert np.abs(pm.sample_u(beta, o_r, tau, alpha, SEED, ch, it) - u_r).max() < 1e-14
            b_r = rm.sample_beta(beta, u_r, o_r, sig2, tau, alpha, SEED, ch, it, betaburn=1)
            assert rel_err(pm.sample_beta(beta, u_r, o_r, sig2, tau, alpha, SEED, ch, it, betaburn=1, as_written=1), b_r) < 1e-11
            assert rel_err(pm.sample_beta(beta, u_r, o_r, sig2, tau, alpha, SEED, ch, it, betaburn=1, as_written=0), b_r) < 1e-10
            bo_r = rm.sample_beta_ortho(beta, u_r, o_r, sig2, tau, alpha, SEED, ch, it)
            assert rel_err(pm.sample_beta_ortho(beta, u_r, o_r, sig2, tau, alpha, SEED, ch, it), bo_r) < 1e-11
        l_r = rm.sample_lambda(beta, alpha, tau, SEED, ch, it)
        assert np.abs(pm.sample_lambda(beta, alpha, tau, SEED, ch, it) / l_r - 1).max() < 1e-12
        g_r = rm.sample_beta_stable(l_r, sig2, tau, SEED, ch, it)
        assert rel_err(pm.sample_beta_stable(l_r, sig2, tau, SEED, ch, it), g_r) < 1e-10
        g_r = rm.sample_beta_stable_ortho(l_r, sig2, tau, SEED, ch, it)
        assert rel_err(pm.sample_beta_stable_ortho(l_r, sig2, tau, SEED, ch, it), g_r) < 1e-13
        a_r = rm.sample_alpha_marg(0.5, beta + 1e-3, tau, 1.0, 1.0, SEED, ch, it)
        assert abs(pm.sample_alpha_marg(0.5, beta + 1e-3, tau, 1.0, 1.0, SEED, ch, it) - a_r) < 1e-14


def test_factors_agree(port, ref):
    """the port's Jacobi SVD against the reference's LAPACK SVD: same d, same |V|, same |a|, same bhat."""
    X, y, rm, _, f = _case(port, ref, 120, 20, 5)
    g = port.model(X, y).factors()
    assert rel_err(g["d"], f["d"]) < 1e-12
    assert rel_err(g["XX"], f["XX"]) < 1e-13 and rel_err(g["Xy"], f["Xy"]) < 1e-13
    assert rel_err(g["bhat"], f["bhat"]) < 1e-10
    assert rel_err(np.abs(g["tV"]), np.abs(f["tV"])) < 1e-9
    assert rel_err(np.abs(g["a"]), np.abs(f["a"])) < 1e-9


def test_retstable_helpers(ref):
    # sinc_MM branches (retstable.cpp:18-29)
    for x in (0.0, 1e-5, 1e-3, 0.005, 0.1, 2.0):
        want = 1.0 if x == 0 else np.sin(x) / x
        assert abs(ref.L.ref_sinc_MM(x) - want) < 1e-12


def test_golden_vectors(port):
    """port == golden outputs of the compiled reference (generated once by tests/golden/make_golden.py)."""
    g = np.load(GOLD)
    X, y = g["X"], g["y"]
    pm = port.model(X, y, (g["tV"], g["d"], g["a"], g["bhat"]))
    seed, ch, it, alpha = int(g["seed"]), int(g["chain"]), int(g["iter"]), float(g["alpha"])
    beta, u = g["beta"], g["u"]
    assert rel_err(pm.sample_tau_marg(beta, alpha, 2.0, 2.0, seed, ch, it), g["tau"]) < 1e-13
    assert rel_err(pm.sample_sig2(beta, 0.0, 0.0, seed, ch, it), g["sig2"]) < 1e-12
    o, s = pm.sample_omega(beta, u, float(g["tau"]), alpha, seed, ch, it)
    assert rel_err(o, g["omega"]) < 1e-13 and np.array_equal(s, g["shape"])
    assert np.abs(pm.sample_u(beta, g["omega"], float(g["tau"]), alpha, seed, ch, it) - g["u_new"]).max() < 1e-14
    assert rel_err(pm.sample_beta(beta, g["u_new"], g["omega"], float(g["sig2"]), float(g["tau"]), alpha, seed, ch, it), g["beta_tri"]) < 1e-10
    assert np.abs(pm.sample_lambda(beta, alpha, float(g["tau"]), seed, ch, it) / g["lambda"] - 1).max() < 1e-12
    assert rel_err(pm.sample_beta_stable(g["lambda"], float(g["sig2"]), float(g["tau"]), seed, ch, it), g["beta_stb"]) < 1e-10
    # the reference's drivers end to end (sequential stream): posterior means recorded for the statistical tests
    assert g["drv_tri_mean"].shape == beta.shape and g["drv_stb_mean"].shape == beta.shape
