"""-m gpu: the CUDA sampler against the oracle, conditional by conditional, on identical Philox
blocks (north_star: "each conditional draw matches the reference C++ sampler to 1e-10 relative").

The device runs one stage-masked launch over all chains; the oracle (C port, and the compiled
reference where oracle/_ref exists) computes the same conditional per chain from the same inputs.
Everything goes through the C ABI (bayesbridge_b200/_lib.py -> libbayesbridge_b200.so).
"""
import numpy as np
import pytest

import bb_oracle
from helpers import check, engines as _engines, rel_err, small_problem, teacher_forced_sweep_error

pytestmark = pytest.mark.gpu

# north_star asks for 1e-10 relative per conditional; the achieved errors (PARITY.md, written from these checks) are
# 1e-16 .. 5e-15 for the closed-form conditionals and <= 2e-13 for a single truncated-normal draw of the rotated sweep, so the
# gates sit at 1e-12 / 1e-11: 100x tighter than the contract and still >= 50x above what was measured.
TOL = 1e-12
DRAW_TOL = 1e-11
SWEEP_TOL = 2e-8

SEED = 0x1234ABCD5678


def _states(f, nchains, P, seed):
    rng = np.random.default_rng(seed)
    beta = f["bhat"][None, :] * (1 + 0.05 * rng.standard_normal((nchains, P))) + 0.02 * rng.standard_normal((nchains, P))
    u = rng.uniform(0.05, 0.9, (nchains, P))
    return beta, u


@pytest.mark.parametrize("N,P,alpha,nchains", [(60, 8, 0.5, 5), (120, 33, 0.5, 3), (442, 64, 0.5, 4), (300, 100, 0.9, 2),
                                               (400, 130, 0.125, 3), (500, 257, 0.5, 2)])
def test_tri_conditionals(port, N, P, alpha, nchains):
    import bayesbridge_b200 as bb
    from bayesbridge_b200._lib import ST_BETA, ST_OMEGA, ST_SIG2, ST_TAU, ST_U
    X, y, _ = small_problem(N, P, seed=P)
    rm, pm, f = _engines(port, X, y)
    model = bb.Model(X, y, factors=f)
    ch = model.chains("triangle", nchains, seed=SEED, chain0=7)
    beta0, u0 = _states(f, nchains, P, P)
    hyp = bb.make_hyper(alpha=alpha, betaburn=1)
    it = 11
    ch.set(beta=beta0, u=u0)
    # tau | beta
    ch.run(hyp, 1, it, ST_TAU)
    tau = ch.get()["tau"]
    cfg = f"tri {N}x{P} a={alpha}"
    engines = [(e, n) for e, n in ((pm, "port"), (rm, "reference")) if e is not None]
    for c in range(nchains):
        for eng, en in engines:
            check("a3 sample_tau_marg", cfg, rel_err(tau[c], eng.sample_tau_marg(beta0[c], alpha, 2.0, 2.0, SEED, 7 + c, it)), TOL, en)
    # sig2 | beta
    ch.run(hyp, 1, it, ST_SIG2)
    sig2 = ch.get()["sig2"]
    for c in range(nchains):
        for eng, en in engines:
            check("a4 sample_sig2", cfg, rel_err(sig2[c], eng.sample_sig2(beta0[c], 0.0, 0.0, SEED, 7 + c, it)), TOL, en)
    # omega | beta, u, tau
    ch.run(hyp, 1, it, ST_OMEGA)
    st = ch.get()
    om, sh = st["aux"], st["shape"]
    for c in range(nchains):
        for eng, en in engines:
            o, s = eng.sample_omega(beta0[c], u0[c], tau[c], alpha, SEED, 7 + c, it)
            check("a5 sample_omega", cfg, rel_err(om[c], o), TOL, en)
            assert np.array_equal(sh[c], s)
    # u | beta, omega, tau
    ch.run(hyp, 1, it, ST_U)
    u1 = ch.get()["u"]
    for c in range(nchains):
        for eng, en in engines:
            check("a6 sample_u", cfg, np.abs(u1[c] - eng.sample_u(beta0[c], om[c], tau[c], alpha, SEED, 7 + c, it)).max(), TOL, en)
    # beta | beta_prev, u, omega, sig2, tau.  (1) one sweep, every coordinate's draw checked on its own
    # against the oracle fed with the device's state before that coordinate (teacher forcing): 1e-10.
    hyp0 = bb.make_hyper(alpha=alpha, betaburn=0)
    ch.run(hyp0, 1, it, ST_BETA)
    b1 = ch.get()["beta"]
    for c in range(nchains):
        bbox = (1.0 - u1[c]) * np.exp(np.log(om[c]) / alpha) * tau[c]
        check("a7 sample_beta (per draw)", cfg, teacher_forced_sweep_error(f, beta0[c], bbox, sig2[c], SEED, 7 + c, it, b1[c]), DRAW_TOL, "port draws")
    # (2) betaburn = 1 (two sweeps, items P..2P-1 in the second) against the oracle's free-running sweep.
    # 2P sequentially dependent draws through a box whose active faces move with every draw: the map is
    # ill-conditioned, and two CPU evaluations of the SAME formula already drift apart (the port written
    # exactly like BridgeRegression.cpp:250-265 vs the compiled reference, which differ only in the
    # summation order of BLAS ddot: 1e-13 at P=64, up to 2e-4 at P=257).  The tolerance is therefore
    # calibrated per case by that CPU-vs-CPU drift; the 1e-10 gate is the teacher-forced check above.
    ch.set(beta=beta0)
    ch.run(hyp, 1, it, ST_BETA)
    b2 = ch.get()["beta"]
    for c in range(nchains):
        args = (beta0[c], u1[c], om[c], sig2[c], tau[c], alpha, SEED, 7 + c, it)
        want = pm.sample_beta(*args, betaburn=1)
        drift = rel_err(pm.sample_beta(*args, betaburn=1, as_written=1), want)
        if rm is not None:
            got = rm.sample_beta(*args, betaburn=1)
            drift = max(drift, rel_err(pm.sample_beta(*args, betaburn=1, as_written=1), got), rel_err(want, got))
            assert rel_err(b2[c], got) < max(SWEEP_TOL, 100 * drift)
        assert rel_err(b2[c], want) < max(SWEEP_TOL, 100 * drift)
    assert not ch.get()["status"].any()


@pytest.mark.parametrize("N,P,alpha,nchains", [(60, 8, 0.5, 5), (442, 64, 0.5, 4), (200, 50, 0.25, 3), (300, 100, 0.9, 2),
                                               (400, 128, 0.5, 2), (300, 150, 0.5, 3), (120, 200, 0.25, 2),
                                               (700, 321, 0.5, 2), (800, 512, 0.5, 2), (700, 530, 0.9, 2)])
def test_stb_conditionals(port, N, P, alpha, nchains):
    import bayesbridge_b200 as bb
    from bayesbridge_b200._lib import ST_BETA, ST_LAMBDA, ST_SIG2, ST_TAU
    X, y, _ = small_problem(N, P, seed=100 + P)
    rm, pm, f = _engines(port, X, y)
    model = bb.Model(X, y, factors=f)
    ch = model.chains("stable", nchains, seed=SEED, chain0=3)
    rng = np.random.default_rng(P)
    base = f["bhat"] if np.abs(f["bhat"]).max() > 0 else 0.1 * rng.standard_normal(P)
    beta0 = base[None, :] + 0.05 * rng.standard_normal((nchains, P))
    hyp = bb.make_hyper(alpha=alpha)
    it = 5
    ch.set(beta=beta0)
    ch.run(hyp, 1, it, ST_TAU)
    tau = ch.get()["tau"]
    ch.run(hyp, 1, it, ST_SIG2)
    sig2 = ch.get()["sig2"]
    ch.run(hyp, 1, it, ST_LAMBDA)
    lam = ch.get()["aux"]
    ch.run(hyp, 1, it, ST_BETA)
    st = ch.get()
    b1 = st["beta"]
    assert not st["status"].any()
    cfg = f"stb {N}x{P} a={alpha}"
    EPS = np.finfo(float).eps
    for c in range(nchains):
        # conditioning of the Gaussian draw: beta = VInv^-1 Xy + sig U^-1 eps with VInv = X'X + diag(lambda sig2/tau^2)
        # (BridgeRegression.cpp:554-556).  Two backward-stable Cholesky solves of the same system agree to ~kappa(VInv) eps;
        # the achieved error stays below kappa eps everywhere (ratio recorded, gate 8) and below 1e-12 on every tested shape.
        kappa = np.linalg.cond(f["XX"] + np.diag(lam[c] * sig2[c] / tau[c] ** 2))
        for eng, en in ((e, n) for e, n in ((pm, "port"), (rm, "reference")) if e is not None):
            check("a3 sample_tau_marg", cfg, rel_err(tau[c], eng.sample_tau_marg(beta0[c], alpha, 2.0, 2.0, SEED, 3 + c, it)), TOL, en)
            check("a4 sample_sig2", cfg, rel_err(sig2[c], eng.sample_sig2(beta0[c], 0.0, 0.0, SEED, 3 + c, it)), TOL, en)
            want = eng.sample_lambda(beta0[c], alpha, tau[c], SEED, 3 + c, it)
            check("a9 sample_lambda", cfg, np.abs(lam[c] / want - 1).max(), TOL, en)
            err = rel_err(b1[c], eng.sample_beta_stable(lam[c], sig2[c], tau[c], SEED, 3 + c, it))
            check("a10 sample_beta_stable", cfg, err, TOL, en)
            check("a10 sample_beta_stable: err / (kappa eps)", cfg, err / (kappa * EPS), 8.0, en)


def test_tri_single_coordinate_draws(port):
    """every individual truncated-normal draw of the sweep to 1e-10: with P = 1 the sweep is one draw,
    so 64 chains x 40 iterations give 2560 independent single draws over a wide range of bounds."""
    import bayesbridge_b200 as bb
    from bayesbridge_b200._lib import ST_BETA
    rng = np.random.default_rng(3)
    X = rng.standard_normal((30, 1)); X /= np.linalg.norm(X)
    y = 2.0 * X[:, 0] + 0.5 * rng.standard_normal(30)
    rm, pm, f = _engines(port, X, y)
    model = bb.Model(X, y, factors=f)
    n = 64
    ch = model.chains("triangle", n, seed=SEED)
    hyp = bb.make_hyper(alpha=0.5)
    for it in range(40):
        beta0 = rng.normal(f["bhat"][0], 1.0, (n, 1))
        u = rng.uniform(0, 0.9, (n, 1)); om = rng.gamma(2.0, 1.0, (n, 1)) + np.abs(beta0) + 0.1
        sig2 = rng.gamma(2.0, 0.2, n) ** (3 * rng.uniform(0, 1, n)); tau = np.abs(beta0[:, 0]) / ((1 - u[:, 0]) * om[:, 0] ** 2) * (1 + rng.gamma(1.0, 1.0, n))
        ch.set(beta=beta0, u=u, aux=om, sig2=sig2, tau=tau)
        ch.run(hyp, 1, it, ST_BETA)
        got = ch.get()["beta"]
        for c in range(n):
            for eng, en in ((e, n) for e, n in ((pm, "port"), (rm, "reference")) if e is not None):
                want = eng.sample_beta(beta0[c], u[c], om[c], sig2[c], tau[c], 0.5, SEED, c, it)
                check("a7 sample_beta (single draw, P=1)", "tri 30x1", rel_err(got[c], want), DRAW_TOL, en)


@pytest.mark.parametrize("N,P,alpha,nchains", [(60, 8, 0.5, 4), (120, 12, 0.5, 3), (442, 64, 0.5, 3), (300, 100, 0.9, 2), (400, 130, 0.125, 2)])
def test_ortho_conditionals(port, N, P, alpha, nchains):
    """rows a8 / a11: the ortho=TRUE conditionals on their own, against ref_sample_beta_ortho /
    ref_sample_beta_stable_ortho (BridgeRegression.cpp:362-403, :514-521) and the port, from identical inputs."""
    import bayesbridge_b200 as bb
    from bayesbridge_b200._lib import ST_BETA
    X, y, _ = small_problem(N, P, seed=300 + P)
    rm, pm, f = _engines(port, X, y)
    engines = [(e, n) for e, n in ((pm, "port"), (rm, "reference")) if e is not None]
    model = bb.Model(X, y, factors=f)
    rng = np.random.default_rng(P)
    it = 9
    hyp = bb.make_hyper(alpha=alpha)
    beta0, u0 = _states(f, nchains, P, P + 1)
    tau = np.abs(beta0).max(axis=1) * rng.uniform(0.8, 2.0, nchains)
    sig2 = rng.uniform(0.05, 0.5, nchains)
    # triangle, ortho: omega must keep beta inside the box b_j = (1-u_j) omega_j^(1/alpha) tau
    om = (np.abs(beta0) / ((1 - u0) * tau[:, None])) ** alpha + rng.gamma(1.5, 1.0, (nchains, P))
    ch = model.chains("triangle", nchains, seed=SEED, chain0=2, ortho=True)
    ch.set(beta=beta0, u=u0, aux=om, sig2=sig2, tau=tau)
    ch.run(hyp, 1, it, ST_BETA)
    st = ch.get()
    assert not st["status"].any()
    cfg = f"ortho {N}x{P} a={alpha}"
    for c in range(nchains):
        for eng, en in engines:
            want = eng.sample_beta_ortho(beta0[c], u0[c], om[c], sig2[c], tau[c], alpha, SEED, 2 + c, it)
            check("a8 sample_beta_ortho", cfg, rel_err(st["beta"][c], want), TOL, en)
    # stable, ortho: independent normals from diag(X'X)
    lam = rng.gamma(1.0, 2.0, (nchains, P)) + 1e-3
    ch = model.chains("stable", nchains, seed=SEED, chain0=2, ortho=True)
    ch.set(beta=beta0, aux=lam, sig2=sig2, tau=tau)
    ch.run(hyp, 1, it, ST_BETA)
    st = ch.get()
    assert not st["status"].any()
    for c in range(nchains):
        for eng, en in engines:
            want = eng.sample_beta_stable_ortho(lam[c], sig2[c], tau[c], SEED, 2 + c, it)
            check("a11 sample_beta_stable_ortho", cfg, rel_err(st["beta"][c], want), TOL, en)


@pytest.mark.parametrize("alpha", [0.125, 0.5, 0.9])
def test_c5_shape_stb_conditionals_alpha_sweep(port, alpha):
    """P = 256 on the blocked HBM path (stbl_* kernels), the c5 shape with fewer rows, at the three alphas of BASELINE's
    sweep: tau, sig2, lambda and the Cholesky draw against the port and the compiled reference."""
    import bayesbridge_b200 as bb
    from bayesbridge_b200 import synth
    from bayesbridge_b200._lib import ST_BETA, ST_LAMBDA, ST_SIG2, ST_TAU
    X, y, _ = synth.gaussian_design(4000, 256, 30, 0.3, 0.8, synth.BASE_SEED + 5)
    N, P = X.shape
    rm, pm, f = _engines(port, X, y)
    model = bb.Model(X, y, factors=f)
    nchains, it = 3, 6
    ch = model.chains("stable", nchains, seed=SEED, chain0=9)
    rng = np.random.default_rng(5)
    beta0 = f["bhat"][None, :] * (1 + 0.05 * rng.standard_normal((nchains, P))) + 0.02 * rng.standard_normal((nchains, P))
    hyp = bb.make_hyper(alpha=alpha)
    ch.set(beta=beta0)
    ch.run(hyp, 1, it, ST_TAU); tau = ch.get()["tau"]
    ch.run(hyp, 1, it, ST_SIG2); sig2 = ch.get()["sig2"]
    ch.run(hyp, 1, it, ST_LAMBDA); lam = ch.get()["aux"]
    ch.run(hyp, 1, it, ST_BETA); st = ch.get()
    assert not st["status"].any()
    cfg = f"c5s stb {N}x{P} a={alpha}"
    EPS = np.finfo(float).eps
    for c in range(nchains):
        kappa = np.linalg.cond(f["XX"] + np.diag(lam[c] * sig2[c] / tau[c] ** 2))
        for eng, en in ((e, n) for e, n in ((pm, "port"), (rm, "reference")) if e is not None):
            check("a3 sample_tau_marg", cfg, rel_err(tau[c], eng.sample_tau_marg(beta0[c], alpha, 2.0, 2.0, SEED, 9 + c, it)), TOL, en)
            check("a4 sample_sig2", cfg, rel_err(sig2[c], eng.sample_sig2(beta0[c], 0.0, 0.0, SEED, 9 + c, it)), TOL, en)
            want = eng.sample_lambda(beta0[c], alpha, tau[c], SEED, 9 + c, it)
            check("a9 sample_lambda", cfg, np.abs(lam[c] / want - 1).max(), TOL, en)
            err = rel_err(st["beta"][c], eng.sample_beta_stable(lam[c], sig2[c], tau[c], SEED, 9 + c, it))
            check("a10 sample_beta_stable", cfg, err, TOL, en)
            check("a10 sample_beta_stable: err / (kappa eps)", cfg, err / (kappa * EPS), 8.0, en)


def test_c1_c2_shape_conditionals_from_chain_states(port):
    """The diabetes-shaped 442 x 64 design of configs c1 / c2 (interaction columns: X'X is far worse conditioned than the
    random designs above), with the chains' OWN states after 60 iterations as inputs: the regime the samplers work in.
    sample_sig2 is the row that could suffer here (rss by expansion against the reference's direct residual)."""
    import bayesbridge_b200 as bb
    from bayesbridge_b200 import synth
    from bayesbridge_b200._lib import ST_ALL, ST_BETA, ST_LAMBDA, ST_OMEGA, ST_SIG2, ST_TAU, ST_U
    X, y, _ = synth.make("c2")
    N, P = X.shape
    rm, pm, f = _engines(port, X, y)
    engines = [(e, n) for e, n in ((pm, "port"), (rm, "reference")) if e is not None]
    model = bb.Model(X, y, factors=f)
    alpha, nchains, it = 0.5, 4, 60
    hyp = bb.make_hyper(alpha=alpha)
    EPS = np.finfo(float).eps
    # stable
    ch = model.chains("stable", nchains, seed=SEED, chain0=1)
    ch.run(hyp, it, 0, ST_ALL, pre_tau=1)
    beta0 = ch.get()["beta"]
    ch.run(hyp, 1, it, ST_TAU); tau = ch.get()["tau"]
    ch.run(hyp, 1, it, ST_SIG2); sig2 = ch.get()["sig2"]
    ch.run(hyp, 1, it, ST_LAMBDA); lam = ch.get()["aux"]
    ch.run(hyp, 1, it, ST_BETA); b1 = ch.get()["beta"]
    cfg = f"c2 stb {N}x{P} (diabetes-shaped, chain states)"
    for c in range(nchains):
        kappa = np.linalg.cond(f["XX"] + np.diag(lam[c] * sig2[c] / tau[c] ** 2))
        for eng, en in engines:
            check("a3 sample_tau_marg", cfg, rel_err(tau[c], eng.sample_tau_marg(beta0[c], alpha, 2.0, 2.0, SEED, 1 + c, it)), TOL, en)
            check("a4 sample_sig2", cfg, rel_err(sig2[c], eng.sample_sig2(beta0[c], 0.0, 0.0, SEED, 1 + c, it)), TOL, en)
            check("a9 sample_lambda", cfg, np.abs(lam[c] / eng.sample_lambda(beta0[c], alpha, tau[c], SEED, 1 + c, it) - 1).max(), TOL, en)
            err = rel_err(b1[c], eng.sample_beta_stable(lam[c], sig2[c], tau[c], SEED, 1 + c, it))
            check("a10 sample_beta_stable", cfg, err, TOL, en)
            check("a10 sample_beta_stable: err / (kappa eps)", cfg, err / (kappa * EPS), 8.0, en)
    # triangle
    ch = model.chains("triangle", nchains, seed=SEED, chain0=1)
    ch.run(hyp, it, 0, ST_ALL, pre_tau=1)
    st = ch.get()
    beta0, u0 = st["beta"], st["u"]
    ch.run(hyp, 1, it, ST_TAU); tau = ch.get()["tau"]
    ch.run(hyp, 1, it, ST_SIG2); sig2 = ch.get()["sig2"]
    ch.run(hyp, 1, it, ST_OMEGA); om = ch.get()["aux"]
    ch.run(hyp, 1, it, ST_U); u1 = ch.get()["u"]
    ch.run(hyp, 1, it, ST_BETA); b1 = ch.get()["beta"]
    cfg = f"c1 tri {N}x{P} (diabetes-shaped, chain states)"
    for c in range(nchains):
        for eng, en in engines:
            check("a3 sample_tau_marg", cfg, rel_err(tau[c], eng.sample_tau_marg(beta0[c], alpha, 2.0, 2.0, SEED, 1 + c, it)), TOL, en)
            check("a4 sample_sig2", cfg, rel_err(sig2[c], eng.sample_sig2(beta0[c], 0.0, 0.0, SEED, 1 + c, it)), TOL, en)
            o, _ = eng.sample_omega(beta0[c], u0[c], tau[c], alpha, SEED, 1 + c, it)
            check("a5 sample_omega", cfg, rel_err(om[c], o), TOL, en)
            check("a6 sample_u", cfg, np.abs(u1[c] - eng.sample_u(beta0[c], om[c], tau[c], alpha, SEED, 1 + c, it)).max(), TOL, en)
        bbox = (1.0 - u1[c]) * np.exp(np.log(om[c]) / alpha) * tau[c]
        check("a7 sample_beta (per draw)", cfg, teacher_forced_sweep_error(f, beta0[c], bbox, sig2[c], SEED, 1 + c, it, b1[c]), DRAW_TOL, "port draws")
    print("cond(X) of the diabetes-shaped design:", f["d"].max() / f["d"].min())


@pytest.mark.parametrize("kappa", [1e2, 1e4, 1e6])
def test_sig2_against_direct_residual_under_ill_conditioning(port, kappa):
    """sample_sig2 never touches X on the device (rss by expansion) while the reference forms
    ||y - X beta||^2 directly (BridgeRegression.cpp:440-443).  Designs with cond(X) = kappa, states near and far from
    the least-squares point: the achieved error is recorded together with its ratio to kappa eps."""
    import bayesbridge_b200 as bb
    from bayesbridge_b200._lib import ST_SIG2
    rng = np.random.default_rng(int(np.log10(kappa)))
    N, P = 300, 40
    Q, _ = np.linalg.qr(rng.standard_normal((N, P)))
    W, _ = np.linalg.qr(rng.standard_normal((P, P)))
    X = np.asfortranarray((Q * np.logspace(0, -np.log10(kappa), P)) @ W.T)
    y = X @ rng.standard_normal(P) + 0.1 * rng.standard_normal(N)
    rm, pm, f = _engines(port, X, y)
    model = bb.Model(X, y, factors=f)
    nchains, it = 6, 3
    scale = np.array([1e-6, 1e-3, 1e-1, 1.0, 10.0, 100.0])[:, None]
    beta0 = f["bhat"][None, :] + scale * rng.standard_normal((nchains, P)) * np.abs(f["bhat"]).max()
    hyp = bb.make_hyper(alpha=0.5)
    EPS = np.finfo(float).eps
    for method in ("triangle", "stable"):
        ch = model.chains(method, nchains, seed=SEED)
        ch.set(beta=beta0)
        ch.run(hyp, 1, it, ST_SIG2)
        sig2 = ch.get()["sig2"]
        cfg = f"{method[:3]} {N}x{P} cond(X)={kappa:.0e}"
        for c in range(nchains):
            for eng, en in ((e, n) for e, n in ((pm, "port"), (rm, "reference")) if e is not None):
                err = rel_err(sig2[c], eng.sample_sig2(beta0[c], 0.0, 0.0, SEED, c, it))
                check("a4 sample_sig2 (ill-conditioned)", cfg, err, 1e-11, en)      # measured <= 6e-13 at cond(X) = 1e6
                check("a4 sample_sig2 (ill-conditioned): err / (kappa eps)", cfg, err / (kappa * EPS), 4.0, en)
