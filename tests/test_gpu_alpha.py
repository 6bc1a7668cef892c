"""-m gpu: the unknown-alpha path (`alpha <= 0` at the .C boundary means "estimate alpha",
BridgeWrapper.cpp:109,152,185; man/bridge.reg.Rd:33-35) -- SURVEY.md section 8, row f1.

  * the Metropolis step itself (BridgeRegression.cpp:477-503) in every kernel family, same Philox blocks as the
    C port and the compiled reference: the proposal must agree to 1e-12 and the accept/reject decision exactly;
  * the four drivers with alpha unknown: first stored samples against the port driver, including the
    (alpha_b, alpha_b) argument pair the stable driver's sampling loop passes (BridgeWrapper.cpp:294);
  * both samplers target the same posterior of alpha (the reference's acceptance criterion, Trial.R:1,98-99)."""
import numpy as np
import pytest

from helpers import rel_err, small_problem

pytestmark = pytest.mark.gpu
SEED = 0xA1FA5EED


def _engines(port, X, y):
    import bb_oracle as O
    pm0 = port.model(X, y)
    f = pm0.factors()
    rm = O.Ref().model(X, y) if O.ref_available() else None
    return rm, port.model(X, y, (f["tV"], f["d"], f["a"], f["bhat"])), f


# (method, ortho, N, P): tri one warp per chain / several warps per chain, stb64, stb_small, stb_large, the ortho kernels
FAMILIES = [("triangle", False, 200, 40), ("triangle", False, 600, 300), ("stable", False, 200, 50), ("stable", False, 300, 100),
            ("stable", False, 300, 150), ("triangle", True, 150, 30), ("stable", True, 150, 30)]


@pytest.mark.parametrize("method,ortho,N,P", FAMILIES)
def test_alpha_step_matches_oracle(port, method, ortho, N, P):
    import bayesbridge_b200 as bb
    from bayesbridge_b200._lib import ST_ALPHA
    X, y, _ = small_problem(N, P, seed=300 + P)
    rm, pm, f = _engines(port, X, y)
    model = bb.Model(X, y, factors=f)
    a0 = np.array([0.5, 0.04, 0.97, 0.3, 0.75, 0.12])          # both clamps of the proposal window are exercised
    nch = a0.size
    rng = np.random.default_rng(P)
    beta0 = 0.3 * rng.standard_normal((nch, P)) * rng.uniform(0.01, 3.0, (nch, 1))
    beta0[1, 0] = 0.0                                            # log|0| = -inf: the term vanishes, as in the reference
    tau0 = rng.uniform(0.05, 2.0, nch)
    pr_a, pr_b = 2.0, 3.5
    hyp = bb.make_hyper(alpha=0.0, alpha_a=pr_a, alpha_b=pr_b)   # alpha unknown
    ch = model.chains(method, nch, seed=SEED, chain0=7, ortho=ortho)
    moved = kept = 0
    for it in (0, 1, 2, 11, 12, 13):
        ch.set(beta=beta0, tau=tau0, alpha=a0)
        ch.run(hyp, 1, it, ST_ALPHA)
        st = ch.get()
        assert not st["status"].any()
        assert np.array_equal(st["beta"], beta0) and np.array_equal(st["tau"], tau0)     # nothing else moves
        for c in range(nch):
            for eng in filter(None, (pm, rm)):
                want = eng.sample_alpha_marg(a0[c], beta0[c], tau0[c], pr_a, pr_b, SEED, 7 + c, it)
                assert abs(st["alpha"][c] - want) < 1e-12, (it, c, st["alpha"][c], want)
            moved += st["alpha"][c] != a0[c]
            kept += st["alpha"][c] == a0[c]
    assert moved > 0 and kept > 0                                # both outcomes of the test were seen


def test_known_alpha_ignores_alpha_stage(port):
    import bayesbridge_b200 as bb
    from bayesbridge_b200._lib import ST_ALL
    X, y, _ = small_problem(120, 12, seed=9)
    model = bb.Model(X, y)
    for method in ("triangle", "stable"):
        ch = model.chains(method, 3, seed=SEED)
        ch.run(bb.make_hyper(alpha=0.37), 5, 0, ST_ALL, pre_tau=1)
        assert np.array_equal(ch.get()["alpha"], np.full(3, 0.37))


@pytest.mark.parametrize("method,ortho", [("tri", False), ("stb", False), ("tri", True), ("stb", True)])
def test_unknown_alpha_first_samples_match_port_driver(port, method, ortho):
    import bayesbridge_b200 as bb
    from bayesbridge_b200._lib import ST_ALL
    X, y, _ = small_problem(120, 12, seed=9)
    f = port.model(X, y).factors()
    pm = port.model(X, y, (f["tV"], f["d"], f["a"], f["bhat"]))
    burn, M, seed = 3, 6, 777
    pr_a, pr_b = 2.0, 3.5
    model = bb.Model(X, y, factors=f)
    ch = model.chains("triangle" if method == "tri" else "stable", 1, seed=seed, chain0=0, ortho=ortho)
    names = ("beta", "u", "w", "shape", "sig2", "tau", "alpha") if method == "tri" else ("beta", "lambda", "sig2", "tau", "alpha")
    ch.reserve(M, names)
    nburn = burn if method == "tri" else burn + 1
    pre = (3 if ortho else 1) if method == "tri" else (0 if ortho else 1)
    ch.run(bb.make_hyper(alpha=0.0, alpha_a=pr_a, alpha_b=pr_b), nburn, 0, ST_ALL, pre_tau=pre)
    ch.store_state(0)
    # sampling loop: the non-orthogonal stable driver passes (alpha_b, alpha_b) (BridgeWrapper.cpp:294)
    pa = pr_b if (method == "stb" and not ortho) else pr_a
    ch.run(bb.make_hyper(alpha=0.0, alpha_a=pa, alpha_b=pr_b), M - 1, nburn, ST_ALL, thin=1, slot0=1)
    got = {n: ch.fetch(n)[0] for n in names}
    kw = dict(burn=burn, seed=seed, ortho=int(ortho), true_alpha=0.0, alpha_a=pr_a, alpha_b=pr_b)
    want = pm.bridge_regression(M, **kw) if method == "tri" else pm.bridge_reg_stable(M, **kw)
    assert len(set(np.round(want["alpha"], 12))) > 1             # alpha really moves in this run
    for n in names:
        assert rel_err(got[n], want[n]) < 1e-6, n


def test_dropin_stb_unknown_alpha_first_samples(port):
    """bridge_reg_stable through the .C-shaped symbol with alpha = 0: device Gram instead of the port's, so the
    comparison is to 1e-6 on the first samples; checks that the drop-in driver applies the reference's schedule
    (burn+1 iterations, alpha step last, (alpha_b, alpha_b) in the sampling loop)."""
    import bayesbridge_b200 as bb
    X, y, _ = small_problem(120, 12, seed=9)
    pm = port.model(X, y)
    M, burn, seed = 6, 3, 4321
    out = bb.bridge_reg_stb(y, X, M, alpha=0.0, alpha_a=2.0, alpha_b=3.5, burn=burn, seed=seed)
    want = pm.bridge_reg_stable(M, burn=burn, seed=seed, ortho=0, true_alpha=0.0, alpha_a=2.0, alpha_b=3.5)
    for n in ("alpha", "tau", "sig2", "beta", "lambda"):
        assert rel_err(out[n], want[n]) < 1e-6, n


def test_alpha_posterior_tri_vs_stb():
    """with alpha unknown the two samplers still target one posterior: means of alpha (and tau) over 128 chains
    agree within 5 Monte-Carlo standard errors (per-chain means are independent replicates)."""
    import bayesbridge_b200 as bb
    from bayesbridge_b200 import synth
    from bayesbridge_b200._lib import ST_ALL
    X, y, _ = synth.make("c2")
    model = bb.Model(X, y)
    hyp = bb.make_hyper(alpha=0.0, alpha_a=1.0, alpha_b=1.0)
    stats = {}
    for method in ("triangle", "stable"):
        burn = 4000 if method == "triangle" else 1000
        ch = model.chains(method, 128, seed=SEED + 3)
        ch.run(hyp, burn, 0, ST_ALL, pre_tau=1)
        ch.reserve(500, ("alpha", "tau"))
        ch.run(hyp, 500 * 8, burn, ST_ALL, thin=8, slot0=0)
        assert not ch.get()["status"].any()
        for k in ("alpha", "tau"):
            cm = ch.fetch(k).mean(axis=1)
            stats[(method, k)] = (cm.mean(), cm.std(ddof=1) / np.sqrt(cm.size))
    for k in ("alpha", "tau"):
        (m0, s0), (m1, s1) = stats[("triangle", k)], stats[("stable", k)]
        assert abs(m0 - m1) / np.hypot(s0, s1) < 5.0, (k, m0, m1, s0, s1)
    assert 0.0 < stats[("stable", "alpha")][0] < 1.0
