"""CPU: the C-ABI library loads and exports every symbol include/bayesbridge_b200.h declares, the host
mirror refuses bad parameters like the reference's R front end, and nothing in the product imports
the oracle.  No compute call is made (there is no GPU here and no CPU fallback)."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_header_symbols():
    import bayesbridge_b200 as bb
    L = bb._lib.load()
    hdr = open(os.path.join(ROOT, "include", "bayesbridge_b200.h")).read()
    body = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    # every function prototype of the header: `<type> [*]name(` at the start of a declaration
    declared = set(re.findall(r"^(?:const\s+)?(?:unsigned\s+)?(?:long\s+)*\w+\s+\**([A-Za-z_]\w*)\s*\(", body, flags=re.M))
    assert declared == set(bb._lib.EXPORTS)
    for s in declared:
        assert hasattr(L, s), s


def test_no_cpu_fallback_without_device():
    import bayesbridge_b200 as bb
    L = bb._lib.load()
    if L.bb_device_count() > 0:
        pytest.skip("a GPU is present")
    X = np.asfortranarray(np.random.default_rng(0).standard_normal((20, 3))); y = np.zeros(20)
    with pytest.raises(RuntimeError, match="no CUDA device"):
        bb.Model(X, y)
    out = bb.bridge_reg_stb(y, X, 5, burn=1)      # the drop-in symbol: NaN-filled outputs, NaN runtime, no exception
    assert np.isnan(out["beta"]).all() and np.isnan(out["runtime"])


def test_check_parameters_like_reference():
    import bayesbridge_b200 as bb
    X = np.zeros((10, 2)); y = np.zeros(9)
    with pytest.raises(ValueError):
        bb.bridge_reg_tri(y, X, 10)                 # N != R   (BridgeWrapper.R:52)
    with pytest.raises(ValueError):
        bb.bridge_reg_tri(np.zeros(10), X, 0)       # niter < 1 (:54)
    with pytest.raises(ValueError):
        bb.bridge_reg_stb(np.zeros(10), X, 5, nu_rate=-1.0)


def test_product_never_touches_the_oracle():
    pat = re.compile(r"bb_oracle|bbo_|oracle[/.]|import oracle|libbb_ref")
    for dirpath, _, files in os.walk(os.path.join(ROOT, "bayesbridge_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not pat.search(txt), os.path.join(dirpath, f)


def test_ess_estimators():
    from bayesbridge_b200 import ess
    r = np.random.default_rng(1)
    x = np.zeros(20000)
    for i in range(1, x.size):
        x[i] = 0.7 * x[i - 1] + r.standard_normal()
    want = x.size * (1 - 0.7) / (1 + 0.7)
    assert abs(ess.ess_ar(x) / want - 1) < 0.15
    assert abs(ess.ess_batch_means(x) / want - 1) < 0.35
    w = r.standard_normal(5000)
    assert abs(ess.ess_ar(w) / 5000 - 1) < 0.2
    s = ess.summary(np.stack([x[:5000], x[5000:10000]])[:, :, None], runtime=2.0)
    assert s["ESS"].shape == (1,) and s["ESS.sec"][0] == s["ESS"][0] / 2.0


def test_scalar_sampler_front_ends_check_parameters(capsys):
    """the R front ends print and return NA on bad parameters (BridgeWrapper.R:295-537); no device is touched"""
    from bayesbridge_b200 import samplers as S
    assert np.isnan(S.rtnorm(10, sig=-1.0)) and np.isnan(S.rtexp(10, rate=0.0)) and np.isnan(S.rrtgamma(10, shape=-1.0))
    assert np.isnan(S.retstable_ld(10, alpha=1.5)) and np.isnan(S.rtnorm_both(10, left=2.0, right=1.0))
    assert np.isnan(S.rtexp_left(0)) and np.isnan(S.retstable_ld(3, h=-1.0))
    assert "must be" in capsys.readouterr().out


def test_port_vector_entry_points_agree_with_the_bulk_draws():
    """bbo_draw_vec (per-item parameters, the .C symbols' argument order) against bbo_draw_many (one parameter set)"""
    import bb_oracle as O
    n = 200
    one = np.ones(n)
    cases = [("rtnorm_both", [-0.5 * one, 1.5 * one, 0.2 * one, 0.7 * one], 3, (-0.5, 1.5, 0.2, 0.7)),
             ("rtnorm_left", [1.0 * one, 0.0 * one, 2.0 * one], 4, (1.0, 0.0, 2.0)),
             ("rtexpon_rate_left", [0.3 * one, 2.0 * one], 5, (0.3, 2.0)),
             ("rtexpon_rate_both", [0.3 * one, 0.9 * one, 2.0 * one], 6, (0.3, 0.9, 2.0)),
             ("rrtgamma_rate", [2.5 * one, 1.5 * one, 1.0 * one], 7, (2.5, 1.5, 1.0)),
             ("retstable_LD", [0.25 * one, 1.0 * one, 3.0 * one], 8, (3.0, 0.25, 1.0))]
    for name, ps, kind, scalars in cases:
        assert np.array_equal(O.draw_vec(name, ps, seed=11, it=0), O.draw_many(kind, n, scalars, seed=11)), name
    # infinite bounds select the one-sided forms (BridgeWrapper.cpp:830-835, 927-938)
    inf = np.inf * one
    assert np.array_equal(O.draw_vec("rtnorm", [1.0 * one, inf, 0.0 * one, 2.0 * one], seed=11), O.draw_many(4, n, (1.0, 0.0, 2.0), seed=11))
    assert np.array_equal(O.draw_vec("rtnorm", [-inf, -1.0 * one, 0.0 * one, 2.0 * one], seed=11), -O.draw_many(4, n, (1.0, -0.0, 2.0), seed=11))
    assert np.array_equal(O.draw_vec("rtexpon_rate", [0.3 * one, inf, 2.0 * one], seed=11), O.draw_many(5, n, (0.3, 2.0), seed=11))


def test_ess_ar_against_an_independent_statement_of_ar_yw():
    """ess.py (and with it the device summaries, which are checked against it) pinned by an independent statement of what
    coda::effectiveSize computes: R's ar.yw solves the Yule-Walker equations of every order p <= floor(10 log10 n) on the
    mean-centred sample autocovariances (divisor n), takes the order with the smallest AIC = n log v_p + 2 p, rescales the
    innovation variance by n / (n - (p + 1)), and ESS = n var / S(0) with S(0) = v / (1 - sum phi)^2.  Here every order is
    solved DIRECTLY (dense Toeplitz system, numpy) instead of by the Levinson-Durbin recursion ess.py uses; the two must agree
    to rounding on every series, and on a long AR(p) both must recover the known coefficients."""
    from bayesbridge_b200 import ess

    def direct(x):
        x = np.asarray(x, float); n = x.size
        xc = x - x.mean()
        K = min(n - 1, int(np.floor(10 * np.log10(n))))
        g = np.array([xc[: n - k] @ xc[k:] / n for k in range(K + 1)])
        best = (n * np.log(g[0]), 0, np.zeros(0), g[0])
        for p in range(1, K + 1):
            R = np.array([[g[abs(i - j)] for j in range(p)] for i in range(p)])
            phi = np.linalg.solve(R, g[1:p + 1])
            v = g[0] - phi @ g[1:p + 1]
            if v <= 0:
                break
            aic = n * np.log(v) + 2 * p
            if aic < best[0]:
                best = (aic, p, phi, v)
        _, p, phi, v = best
        v *= n / (n - (p + 1))
        return n * x.var(ddof=1) * (1 - phi.sum()) ** 2 / v, p, phi

    r = np.random.default_rng(11)
    for phis, n in (([0.6], 3000), ([0.5, -0.3], 5000), ([0.2, 0.1, 0.4], 20000), ([], 1000)):
        x = np.zeros(n + 200)
        e = r.standard_normal(n + 200)
        for t in range(len(x)):
            x[t] = e[t] + sum(ph * x[t - 1 - j] for j, ph in enumerate(phis) if t - 1 - j >= 0)
        x = x[200:]
        want, p, phi = direct(x)
        np.testing.assert_allclose(ess.ess_ar(x), want, rtol=1e-9)
        if n >= 20000:                       # long series: the fitted model is the generating one
            assert p == len(phis)
            np.testing.assert_allclose(phi, phis, atol=0.03)
            np.testing.assert_allclose(want, n * (1 - sum(phis)) ** 2 * x.var(ddof=1) / 1.0, rtol=0.05)   # S(0) = sigma_e^2 / (1 - sum phi)^2
