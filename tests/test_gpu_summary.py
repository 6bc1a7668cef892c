"""-m gpu: sum.stat on the device (bb_chains_summary) against the host restatement
bayesbridge_b200/ess.py (coda's effectiveSize definition, R's type-7 quantiles) on the same stored draws."""
import numpy as np
import pytest

from helpers import small_problem

pytestmark = pytest.mark.gpu


def _run(method, N, P, nch, nsamp, seed=3):
    import bayesbridge_b200 as bb
    from bayesbridge_b200._lib import ST_ALL
    X, y, _ = small_problem(N, P, seed=seed)
    model = bb.Model(X, y)
    ch = model.chains(method, nch, seed=77)
    hyp = bb.make_hyper(alpha=0.5)
    ch.reserve(nsamp, ("beta",))
    ch.run(hyp, 50, 0, ST_ALL, pre_tau=1)
    ch.run(hyp, nsamp, 50, ST_ALL, thin=1, slot0=0)
    return ch


@pytest.mark.parametrize("method,N,P,nch,nsamp", [("stable", 120, 9, 5, 300), ("triangle", 80, 13, 3, 257), ("stable", 442, 64, 4, 1000)])
def test_device_summary_matches_host(method, N, P, nch, nsamp):
    from bayesbridge_b200 import ess
    ch = _run(method, N, P, nch, nsamp)
    draws = ch.fetch("beta")
    host = ess.summary(draws, runtime=2.0)
    dev = ch.summary("beta", runtime=2.0, per_chain=True)
    for k in ("mean", "sd", "t", "Q.1", "Q.9"):
        np.testing.assert_allclose(dev[k], host[k], rtol=1e-10, atol=1e-12, err_msg=k)
    np.testing.assert_allclose(dev["ESS"], host["ESS"], rtol=1e-7, err_msg="ESS")
    np.testing.assert_allclose(dev["ESS.sec"], host["ESS"] / 2.0, rtol=1e-7)
    # per-chain block: mean, sd, ESS, Q.1, Q.9 of every chain's own series
    pc = dev["chains"]
    for c in range(nch):
        np.testing.assert_allclose(pc[c, 0], draws[c].mean(0), rtol=1e-11, atol=1e-13)
        np.testing.assert_allclose(pc[c, 1], draws[c].std(0, ddof=1), rtol=1e-10)
        np.testing.assert_allclose(pc[c, 2], [ess.ess_ar(draws[c, :, j]) for j in range(P)], rtol=1e-7)
        np.testing.assert_allclose(pc[c, 3], np.quantile(draws[c], 0.1, axis=0), rtol=1e-10, atol=1e-12)
        np.testing.assert_allclose(pc[c, 4], np.quantile(draws[c], 0.9, axis=0), rtol=1e-10, atol=1e-12)


def test_summary_slot_window_and_errors():
    ch = _run("stable", 100, 6, 2, 64)
    draws = ch.fetch("beta")
    s = ch.summary("beta", slot0=10, nslots=40)
    np.testing.assert_allclose(s["mean"], draws[:, 10:50].reshape(-1, 6).mean(0), rtol=1e-11, atol=1e-13)
    np.testing.assert_allclose(s["Q.9"], np.quantile(draws[:, 10:50].reshape(-1, 6), 0.9, axis=0), rtol=1e-10, atol=1e-12)
    assert "ESS.sec" not in s
    with pytest.raises(RuntimeError):
        ch.summary("beta", slot0=60, nslots=10)       # out of range
    with pytest.raises(RuntimeError):
        ch.summary("lambda")                          # not reserved


def test_summary_with_ties_and_constant_series():
    """quantiles with repeated values (the `shape` indicator of the triangle sampler is 1 or 2) and ESS = 0 for a
    constant series follow the host definitions."""
    import bayesbridge_b200 as bb
    from bayesbridge_b200 import ess
    from bayesbridge_b200._lib import ST_ALL
    X, y, _ = small_problem(90, 7, seed=5)
    model = bb.Model(X, y)
    ch = model.chains("triangle", 3, seed=5)
    hyp = bb.make_hyper(alpha=0.5)
    ch.reserve(200, ("beta", "shape"))
    ch.run(hyp, 200, 0, ST_ALL, pre_tau=1, thin=1, slot0=0)
    sh = ch.fetch("shape")
    dev = ch.summary("shape")
    flat = sh.reshape(-1, 7)
    np.testing.assert_allclose(dev["Q.1"], np.quantile(flat, 0.1, axis=0), rtol=0, atol=1e-12)
    np.testing.assert_allclose(dev["Q.9"], np.quantile(flat, 0.9, axis=0), rtol=0, atol=1e-12)
    np.testing.assert_allclose(dev["mean"], flat.mean(0), rtol=1e-12)
    host_ess = np.array([sum(ess.ess_ar(sh[c, :, j]) for c in range(3)) for j in range(7)])
    np.testing.assert_allclose(dev["ESS"], host_ess, rtol=1e-7)


def test_summary_mode_driver_equals_draws_mode():
    """bb_bridge_batched_summary == sum.stat of the draws bb_bridge_batched returns for the same seed."""
    import bayesbridge_b200 as bb
    from bayesbridge_b200 import ess
    X, y, _ = small_problem(150, 8, seed=9)
    kw = dict(method="stable", nchains=6, seed=99, burn=30, alpha=0.5)
    full = bb.bridge_batched(y, X, 120, want=("beta", "sig2", "tau"), **kw)
    sm = bb.bridge_batched_summary(y, X, 120, per_chain=True, **kw)
    host = ess.summary(full["beta"])
    for k in ("mean", "sd", "t"):
        np.testing.assert_allclose(sm[k], host[k], rtol=1e-10, atol=1e-12, err_msg=k)
    for k in ("Q.1", "Q.9"):      # the driver keeps RUNNING summaries: quantiles from the 2048-bin sketch (0.01 sd resolution)
        assert (np.abs(sm[k] - host[k]) <= 0.02 * host["sd"]).all(), k
    np.testing.assert_allclose(sm["ESS"], host["ESS"], rtol=1e-7)
    np.testing.assert_allclose(sm["ESS.sec"] * sm["runtime"], sm["ESS"], rtol=1e-12)
    assert np.array_equal(sm["sig2"], full["sig2"]) and np.array_equal(sm["tau"], full["tau"])
    assert sm["chains"].shape == (6, 5, 8) and not sm["status"].any()


@pytest.mark.parametrize("method,N,P,nch,nsamp,chunks", [("stable", 120, 9, 5, 1500, (2, 700, 1, 47, 750)), ("triangle", 80, 13, 3, 900, (300, 300, 300)),
                                                          ("stable", 442, 64, 4, 2000, (1024, 976))])
def test_running_summary_matches_stored_draw_summary(method, N, P, nch, nsamp, chunks):
    """bb_chains_summary_begin / add / end (running accumulators, draws folded chunk by chunk -- ragged chunks on purpose,
    including one of a single slot) against bb_chains_summary on the same stored draws: mean, sd, ESS are the same
    quantities recovered from lagged products (1e-9), Q.1 / Q.9 come from the 2048-bin sketch (0.02 sd)."""
    assert sum(chunks) == nsamp
    ch = _run(method, N, P, nch, nsamp)
    want = ch.summary("beta", runtime=3.0, per_chain=True)
    ch.running_begin("beta")
    s0 = 0
    for m in chunks:
        ch.running_add(s0, m); s0 += m
    got = ch.running_end(runtime=3.0, per_chain=True)
    for k in ("mean", "sd", "t"):
        np.testing.assert_allclose(got[k], want[k], rtol=1e-10, atol=1e-12, err_msg=k)
    np.testing.assert_allclose(got["ESS"], want["ESS"], rtol=1e-8, err_msg="ESS")
    np.testing.assert_allclose(got["ESS.sec"], want["ESS.sec"], rtol=1e-8)
    for k in ("Q.1", "Q.9"):
        assert (np.abs(got[k] - want[k]) <= 0.02 * want["sd"]).all(), (k, np.abs(got[k] - want[k]).max() / want["sd"].max())
    np.testing.assert_allclose(got["chains"][:, :3], want["chains"][:, :3], rtol=1e-8, atol=1e-12)
    assert np.isnan(got["chains"][:, 3:]).all()


def test_batched_summary_driver_has_no_slot_limit():
    """bb_bridge_batched_summary over 20000 samples per chain (above the 16384 slots the stored-draw summary can hold): the
    running summaries against the host summary of the same run fetched as draws."""
    import bayesbridge_b200 as bb
    from bayesbridge_b200 import ess
    X, y, _ = small_problem(100, 8, seed=5)
    kw = dict(method="stable", nchains=3, seed=41, burn=100, alpha=0.5)
    nsamp = 20000
    s = bb.bridge_batched_summary(y, X, nsamp, per_chain=True, **kw)
    d = bb.bridge_batched(y, X, nsamp, want=("beta", "sig2", "tau"), **kw)
    host = ess.summary(d["beta"], runtime=1.0)
    np.testing.assert_allclose(s["mean"], host["mean"], rtol=1e-10, atol=1e-12)
    np.testing.assert_allclose(s["sd"], host["sd"], rtol=1e-10)
    # coda's lag window is floor(10 log10 n) = 43 here (below the cap of 48): the same estimator
    np.testing.assert_allclose(s["ESS"], host["ESS"], rtol=1e-7)
    for k in ("Q.1", "Q.9"):
        assert (np.abs(s[k] - host[k]) <= 0.02 * host["sd"]).all()
    assert np.array_equal(s["sig2"], d["sig2"]) and np.array_equal(s["tau"], d["tau"])     # the scalar draws still come back, chunk by chunk
    assert s["runtime"] > 0 and not s["status"].any()
