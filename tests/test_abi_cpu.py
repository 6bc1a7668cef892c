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
    declared = set(re.findall(r"\b(bb_[a-z0-9_]+|bridge_regression|bridge_reg_stable)\s*\(", body))
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
