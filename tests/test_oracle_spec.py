"""CPU: the draw specification (oracle/port/bbo_draws.h) checked independently of the sampler.

The reference's RNG library is not in its tree (parity unpinned at the draw level), so the derived
draws are OUR definition; a same-definition comparison cannot catch a shared mistake, hence these
distributional checks against scipy.stats, plus the Random123 known-answer vectors for Philox and
the only deterministic identity the reference states (TestSamplers.R:24-39)."""
import numpy as np
import pytest
from scipy import stats

import bb_oracle as O

N = 40000


def test_philox_known_answers():
    # Random123 kat_vectors, philox4x32-10
    assert O.philox([0, 0, 0, 0], [0, 0]) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert O.philox([0xffffffff] * 4, [0xffffffff] * 2) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert O.philox([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0]) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


def _ks(x, cdf):
    return stats.kstest(x, cdf).pvalue


def test_base_draws():
    u = O.draw_many(0, N)
    assert 0 < u.min() and u.max() < 1
    assert _ks(u, "uniform") > 1e-3
    z = O.draw_many(1, N, seed=2)
    assert _ks(z, "norm") > 1e-3


@pytest.mark.parametrize("shape,rate", [(1.0, 1.0), (2.0, 1.0), (0.5, 2.0), (3.7, 0.5), (130.0, 40.0), (8002.0, 3.0)])
def test_gamma(shape, rate):
    x = O.draw_many(2, N, (shape, rate), seed=3)
    assert _ks(x, stats.gamma(shape, scale=1 / rate).cdf) > 1e-3


def test_igamma():
    x = O.draw_many(9, N, (221.0, 35.0), seed=4)
    assert _ks(x, stats.invgamma(221.0, scale=35.0).cdf) > 1e-3


@pytest.mark.parametrize("l,r,m,s", [(-1, 1, 0, 1), (-0.3, 0.2, 5, 2), (2, 3, 0, 1), (2, 50, 0, 1), (-60, -40, 0, 1),
                                      (8, 8.5, 0, 1), (-5, 7, 1, 1.5), (0, 1e-3, 0, 1), (-1, 4, 10, 0.5), (37, 39, 0, 1)])
def test_tnorm(l, r, m, s):
    x = O.draw_many(3, N, (l, r, m, s), seed=5)
    assert x.min() >= l and x.max() <= r
    a, b = (l - m) / s, (r - m) / s
    if a > 30 or b < -30:   # scipy's truncnorm cdf loses accuracy in the extreme tail: compare with the exponential limit
        t = (x - m) / s
        e = (t - a) * a if a > 0 else (b - t) * (-b)
        assert abs(e.mean() - 1.0) < 0.1
    else:
        assert _ks(x, stats.truncnorm(a, b, loc=m, scale=s).cdf) > 1e-3


def test_tnorm_left():
    x = O.draw_many(4, N, (1.5, 0.0, 1.0), seed=6)
    assert x.min() >= 1.5
    assert _ks(x, stats.truncnorm(1.5, np.inf).cdf) > 1e-3
    x = O.draw_many(4, N, (-1.0, 0.5, 2.0), seed=7)
    assert _ks(x, stats.truncnorm(-0.75, np.inf, loc=0.5, scale=2.0).cdf) > 1e-3


def test_texpon():
    # TestSamplers.R:24-39: rtexp.left == left + rexp(rate) on the same stream
    left, rate = 0.7, 2.5
    x = O.draw_many(5, 1000, (left, rate), seed=8)
    u = O.draw_many(0, 1000, seed=8)
    assert np.allclose(x, left - np.log(u) / rate, rtol=0, atol=1e-14)
    x = O.draw_many(6, N, (1.0, 3.0, 0.8), seed=9)
    assert x.min() >= 1.0 and x.max() <= 3.0
    cdf = lambda t: (1 - np.exp(-0.8 * (t - 1.0))) / (1 - np.exp(-0.8 * 2.0))
    assert _ks(x, cdf) > 1e-3


def test_rtgamma():
    x = O.draw_many(7, N, (2.5, 1.5, 1.2), seed=10)
    assert x.max() <= 1.2
    g = stats.gamma(2.5, scale=1 / 1.5)
    assert _ks(x, lambda t: g.cdf(t) / g.cdf(1.2)) > 1e-3


@pytest.mark.parametrize("h,alpha", [(0.0, 0.25), (0.5, 0.25), (3.0, 0.25), (40.0, 0.25), (2.0, 0.0625), (1.0, 0.45), (500.0, 0.125)])
def test_retstable_laplace(h, alpha):
    """E exp(-t S) = exp(-((h+t)^alpha - h^alpha))  (retstable.cpp:80-84, V0 = 1)."""
    s = O.draw_many(8, N, (h, alpha, 1.0), seed=11)
    assert np.all(s > 0) and np.all(np.isfinite(s))
    for t in (0.3, 1.0, 4.0):
        want = np.exp(-((h + t) ** alpha - h ** alpha))
        got = np.exp(-t * s)
        se = got.std() / np.sqrt(N)
        assert abs(got.mean() - want) < 5 * se + 1e-12
