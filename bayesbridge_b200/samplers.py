"""Python mirror of the reference's stand-alone sampler front ends (BridgeWrapper.R:295-537): same names
(dots -> underscores), argument order, defaults, recycling and parameter checks (which print and return NaN,
like the R code returns NA).  Each call is one launch of `scalar_kernel` (csrc/bb_scalar.cu) through the
.C-shaped symbols of include/bayesbridge_b200.h; there is no CPU fallback."""
import ctypes as C

import numpy as np

from . import _lib


def _rec(a, num):
    """array(a, num): R's recycling (BridgeWrapper.R:311-312)."""
    a = np.atleast_1d(np.asarray(a, dtype=np.float64)).ravel()
    return np.ascontiguousarray(np.resize(a, num))


def _call(name, num, *arrs, seed=None):
    L = _lib.load()
    if seed is not None:
        L.bb_set_seed(int(seed))
    x = np.zeros(num)
    dp = C.POINTER(C.c_double)
    getattr(L, name)(x.ctypes.data_as(dp), *[a.ctypes.data_as(dp) for a in arrs], C.byref(C.c_int(num)))
    return x


def _bad(msg):
    print(msg)
    return float("nan")


def rtexp_left(num=1, left=0.0, rate=1.0, seed=None):
    """rtexp.left (BridgeWrapper.R:295-315)"""
    if np.any(np.asarray(rate) <= 0):
        return _bad("rate must be >= 0.")
    if not num > 0:
        return _bad("num must be greater than zero.")
    return _call("rtexpon_rate_left", num, _rec(left, num), _rec(rate, num), seed=seed)


def rtexp_both(num=1, left=0.0, right=1.0, rate=1.0, seed=None):
    """rtexp.both (BridgeWrapper.R:317-345)"""
    if np.any(np.asarray(rate) <= 0):
        return _bad("rate must be >= 0.")
    if not num > 0:
        return _bad("num must be greater than zero.")
    if np.any(np.asarray(left) > np.asarray(right)):
        print("left must be <= right.")
    if np.any(np.isposinf(right)):
        print("right must be < infinity.")
    return _call("rtexpon_rate_both", num, _rec(left, num), _rec(right, num), _rec(rate, num), seed=seed)


def rtexp(num=1, left=0.0, right=np.inf, rate=1.0, seed=None):
    """rtexp (BridgeWrapper.R:347-375)"""
    if np.any(np.asarray(rate) <= 0):
        return _bad("rate must be > 0.")
    if not num > 0:
        return _bad("num must be greater than zero.")
    if np.any(np.asarray(left) > np.asarray(right)):
        return _bad("left must be <= right.")
    return _call("rtexpon_rate", num, _rec(left, num), _rec(right, num), _rec(rate, num), seed=seed)


def rtnorm_left(num=1, left=0.0, mu=0.0, sig=1.0, seed=None):
    """rtnorm.left (BridgeWrapper.R:380-402)"""
    if np.any(np.asarray(sig) <= 0):
        return _bad("sig must be greater than zero.")
    if not num > 0:
        return _bad("num must be greater than zero.")
    return _call("rtnorm_left", num, _rec(left, num), _rec(mu, num), _rec(sig, num), seed=seed)


def rtnorm_right(num=1, right=0.0, mu=0.0, sig=1.0, seed=None):
    """rtnorm.right (BridgeWrapper.R:404-407)"""
    return -1.0 * rtnorm_left(num=num, left=-1.0 * np.asarray(right, dtype=float), mu=-1.0 * np.asarray(mu, dtype=float),
                              sig=sig, seed=seed)


def rtnorm_both(num=1, left=-1.0, right=1.0, mu=0.0, sig=1.0, seed=None):
    """rtnorm.both (BridgeWrapper.R:409-438)"""
    if np.any(np.asarray(sig) <= 0):
        return _bad("sig must be greater than zero.")
    if np.any(np.asarray(left) >= np.asarray(right)):
        return _bad("rtnorm: left must be less than right.")
    if not num > 0:
        return _bad("rtnorm: num must be greater than zero.")
    return _call("rtnorm_both", num, _rec(left, num), _rec(right, num), _rec(mu, num), _rec(sig, num), seed=seed)


def rtruncated_norm(num=1, left=-np.inf, right=np.inf, mu=0.0, sig=1.0, seed=None):
    """rtruncated.norm (BridgeWrapper.R:440-476)"""
    if np.any(np.asarray(sig) <= 0):
        return _bad("sig must be greater than zero.")
    if np.any(np.asarray(left) > np.asarray(right)):
        return _bad("left must be less than or equal to right.")
    if not num > 0:
        return _bad("num must be greater than zero.")
    return _call("rtnorm", num, _rec(left, num), _rec(right, num), _rec(mu, num), _rec(sig, num), seed=seed)


def rtnorm(num=1, mu=0.0, sig=1.0, left=-np.inf, right=np.inf, seed=None):
    """rtnorm (BridgeWrapper.R:478-482)"""
    return rtruncated_norm(num=num, left=left, right=right, mu=mu, sig=sig, seed=seed)


def rrtgamma(num=1, shape=1.0, rate=1.0, rtrunc=1.0, scale=None, seed=None):
    """rrtgamma (BridgeWrapper.R:484-511); `scale`, when given, overrides rate = 1/scale as in the R code."""
    if scale is not None:
        rate = 1.0 / np.asarray(scale, dtype=float)
    if not np.all(np.asarray(shape) > 0):
        return _bad("shape must be greater than zero.")
    if not np.all(np.asarray(rate) > 0):
        return _bad("scale/rate must be greater than zero.")
    if not np.all(np.asarray(rtrunc) > 0):
        return _bad("rtrunc must be greater than zero.")
    return _call("rrtgamma_rate", num, _rec(shape, num), _rec(rate, num), _rec(rtrunc, num), seed=seed)


def retstable_ld(num=1, alpha=1.0, V0=1.0, h=1.0, seed=None):
    """retstable.ld (BridgeWrapper.R:513-537): S with Laplace transform exp(-V0((h+t)^alpha - h^alpha))."""
    if not np.all(np.asarray(V0) > 0):
        return _bad("V0 must be > 0.")
    if not np.all(np.asarray(h) >= 0):
        return _bad("h must be >= 0")
    if not (np.all(np.asarray(alpha) > 0) and np.all(np.asarray(alpha) <= 1)):
        return _bad("alpha must be in (0,1].")
    return _call("retstable_LD", num, _rec(alpha, num), _rec(V0, num), _rec(h, num), seed=seed)
