"""ctypes bindings for the two CPU checkers -- ORACLE, test infrastructure only.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  Nothing under bayesbridge_b200/ does.

  Port  : oracle/port/libbbo_port.so  (plain-C restatement, bbo_port.c)
  Ref   : oracle/_ref/libbb_ref.so    (the reference's own C++ compiled unmodified against
                                        oracle/compat shims; absent if never built)
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_dp = C.POINTER(C.c_double)

ST_TAU, ST_SIG2, ST_OMEGA, ST_U, ST_BETA, ST_LAMBDA, ST_BETA_STB, ST_TAU_INIT, ST_ALPHA, ST_SCALAR = range(1, 11)


def _p(a):
    return None if a is None else a.ctypes.data_as(_dp)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def build(force=False):
    """(Re)build the checkers with oracle/Makefile (needs gcc/g++; _ref only where /root/reference exists)."""
    if force or not os.path.exists(os.path.join(_HERE, "port", "libbbo_port.so")) or (
            os.path.isdir("/root/reference/Code/C") and not os.path.exists(os.path.join(_HERE, "_ref", "libbb_ref.so"))):
        subprocess.run(["make", "-C", _HERE, "all"], check=True, stdout=subprocess.DEVNULL)


class _Opts(C.Structure):
    _fields_ = [("sig2_shape", C.c_double), ("sig2_scale", C.c_double), ("nu_shape", C.c_double), ("nu_rate", C.c_double),
                ("alpha_a", C.c_double), ("alpha_b", C.c_double), ("true_sig2", C.c_double), ("true_tau", C.c_double),
                ("true_alpha", C.c_double), ("burn", C.c_int), ("betaburn", C.c_int), ("ortho", C.c_int),
                ("as_written", C.c_int), ("seed", C.c_uint64), ("chain", C.c_uint32)]


class Port:
    """The plain-C restatement."""

    def __init__(self):
        build()
        L = C.CDLL(os.path.join(_HERE, "port", "libbbo_port.so"))
        L.bbo_model_create.restype = C.c_void_p
        L.bbo_model_create.argtypes = [_dp, _dp, C.c_int, C.c_int]
        L.bbo_model_create_with_factors.restype = C.c_void_p
        L.bbo_model_create_with_factors.argtypes = [_dp, _dp, C.c_int, C.c_int, _dp, _dp, _dp, _dp]
        L.bbo_model_free.argtypes = [C.c_void_p]
        L.bbo_model_get.argtypes = [C.c_void_p] + [_dp] * 6
        for name in ("bbo_sample_tau_marg", "bbo_sample_sig2", "bbo_retstable_LD", "bbo_sample_alpha_marg",
                     "bbo_bridge_regression", "bbo_bridge_reg_stable"):
            getattr(L, name).restype = C.c_double
        L.bbo_sample_tau_marg.argtypes = [C.c_void_p, _dp, C.c_double, C.c_double, C.c_double, C.c_void_p]
        L.bbo_sample_sig2.argtypes = [C.c_void_p, _dp, C.c_double, C.c_double, C.c_void_p]
        L.bbo_sample_omega.argtypes = [C.c_void_p, _dp, _dp, _dp, _dp, C.c_double, C.c_double, C.c_void_p]
        L.bbo_sample_u.argtypes = [C.c_void_p, _dp, _dp, _dp, C.c_double, C.c_double, C.c_void_p]
        L.bbo_rtnorm_gibbs.argtypes = [C.c_void_p, _dp, _dp, C.c_double, C.c_int, C.c_uint32, C.c_void_p]
        L.bbo_sample_beta.argtypes = [C.c_void_p, _dp, _dp, _dp, _dp, C.c_double, C.c_double, C.c_double, C.c_int, C.c_int,
                                      C.c_uint32, C.c_void_p]
        L.bbo_sample_beta_ortho.argtypes = [C.c_void_p, _dp, _dp, _dp, _dp, C.c_double, C.c_double, C.c_double, C.c_void_p]
        L.bbo_retstable_LD.argtypes = [C.c_double, C.c_double, C.c_double, C.c_void_p]
        L.bbo_sample_lambda.argtypes = [C.c_void_p, _dp, _dp, C.c_double, C.c_double, C.c_void_p]
        L.bbo_sample_beta_stable.restype = C.c_int
        L.bbo_sample_beta_stable.argtypes = [C.c_void_p, _dp, _dp, C.c_double, C.c_double, C.c_void_p]
        L.bbo_sample_beta_stable_ortho.argtypes = [C.c_void_p, _dp, _dp, C.c_double, C.c_double, C.c_void_p]
        L.bbo_sample_alpha_marg.argtypes = [C.c_void_p, C.c_double, _dp, C.c_double, C.c_double, C.c_double, C.c_void_p]
        L.bbo_bridge_regression.argtypes = [C.c_void_p, C.POINTER(_Opts), C.c_int] + [_dp] * 7
        L.bbo_bridge_reg_stable.argtypes = [C.c_void_p, C.POINTER(_Opts), C.c_int] + [_dp] * 5
        self.L = L

    class _Rng(C.Structure):
        _fields_ = [("seed_lo", C.c_uint32), ("seed_hi", C.c_uint32), ("chain", C.c_uint32), ("iter", C.c_uint32),
                    ("stage", C.c_uint32), ("item", C.c_uint32), ("k", C.c_uint32), ("seq", C.c_uint64),
                    ("sequential", C.c_int)]

    def rng(self, seed, chain, it, stage, item=0):
        r = Port._Rng(seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF, chain, it, stage, item, 0, 0, 0)
        return r

    def model(self, X, y, factors=None):
        return PortModel(self, X, y, factors)


class PortModel:
    def __init__(self, port, X, y, factors=None):
        self.port, self.L = port, port.L
        self.X = np.asfortranarray(X, dtype=np.float64)
        self.y = _f64(y)
        self.N, self.P = self.X.shape
        if factors is None:
            self.h = self.L.bbo_model_create(_p(self.X), _p(self.y), self.N, self.P)
        else:
            tV, d, a, bhat = [None if f is None else np.asfortranarray(f, dtype=np.float64) for f in factors]
            self._keep = (tV, d, a, bhat)
            self.h = self.L.bbo_model_create_with_factors(_p(self.X), _p(self.y), self.N, self.P, _p(tV), _p(d), _p(a), _p(bhat))

    def __del__(self):
        if getattr(self, "h", None):
            self.L.bbo_model_free(self.h)
            self.h = None

    def factors(self):
        P = self.P
        XX = np.zeros((P, P), order="F"); tV = np.zeros((P, P), order="F")
        Xy, bhat, d, a = (np.zeros(P) for _ in range(4))
        self.L.bbo_model_get(self.h, _p(XX), _p(Xy), _p(bhat), _p(tV), _p(d), _p(a))
        return dict(XX=XX, Xy=Xy, bhat=bhat, tV=tV, d=d, a=a)

    def _r(self, seed, chain, it, stage):
        return self.port.rng(seed, chain, it, stage)

    def sample_tau_marg(self, beta, alpha, nu_shape, nu_rate, seed, chain, it, stage=ST_TAU):
        r = self._r(seed, chain, it, stage)
        return self.L.bbo_sample_tau_marg(self.h, _p(_f64(beta)), alpha, nu_shape, nu_rate, C.byref(r))

    def sample_sig2(self, beta, sig2_shape, sig2_scale, seed, chain, it):
        r = self._r(seed, chain, it, ST_SIG2)
        return self.L.bbo_sample_sig2(self.h, _p(_f64(beta)), sig2_shape, sig2_scale, C.byref(r))

    def sample_omega(self, beta, u, tau, alpha, seed, chain, it):
        r = self._r(seed, chain, it, ST_OMEGA)
        om, sh = np.zeros(self.P), np.zeros(self.P)
        self.L.bbo_sample_omega(self.h, _p(om), _p(sh), _p(_f64(beta)), _p(_f64(u)), tau, alpha, C.byref(r))
        return om, sh

    def sample_u(self, beta, omega, tau, alpha, seed, chain, it):
        r = self._r(seed, chain, it, ST_U)
        u = np.zeros(self.P)
        self.L.bbo_sample_u(self.h, _p(u), _p(_f64(beta)), _p(_f64(omega)), tau, alpha, C.byref(r))
        return u

    def rtnorm_gibbs(self, beta, b, sig2, seed, chain, it, as_written=0, item_base=0):
        r = self._r(seed, chain, it, ST_BETA)
        out = _f64(beta).copy()
        self.L.bbo_rtnorm_gibbs(self.h, _p(out), _p(_f64(b)), sig2, as_written, item_base, C.byref(r))
        return out

    def sample_beta(self, beta_prev, u, omega, sig2, tau, alpha, seed, chain, it, betaburn=0, as_written=0):
        r = self._r(seed, chain, it, ST_BETA)
        out = np.zeros(self.P)
        self.L.bbo_sample_beta(self.h, _p(out), _p(_f64(beta_prev)), _p(_f64(u)), _p(_f64(omega)), sig2, tau, alpha, betaburn,
                               as_written, it, C.byref(r))
        return out

    def sample_beta_ortho(self, beta_prev, u, omega, sig2, tau, alpha, seed, chain, it):
        r = self._r(seed, chain, it, ST_BETA)
        out = np.zeros(self.P)
        self.L.bbo_sample_beta_ortho(self.h, _p(out), _p(_f64(beta_prev)), _p(_f64(u)), _p(_f64(omega)), sig2, tau, alpha, C.byref(r))
        return out

    def sample_lambda(self, beta, alpha, tau, seed, chain, it):
        r = self._r(seed, chain, it, ST_LAMBDA)
        lam = np.zeros(self.P)
        self.L.bbo_sample_lambda(self.h, _p(lam), _p(_f64(beta)), alpha, tau, C.byref(r))
        return lam

    def sample_beta_stable(self, lam, sig2, tau, seed, chain, it):
        r = self._r(seed, chain, it, ST_BETA_STB)
        out = np.zeros(self.P)
        rc = self.L.bbo_sample_beta_stable(self.h, _p(out), _p(_f64(lam)), sig2, tau, C.byref(r))
        if rc:
            raise FloatingPointError("Cholesky failed")
        return out

    def sample_beta_stable_ortho(self, lam, sig2, tau, seed, chain, it):
        r = self._r(seed, chain, it, ST_BETA_STB)
        out = np.zeros(self.P)
        self.L.bbo_sample_beta_stable_ortho(self.h, _p(out), _p(_f64(lam)), sig2, tau, C.byref(r))
        return out

    def sample_alpha_marg(self, alpha_prev, beta, tau, pr_a, pr_b, seed, chain, it):
        r = self._r(seed, chain, it, ST_ALPHA)
        return self.L.bbo_sample_alpha_marg(self.h, alpha_prev, _p(_f64(beta)), tau, pr_a, pr_b, C.byref(r))

    def _opts(self, kw):
        d = dict(sig2_shape=0.0, sig2_scale=0.0, nu_shape=2.0, nu_rate=2.0, alpha_a=1.0, alpha_b=1.0, true_sig2=0.0,
                 true_tau=0.0, true_alpha=0.5, burn=500, betaburn=0, ortho=0, as_written=0, seed=1, chain=0)
        d.update(kw)
        return _Opts(**d)

    def bridge_regression(self, M, **kw):
        """BridgeWrapper.cpp:80-204 / 320-431.  Returns dict of (M,P)/(M,) arrays + runtime."""
        o = self._opts(kw); P = self.P
        beta, u, om, sh = (np.zeros((M, P)) for _ in range(4))
        s2, tau, al = (np.zeros(M) for _ in range(3))
        rt = self.L.bbo_bridge_regression(self.h, C.byref(o), M, _p(beta), _p(u), _p(om), _p(sh), _p(s2), _p(tau), _p(al))
        return dict(beta=beta, u=u, w=om, shape=sh, sig2=s2, tau=tau, alpha=al, runtime=rt)

    def bridge_reg_stable(self, M, **kw):
        o = self._opts(kw); P = self.P
        beta, lam = np.zeros((M, P)), np.zeros((M, P))
        s2, tau, al = (np.zeros(M) for _ in range(3))
        rt = self.L.bbo_bridge_reg_stable(self.h, C.byref(o), M, _p(beta), _p(lam), _p(s2), _p(tau), _p(al))
        return dict(beta=beta, **{"lambda": lam}, sig2=s2, tau=tau, alpha=al, runtime=rt)

    def retstable_LD(self, h, alpha, V0, seed, chain, it, item):
        r = self.port.rng(seed, chain, it, ST_LAMBDA, item)
        return self.L.bbo_retstable_LD(h, alpha, V0, C.byref(r))


def ref_available():
    return os.path.exists(os.path.join(_HERE, "_ref", "libbb_ref.so"))


class Ref:
    """The reference's own C++ (unmodified sources + compat shims), single-threaded BLAS."""

    def __init__(self):
        build()
        os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")
        L = C.CDLL(os.path.join(_HERE, "_ref", "libbb_ref.so"))
        ull, ui = C.c_ulonglong, C.c_uint
        L.ref_create.restype = C.c_void_p
        L.ref_create.argtypes = [_dp, _dp, C.c_int, C.c_int]
        L.ref_free.argtypes = [C.c_void_p]
        L.ref_get_factors.argtypes = [C.c_void_p] + [_dp] * 6
        for n in ("ref_sample_tau_marg", "ref_sample_sig2", "ref_sample_alpha_marg", "ref_retstable_LD", "ref_sinc_MM", "ref_A_", "ref_BdB0"):
            getattr(L, n).restype = C.c_double
        L.ref_sample_tau_marg.argtypes = [C.c_void_p, _dp, C.c_double, C.c_double, C.c_double, ull, ui, ui, ui]
        L.ref_sample_sig2.argtypes = [C.c_void_p, _dp, C.c_double, C.c_double, ull, ui, ui]
        L.ref_sample_omega.argtypes = [C.c_void_p, _dp, _dp, _dp, _dp, C.c_double, C.c_double, ull, ui, ui]
        L.ref_sample_u.argtypes = [C.c_void_p, _dp, _dp, _dp, C.c_double, C.c_double, ull, ui, ui]
        L.ref_sample_beta.argtypes = [C.c_void_p, _dp, _dp, _dp, _dp, C.c_double, C.c_double, C.c_double, C.c_int, ull, ui, ui]
        L.ref_rtnorm_gibbs.argtypes = [_dp, _dp, _dp, _dp, _dp, C.c_double, C.c_int, ull, ui, ui]
        L.ref_sample_beta_ortho.argtypes = [C.c_void_p, _dp, _dp, _dp, _dp, C.c_double, C.c_double, C.c_double, ull, ui, ui]
        L.ref_sample_lambda.argtypes = [C.c_void_p, _dp, _dp, C.c_double, C.c_double, ull, ui, ui]
        L.ref_sample_beta_stable.restype = C.c_int
        L.ref_sample_beta_stable.argtypes = [C.c_void_p, _dp, _dp, C.c_double, C.c_double, C.c_double, ull, ui, ui]
        L.ref_sample_beta_stable_ortho.argtypes = [C.c_void_p, _dp, _dp, C.c_double, C.c_double, C.c_double, ull, ui, ui]
        L.ref_sample_alpha_marg.argtypes = [C.c_void_p, C.c_double, _dp, C.c_double, C.c_double, C.c_double, ull, ui, ui]
        L.ref_retstable_LD.argtypes = [C.c_double, C.c_double, C.c_double, ull, ui, ui, ui]
        L.ref_sinc_MM.argtypes = [C.c_double]
        L.ref_A_.argtypes = [C.c_double, C.c_double]
        L.ref_BdB0.argtypes = [C.c_double, C.c_double]
        L.bb_compat_set_seed.argtypes = [ull, ui]
        ip = C.POINTER(C.c_int)
        L.bridge_regression.argtypes = [_dp] * 18 + [ip] * 4 + [_dp, ip, ip, ip]
        L.bridge_reg_stable.argtypes = [_dp] * 16 + [ip] * 4 + [_dp, ip]
        self.L = L

    def model(self, X, y):
        return RefModel(self, X, y)

    # ---- the reference's extern "C" entry points, exactly as R's .C() calls them
    #      (BridgeWrapper.R:170-179, 220-228; BridgeWrapper.cpp:572-656, 659-732) ----
    def bridge_regression(self, y, X, nsamp, alpha=0.5, sig2_shape=0.0, sig2_scale=0.0, nu_shape=2.0, nu_rate=2.0,
                          alpha_a=1.0, alpha_b=1.0, sig2_true=0.0, tau_true=0.0, burn=500, ortho=False, betaburn=0,
                          seed=1, chain=0, quiet=True):
        X = np.asfortranarray(X, dtype=np.float64); y = _f64(y)
        N, P = X.shape; M = nsamp
        beta, u, om, sh = (np.zeros((M, P)) for _ in range(4))
        s2, tau, al = (np.zeros(M) for _ in range(3))
        d = lambda v: C.byref(C.c_double(v))
        i = lambda v: C.byref(C.c_int(int(v)))
        rt = C.c_double(0)
        self.L.bb_compat_set_seed(seed, chain)
        with _quiet(quiet):
            self.L.bridge_regression(_p(beta), _p(u), _p(om), _p(sh), _p(s2), _p(tau), _p(al), _p(y), _p(X),
                                     d(sig2_shape), d(sig2_scale), d(nu_shape), d(nu_rate), d(alpha_a), d(alpha_b),
                                     d(sig2_true), d(tau_true), d(alpha), i(P), i(N), i(M), i(burn), C.byref(rt),
                                     i(ortho), i(betaburn), i(0))
        return dict(beta=beta, u=u, w=om, shape=sh, sig2=s2, tau=tau, alpha=al, runtime=rt.value)

    def bridge_reg_stable(self, y, X, nsamp, alpha=0.5, sig2_shape=0.0, sig2_scale=0.0, nu_shape=2.0, nu_rate=2.0,
                          alpha_a=1.0, alpha_b=1.0, sig2_true=0.0, tau_true=0.0, burn=500, ortho=False,
                          seed=1, chain=0, quiet=True):
        X = np.asfortranarray(X, dtype=np.float64); y = _f64(y)
        N, P = X.shape; M = nsamp
        beta, lam = np.zeros((M, P)), np.zeros((M, P))
        s2, tau, al = (np.zeros(M) for _ in range(3))
        d = lambda v: C.byref(C.c_double(v))
        i = lambda v: C.byref(C.c_int(int(v)))
        rt = C.c_double(0)
        self.L.bb_compat_set_seed(seed, chain)
        with _quiet(quiet):
            self.L.bridge_reg_stable(_p(beta), _p(lam), _p(s2), _p(tau), _p(al), _p(y), _p(X),
                                     d(sig2_shape), d(sig2_scale), d(nu_shape), d(nu_rate), d(alpha_a), d(alpha_b),
                                     d(sig2_true), d(tau_true), d(alpha), i(P), i(N), i(M), i(burn), C.byref(rt), i(ortho))
        return dict(beta=beta, **{"lambda": lam}, sig2=s2, tau=tau, alpha=al, runtime=rt.value)


class _quiet:
    """Silence the reference's printf banners (BridgeWrapper.cpp:112-117,162-163,200) at fd level."""

    def __init__(self, on):
        self.on = on

    def __enter__(self):
        if self.on:
            import sys
            sys.stdout.flush()
            self._saved = os.dup(1)
            self._null = os.open(os.devnull, os.O_WRONLY)
            os.dup2(self._null, 1)

    def __exit__(self, *a):
        if self.on:
            C.CDLL(None).fflush(None)
            os.dup2(self._saved, 1)
            os.close(self._saved); os.close(self._null)


class RefModel:
    def __init__(self, ref, X, y):
        self.L = ref.L
        self.X = np.asfortranarray(X, dtype=np.float64); self.y = _f64(y)
        self.N, self.P = self.X.shape
        self.h = self.L.ref_create(_p(self.X), _p(self.y), self.N, self.P)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_free(self.h); self.h = None

    def factors(self):
        P = self.P
        XX = np.zeros((P, P), order="F"); tV = np.zeros((P, P), order="F")
        Xy, bhat, d, a = (np.zeros(P) for _ in range(4))
        self.L.ref_get_factors(self.h, _p(XX), _p(Xy), _p(bhat), _p(tV), _p(d), _p(a))
        return dict(XX=XX, Xy=Xy, bhat=bhat, tV=tV, d=d, a=a)

    def sample_tau_marg(self, beta, alpha, nu_shape, nu_rate, seed, chain, it, stage=ST_TAU):
        return self.L.ref_sample_tau_marg(self.h, _p(_f64(beta)), alpha, nu_shape, nu_rate, seed, chain, it, stage)

    def sample_sig2(self, beta, sig2_shape, sig2_scale, seed, chain, it):
        return self.L.ref_sample_sig2(self.h, _p(_f64(beta)), sig2_shape, sig2_scale, seed, chain, it)

    def sample_omega(self, beta, u, tau, alpha, seed, chain, it):
        om, sh = np.zeros(self.P), np.zeros(self.P)
        self.L.ref_sample_omega(self.h, _p(om), _p(sh), _p(_f64(beta)), _p(_f64(u)), tau, alpha, seed, chain, it)
        return om, sh

    def sample_u(self, beta, omega, tau, alpha, seed, chain, it):
        u = np.zeros(self.P)
        self.L.ref_sample_u(self.h, _p(u), _p(_f64(beta)), _p(_f64(omega)), tau, alpha, seed, chain, it)
        return u

    def sample_beta(self, beta_prev, u, omega, sig2, tau, alpha, seed, chain, it, betaburn=0):
        out = np.zeros(self.P)
        self.L.ref_sample_beta(self.h, _p(out), _p(_f64(beta_prev)), _p(_f64(u)), _p(_f64(omega)), sig2, tau, alpha, betaburn, seed, chain, it)
        return out

    def rtnorm_gibbs(self, beta, a, tV, d, b, sig2, seed, chain, it):
        out = _f64(beta).copy(); P = out.size
        tV = np.asfortranarray(tV, dtype=np.float64)
        self.L.ref_rtnorm_gibbs(_p(out), _p(_f64(a)), _p(tV), _p(_f64(d)), _p(_f64(b)), sig2, P, seed, chain, it)
        return out

    def sample_beta_ortho(self, beta_prev, u, omega, sig2, tau, alpha, seed, chain, it):
        out = np.zeros(self.P)
        self.L.ref_sample_beta_ortho(self.h, _p(out), _p(_f64(beta_prev)), _p(_f64(u)), _p(_f64(omega)), sig2, tau, alpha, seed, chain, it)
        return out

    def sample_lambda(self, beta, alpha, tau, seed, chain, it):
        lam = np.zeros(self.P)
        self.L.ref_sample_lambda(self.h, _p(lam), _p(_f64(beta)), alpha, tau, seed, chain, it)
        return lam

    def sample_beta_stable(self, lam, sig2, tau, seed, chain, it, alpha=0.5):
        out = np.zeros(self.P)
        if self.L.ref_sample_beta_stable(self.h, _p(out), _p(_f64(lam)), alpha, sig2, tau, seed, chain, it):
            raise FloatingPointError("Cholesky failed")
        return out

    def sample_beta_stable_ortho(self, lam, sig2, tau, seed, chain, it, alpha=0.5):
        out = np.zeros(self.P)
        self.L.ref_sample_beta_stable_ortho(self.h, _p(out), _p(_f64(lam)), alpha, sig2, tau, seed, chain, it)
        return out

    def sample_alpha_marg(self, alpha_prev, beta, tau, pr_a, pr_b, seed, chain, it):
        return self.L.ref_sample_alpha_marg(self.h, alpha_prev, _p(_f64(beta)), tau, pr_a, pr_b, seed, chain, it)

    def retstable_LD(self, h, alpha, V0, seed, chain, it, item):
        return self.L.ref_retstable_LD(h, alpha, V0, seed, chain, it, item)


def draw_many(kind, n, params=(), seed=1):
    """bulk draws from the draw spec (bbo_draws.h) for distribution tests; see bbo_draw_many."""
    L = Port().L
    L.bbo_draw_many.argtypes = [C.c_int, C.c_int, _dp, C.c_uint64, _dp]
    p = _f64(list(params) + [0.0])
    out = np.zeros(n)
    L.bbo_draw_many(kind, n, _p(p), seed, _p(out))
    return out


SC_KINDS = {"rtexpon_rate_left": 0, "rtexpon_rate_both": 1, "rtexpon_rate": 2, "rtnorm_left": 3, "rtnorm_both": 4,
            "rtnorm": 5, "rrtgamma_rate": 6, "retstable_LD": 7}


def draw_vec(name, params, seed=1, it=0):
    """the stand-alone sampler entry points (BridgeWrapper.cpp:762-984) on per-item parameter arrays, in the
    argument order of the .C symbol `name`; see bbo_draw_vec."""
    L = Port().L
    L.bbo_draw_vec.argtypes = [C.c_int, C.c_int, _dp, _dp, _dp, _dp, C.c_uint64, C.c_uint32, _dp]
    ps = [_f64(p) for p in params]
    n = ps[0].size
    out = np.zeros(n)
    ps4 = ps + [None] * (4 - len(ps))
    L.bbo_draw_vec(SC_KINDS[name], n, *[None if p is None else _p(p) for p in ps4], seed, it, _p(out))
    return out


def philox(ctr, key):
    L = Port().L
    c = (C.c_uint32 * 4)(*ctr); k = (C.c_uint32 * 2)(*key); o = (C.c_uint32 * 4)()
    L.bbo_philox(c, k, o)
    return list(o)


def tnorm_at(seed, chain, it, stage, item, left, right, mu, sd):
    L = Port().L
    L.bbo_tnorm_at.restype = C.c_double
    L.bbo_tnorm_at.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32] + [C.c_double] * 4
    return L.bbo_tnorm_at(seed, chain, it, stage, item, left, right, mu, sd)


def flat_at(seed, chain, it, stage, item, a, b):
    L = Port().L
    L.bbo_flat_at.restype = C.c_double
    L.bbo_flat_at.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32] + [C.c_double] * 2
    return L.bbo_flat_at(seed, chain, it, stage, item, a, b)


def norm_at(seed, chain, it, stage, item):
    """the standard normal of Philox block 0 at (chain, iteration, stage, item): what Stream::norm(0) returns on the device"""
    L = Port().L
    L.bbo_norm_at.restype = C.c_double
    L.bbo_norm_at.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
    return L.bbo_norm_at(seed, chain, it, stage, item)
