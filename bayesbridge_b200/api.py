"""Host-side mirror of the reference's R front end and C++ sampler interface.

R entry points (Code/C/BridgeWrapper.R in the reference):
    bridge.reg.tri  :139-188  -> bridge_reg_tri
    bridge.reg.stb  :194-234  -> bridge_reg_stb
    bridge.reg      :240-276  -> bridge_reg
Same argument names (dots -> underscores), same defaults, same parameter checks
(check.parameters :50-69), same return list.  They call the drop-in C symbols
`bridge_regression` / `bridge_reg_stable` of libbayesbridge_b200.so exactly as .C() would.

`Model` / `Chains` expose class BridgeRegression (Code/C/BridgeRegression.h:81-166): the
constructor's factors live in HBM, the conditionals are stage-masked kernel launches over all
chains.  `bridge_batched` is the many-chain extension.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import (BB_STB, BB_TRI, Hyper, OUT_ALPHA, OUT_BETA, OUT_LAMBDA, OUT_OMEGA, OUT_SHAPE, OUT_SIG2, OUT_TAU, OUT_U,
                   ST_ALL, ST_BETA, ST_LAMBDA, ST_OMEGA, ST_SIG2, ST_TAU, ST_U, dp, up)


def _p(a):
    return None if a is None else a.ctypes.data_as(dp)


def _f(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _check_parameters(N, R, M, sig2_shape, sig2_scale, nu_shape, nu_rate, alpha_a, alpha_b):
    """check.parameters, BridgeWrapper.R:50-69 (prints and refuses, like the reference)."""
    ok = True
    if N != R:
        print("Error: y and X do not conform.")
        ok = False
    for val, lo, name in ((M, 1, "niter"), (sig2_shape, 0, "sig2.shape"), (sig2_scale, 0, "sig2.scale"),
                          (nu_shape, 0, "nu.shape"), (nu_rate, 0, "nu.rate"), (alpha_a, -1, "alpha.a"), (alpha_b, -1, "alpha.b")):
        if not isinstance(val, (int, float, np.integer, np.floating)):
            print(f"Error: {name} is not numeric.")
            ok = False
        elif val < lo:
            print(f"Error: {name}<{lo}")
            ok = False
    return ok


def _d(v):
    return C.byref(C.c_double(float(v)))


def _i(v):
    return C.byref(C.c_int(int(v)))


def bridge_reg_tri(y, X, nsamp, alpha=0.5, sig2_shape=0.0, sig2_scale=0.0, nu_shape=2.0, nu_rate=2.0, alpha_a=1.0, alpha_b=1.0,
                   sig2_true=0.0, tau_true=0.0, burn=500, ortho=False, betaburn=0, extras=False, seed=None, device=None):
    """bridge.reg.tri (BridgeWrapper.R:139-188): mixture-of-triangles sampler, one chain."""
    L = _lib.load()
    X = np.asfortranarray(X, dtype=np.float64)
    y = _f(y)
    N, (R, P), M = y.size, X.shape, int(nsamp)
    if not _check_parameters(N, R, M, sig2_shape, sig2_scale, nu_shape, nu_rate, alpha_a, alpha_b):
        raise ValueError("bridge.reg.tri: invalid parameters")
    if seed is not None:
        L.bb_set_seed(int(seed))
    if device is not None:
        L.bb_set_device(int(device))
    beta, u, omega, shape = (np.zeros((M, P)) for _ in range(4))
    sig2, tau, al = (np.zeros(M) for _ in range(3))
    rtime = C.c_double(0.0)
    L.bridge_regression(_p(beta), _p(u), _p(omega), _p(shape), _p(sig2), _p(tau), _p(al), _p(y), _p(X),
                        _d(sig2_shape), _d(sig2_scale), _d(nu_shape), _d(nu_rate), _d(alpha_a), _d(alpha_b),
                        _d(sig2_true), _d(tau_true), _d(alpha), _i(P), _i(N), _i(M), _i(burn), C.byref(rtime),
                        _i(bool(ortho)), _i(betaburn), _i(0))
    return {"beta": beta, "u": u, "w": omega, "shape": shape, "sig2": sig2, "tau": tau, "alpha": al, "runtime": rtime.value}


def bridge_reg_stb(y, X, nsamp, alpha=0.5, sig2_shape=0.0, sig2_scale=0.0, nu_shape=2.0, nu_rate=2.0, alpha_a=1.0, alpha_b=1.0,
                   sig2_true=0.0, tau_true=0.0, burn=500, ortho=False, seed=None, device=None):
    """bridge.reg.stb (BridgeWrapper.R:194-234): normal-scale-mixture (stable) sampler, one chain."""
    L = _lib.load()
    X = np.asfortranarray(X, dtype=np.float64)
    y = _f(y)
    N, (R, P), M = y.size, X.shape, int(nsamp)
    if not _check_parameters(N, R, M, sig2_shape, sig2_scale, nu_shape, nu_rate, alpha_a, alpha_b):
        raise ValueError("bridge.reg.stb: invalid parameters")
    if seed is not None:
        L.bb_set_seed(int(seed))
    if device is not None:
        L.bb_set_device(int(device))
    beta, lam = np.zeros((M, P)), np.zeros((M, P))
    sig2, tau, al = (np.zeros(M) for _ in range(3))
    rt = C.c_double(0.0)
    L.bridge_reg_stable(_p(beta), _p(lam), _p(sig2), _p(tau), _p(al), _p(y), _p(X),
                        _d(sig2_shape), _d(sig2_scale), _d(nu_shape), _d(nu_rate), _d(alpha_a), _d(alpha_b),
                        _d(sig2_true), _d(tau_true), _d(alpha), _i(P), _i(N), _i(M), _i(burn), C.byref(rt), _i(bool(ortho)))
    return {"beta": beta, "lambda": lam, "sig2": sig2, "tau": tau, "alpha": al, "runtime": rt.value}


def bridge_reg(y, X, nsamp, alpha=0.5, sig2_shape=0.0, sig2_scale=0.0, nu_shape=2.0, nu_rate=2.0, alpha_a=1.0, alpha_b=1.0,
               sig2_true=0.0, tau_true=0.0, burn=500, method="triangle", ortho=False, seed=None, device=None):
    """bridge.reg (BridgeWrapper.R:240-276).  Like the reference it IGNORES its prior arguments and
    hard-codes alpha=0.5, nu.shape=nu.rate=0.5, burn=500 (BridgeWrapper.R:252-259,262-269)."""
    if method == "triangle":
        return bridge_reg_tri(y, X, nsamp, alpha=0.5, nu_shape=0.5, nu_rate=0.5, burn=500, ortho=ortho, seed=seed, device=device)
    if method == "stable":
        return bridge_reg_stb(y, X, nsamp, alpha=0.5, nu_shape=0.5, nu_rate=0.5, burn=500, ortho=ortho, seed=seed, device=device)
    print('Unrecognized method.  Use "triangles" or "stable".')
    return None


def make_hyper(alpha=0.5, sig2_shape=0.0, sig2_scale=0.0, nu_shape=2.0, nu_rate=2.0, alpha_a=1.0, alpha_b=1.0, sig2_true=0.0,
               tau_true=0.0, betaburn=0):
    return Hyper(sig2_shape, sig2_scale, nu_shape, nu_rate, alpha_a, alpha_b, sig2_true, tau_true, alpha, betaburn)


def bridge_batched(y, X, nsamp, method="triangle", nchains=1, seed=20121027, chain0=0, burn=500, thin=1, ortho=False,
                   device=0, want=("beta", "sig2", "tau"), **hyper):
    """Many chains, the drivers' schedule per chain (bb_bridge_batched).  Arrays are [chain, sample, P]."""
    L = _lib.load()
    X = np.asfortranarray(X, dtype=np.float64)
    y = _f(y)
    N, P = X.shape
    meth = BB_TRI if method in ("triangle", "tri") else BB_STB
    h = make_hyper(**hyper)
    out = {}
    def vec(name):
        if name in want:
            out[name] = np.zeros((nchains, nsamp, P))
            return _p(out[name])
        return None
    def sca(name):
        if name in want:
            out[name] = np.zeros((nchains, nsamp))
            return _p(out[name])
        return None
    b = vec("beta")
    if meth == BB_TRI:
        a1, a2, a3 = vec("u"), vec("w"), vec("shape")
    else:
        a1, a2, a3 = vec("lambda"), None, None
    s2, ta, al = sca("sig2"), sca("tau"), sca("alpha")
    status = np.zeros(nchains, dtype=np.uint32)
    rt = C.c_double(0.0)
    rc = L.bb_bridge_batched(meth, int(bool(ortho)), _p(X), _p(y), N, P, C.byref(h), nchains, seed, chain0, burn, nsamp, thin,
                             device, b, a1, a2, a3, s2, ta, al, status.ctypes.data_as(up), C.byref(rt))
    if rc:
        raise RuntimeError("bb_bridge_batched: " + _lib.last_error())
    out["status"] = status
    out["runtime"] = rt.value
    return out


def bridge_batched_summary(y, X, nsamp, method="triangle", nchains=1, seed=20121027, chain0=0, burn=500, thin=1, ortho=False,
                           device=0, per_chain=False, **hyper):
    """bridge_batched in summary mode (bb_bridge_batched_summary): the draws stay on the device, the result is the table
    of sum.stat (Code/R/PublicBenchmark.R:112-134) -- mean, sd, ESS, ESS.sec, t, Q.1, Q.9 per coefficient over all chains."""
    L = _lib.load()
    X = np.asfortranarray(X, dtype=np.float64)
    y = _f(y)
    N, P = X.shape
    meth = BB_TRI if method in ("triangle", "tri") else BB_STB
    h = make_hyper(**hyper)
    pooled = np.zeros((7, P))
    pc = np.zeros((nchains, 5, P)) if per_chain else None
    sig2, tau = np.zeros((nchains, nsamp)), np.zeros((nchains, nsamp))
    status = np.zeros(nchains, dtype=np.uint32)
    rt = C.c_double(0.0)
    rc = L.bb_bridge_batched_summary(meth, int(bool(ortho)), _p(X), _p(y), N, P, C.byref(h), nchains, seed, chain0, burn, nsamp,
                                     thin, device, _p(pooled), _p(pc), _p(sig2), _p(tau), status.ctypes.data_as(up), C.byref(rt))
    if rc:
        raise RuntimeError("bb_bridge_batched_summary: " + _lib.last_error())
    out = dict(zip(("mean", "sd", "ESS", "ESS.sec", "t", "Q.1", "Q.9"), pooled))
    out.update(sig2=sig2, tau=tau, status=status, runtime=rt.value)
    if per_chain:
        out["chains"] = pc
    return out


class Model:
    """BridgeRegression::BridgeRegression(X, y) (BridgeRegression.cpp:13-69) on the device."""

    def __init__(self, X, y, device=0, factors=None, _handle=None, _shape=None):
        self.L = _lib.load()
        self.device = device
        if _handle is not None:
            self.h, (self.N, self.P) = _handle, _shape
            return
        Xf = np.asfortranarray(X, dtype=np.float64)
        yf = _f(y)
        self.N, self.P = Xf.shape
        if factors is None:
            self.h = self.L.bb_model_create(Xf.ctypes.data, yf.ctypes.data, self.N, self.P, device, _lib.MEM_HOST)
        else:
            tV = np.asfortranarray(factors["tV"], dtype=np.float64)
            d, a = _f(factors["d"]), _f(factors["a"])
            bhat = None if factors.get("bhat") is None else _f(factors["bhat"])
            self.h = self.L.bb_model_create_factors(Xf.ctypes.data, yf.ctypes.data, self.N, self.P, _p(tV), _p(d), _p(a), _p(bhat),
                                                    device, _lib.MEM_HOST)
        if not self.h:
            raise RuntimeError("bb_model_create: " + _lib.last_error())

    @classmethod
    def from_device(cls, X_ptr, y_ptr, N, P, device=0):
        """X, y already resident in HBM (raw device pointers, column-major N x P)."""
        L = _lib.load()
        h = L.bb_model_create(X_ptr, y_ptr, N, P, device, _lib.MEM_DEVICE)
        if not h:
            raise RuntimeError("bb_model_create: " + _lib.last_error())
        return cls(None, None, device, _handle=h, _shape=(N, P))

    @classmethod
    def from_gram(cls, G_ptr, N, P, device=0):
        L = _lib.load()
        h = L.bb_model_create_gram(G_ptr, N, P, device)
        if not h:
            raise RuntimeError("bb_model_create_gram: " + _lib.last_error())
        return cls(None, None, device, _handle=h, _shape=(N, P))

    def factors(self):
        P = self.P
        XX = np.zeros((P, P), order="F"); tV = np.zeros((P, P), order="F")
        Xy, bhat, d, a = (np.zeros(P) for _ in range(4))
        if self.L.bb_model_get(self.h, _p(XX), _p(Xy), _p(bhat), _p(tV), _p(d), _p(a)):
            raise RuntimeError(_lib.last_error())
        return dict(XX=XX, Xy=Xy, bhat=bhat, tV=tV, d=d, a=a)

    @property
    def precompute_ms(self):
        return self.L.bb_model_precompute_ms(self.h)

    def chains(self, method="triangle", nchains=1, seed=20121027, chain0=0, ortho=False):
        return Chains(self, method, nchains, seed, chain0, ortho)

    def close(self):
        if getattr(self, "h", None):
            self.L.bb_model_free(self.h)
            self.h = None

    def __del__(self):
        self.close()


_OUT = {"beta": OUT_BETA, "u": OUT_U, "w": OUT_OMEGA, "omega": OUT_OMEGA, "shape": OUT_SHAPE, "lambda": OUT_LAMBDA,
        "sig2": OUT_SIG2, "tau": OUT_TAU, "alpha": OUT_ALPHA}


class Chains:
    """The per-chain slices the drivers carry (BridgeWrapper.cpp:80-100) for `nchains` chains, in HBM."""

    def __init__(self, model, method, nchains, seed, chain0, ortho):
        self.model, self.L = model, model.L
        self.method = BB_TRI if method in ("triangle", "tri", BB_TRI) else BB_STB
        self.nchains, self.P = nchains, model.P
        self.h = self.L.bb_chains_create(model.h, self.method, int(bool(ortho)), nchains, seed, chain0)
        if not self.h:
            raise RuntimeError("bb_chains_create: " + _lib.last_error())
        self.nslots = 0

    def _ck(self, rc):
        if rc:
            raise RuntimeError(_lib.last_error())

    def wide_gaussian(self, on=True):
        """opt into the N x N form of the Gaussian draw of beta (stable sampler, P > N; bb_chains_option): same law as
        sample_beta_stable (BridgeRegression.cpp:552-575), not the same draws"""
        self._ck(self.L.bb_chains_option(self.h, _lib.OPT_WIDE_GAUSSIAN, int(bool(on))))
        return self

    def set(self, beta=None, u=None, aux=None, sig2=None, tau=None, alpha=None):
        arrs = [None if a is None else _f(a) for a in (beta, u, aux, sig2, tau)]
        self._ck(self.L.bb_chains_set(self.h, *[_p(a) for a in arrs]))
        if alpha is not None:
            self._ck(self.L.bb_chains_alpha(self.h, _p(_f(np.broadcast_to(alpha, (self.nchains,)))), None))

    def get(self):
        n, P = self.nchains, self.P
        beta, u, aux, shape = (np.zeros((n, P)) for _ in range(4))
        sig2, tau = np.zeros(n), np.zeros(n)
        status = np.zeros(n, dtype=np.uint32)
        self._ck(self.L.bb_chains_get(self.h, _p(beta), _p(u), _p(aux), _p(shape), _p(sig2), _p(tau), status.ctypes.data_as(up)))
        alpha = np.zeros(n)
        self._ck(self.L.bb_chains_alpha(self.h, None, _p(alpha)))
        return dict(beta=beta, u=u, aux=aux, shape=shape, sig2=sig2, tau=tau, alpha=alpha, status=status)

    def reserve(self, nslots, outputs=("beta",)):
        mask = 0
        for o in outputs:
            mask |= _OUT[o]
        self._ck(self.L.bb_chains_reserve(self.h, nslots, mask))
        self.nslots = nslots

    def run(self, hyper, niter, iter0=0, stages=ST_ALL, pre_tau=0, thin=0, slot0=0):
        """Advance all chains; returns the device time of the launches in ms."""
        ms = C.c_double(0.0)
        self._ck(self.L.bb_chains_run(self.h, C.byref(hyper), iter0, niter, stages, pre_tau, thin, slot0, C.byref(ms)))
        return ms.value

    def store_state(self, slot):
        self._ck(self.L.bb_chains_store_state(self.h, slot))

    def fetch(self, name, slot0=0, nslots=None, out=None):
        """stored draws [chain, slot, P] (or [chain, slot]); `out`: a host buffer to fill instead of a new numpy array --
        anything with .ctypes.data or .data_ptr() (e.g. a pinned torch tensor: the copy then runs at full PCIe speed)."""
        nslots = self.nslots - slot0 if nslots is None else nslots
        w = 1 if name in ("sig2", "tau", "alpha") else self.P
        if out is None:
            out = np.zeros((self.nchains, nslots, w)) if w > 1 else np.zeros((self.nchains, nslots))
        ptr = out.data_ptr() if hasattr(out, "data_ptr") else out.ctypes.data
        self._ck(self.L.bb_chains_fetch(self.h, _OUT[name], slot0, nslots, ptr))
        return out

    def summary(self, name="beta", slot0=0, nslots=None, runtime=None, per_chain=False):
        """sum.stat (Code/R/PublicBenchmark.R:112-134) of the stored draws, computed on the device: dict of
        mean, sd, ESS (coda definition, summed over chains), ESS.sec, t, Q.1, Q.9 per coefficient over all chains
        pooled; with per_chain=True also "chains": [nchains, 5, P] = mean, sd, ESS, Q.1, Q.9 of each chain."""
        nslots = self.nslots - slot0 if nslots is None else nslots
        pooled = np.zeros((7, self.P))
        pc = np.zeros((self.nchains, 5, self.P)) if per_chain else None
        self._ck(self.L.bb_chains_summary(self.h, _OUT[name], slot0, nslots, float(runtime or 0.0),
                                          pc.ctypes.data if per_chain else None, pooled.ctypes.data))
        out = dict(zip(("mean", "sd", "ESS", "ESS.sec", "t", "Q.1", "Q.9"), pooled))
        if runtime is None:
            del out["ESS.sec"]
        if per_chain:
            out["chains"] = pc
        return out

    def running_begin(self, name="beta"):
        """RUNNING sum.stat (bb_chains_summary_begin / add / end): fold chunks of stored draws into per-series accumulators,
        reuse the draw buffer, read the table at the end -- no limit on the run length."""
        self._ck(self.L.bb_chains_summary_begin(self.h, _OUT[name]))

    def running_add(self, slot0, nslots):
        self._ck(self.L.bb_chains_summary_add(self.h, slot0, nslots))

    def running_end(self, runtime=None, per_chain=False):
        pooled = np.zeros((7, self.P))
        pc = np.zeros((self.nchains, 5, self.P)) if per_chain else None
        self._ck(self.L.bb_chains_summary_end(self.h, float(runtime or 0.0), pc.ctypes.data if per_chain else None, pooled.ctypes.data))
        out = dict(zip(("mean", "sd", "ESS", "ESS.sec", "t", "Q.1", "Q.9"), pooled))
        if runtime is None:
            del out["ESS.sec"]
        if per_chain:
            out["chains"] = pc
        return out

    # ---- the conditionals of class BridgeRegression, one launch each over all chains ----
    def sample_tau_marg(self, hyper, it):  # BridgeRegression.cpp:453-465
        return self.run(hyper, 1, it, ST_TAU)

    def sample_sig2(self, hyper, it):      # :436-450
        return self.run(hyper, 1, it, ST_SIG2)

    def sample_omega(self, hyper, it):     # :130-147
        return self.run(hyper, 1, it, ST_OMEGA)

    def sample_u(self, hyper, it):         # :97-110
        return self.run(hyper, 1, it, ST_U)

    def sample_beta(self, hyper, it):      # :405-433 (tri) / :552-575 (stb) / ortho forms
        return self.run(hyper, 1, it, ST_BETA)

    def sample_lambda(self, hyper, it):    # :506-510
        return self.run(hyper, 1, it, ST_LAMBDA)

    def close(self):
        if getattr(self, "h", None):
            self.L.bb_chains_free(self.h)
            self.h = None

    def __del__(self):
        self.close()
