"""the phases of one end-to-end driver call (bb_set_verbose(2)): python scripts/e2e_trace.py [workload]"""
import sys, time, ctypes as C
sys.path.insert(0, ".")
import numpy as np, torch
import bayesbridge_b200 as bb
from bayesbridge_b200 import synth, _lib
name = sys.argv[1] if len(sys.argv) > 1 else "c3"
X, y, _ = synth.make(name)
N, P = X.shape
nch, nsamp = {"c3": (512, 101), "c4": (1024, 5), "c2": (512, 201), "c1": (1, 20001)}.get(name, (256, 21))
method = "triangle" if name in ("c1", "c3", "c5") else "stable"
Xp = torch.from_numpy(np.asfortranarray(X).T.copy()).pin_memory(); yp = torch.from_numpy(y.copy()).pin_memory()
beta = torch.empty((nch, nsamp, P), dtype=torch.float64).pin_memory()
L = _lib.load()
from bayesbridge_b200._lib import BB_STB, BB_TRI, dp, up
hyp = bb.make_hyper(alpha=0.25 if name == "c4" else 0.5); rt = C.c_double(0); status = np.zeros(nch, dtype=np.uint32)
cast = lambda t: C.cast(t.data_ptr(), dp)
for s in range(3):
    if s == 2: L.bb_set_verbose(2)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    rc = L.bb_bridge_batched(BB_TRI if method == "triangle" else BB_STB, 0, cast(Xp), cast(yp), N, P, C.byref(hyp), nch, 7 + s, 0, 0, nsamp, 1, 0,
                             cast(beta), None, None, None, None, None, None, status.ctypes.data_as(up), C.byref(rt))
    torch.cuda.synchronize(); print("call", s, "rc", rc, "wall %.1f ms, runtime %.1f ms" % (1e3 * (time.perf_counter() - t0), 1e3 * rt.value), flush=True)
