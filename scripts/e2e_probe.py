import sys, time, ctypes as C, numpy as np
sys.path.insert(0,'.')
import bayesbridge_b200 as bb
from bayesbridge_b200 import synth
from bayesbridge_b200._lib import BB_STB, BB_TRI, dp, up
import torch
L=bb._lib.load(); L.bb_set_verbose(2)
WL=sys.argv[1] if len(sys.argv)>1 else "c2"; nch=int(sys.argv[2]) if len(sys.argv)>2 else 512; ips=int(sys.argv[3]) if len(sys.argv)>3 else 200
X,y,_=synth.make(WL); N,P=X.shape; nsamp=ips+1; METH=BB_TRI if synth.CONFIGS[WL]["method"]=="triangle" else BB_STB; ALPHA=synth.CONFIGS[WL]["alpha"]
Xp=torch.from_numpy(np.asfortranarray(X).ravel(order="F").copy()).pin_memory(); yp=torch.from_numpy(y.copy()).pin_memory()
beta=torch.empty((nch,nsamp,P),dtype=torch.float64).pin_memory(); sig2=torch.empty((nch,nsamp),dtype=torch.float64).pin_memory(); tau=torch.empty((nch,nsamp),dtype=torch.float64).pin_memory()
status=np.zeros(nch,dtype=np.uint32); hyp=bb.make_hyper(alpha=ALPHA); rt=C.c_double(0)
cast=lambda t: C.cast(t.data_ptr(), dp)
for s in range(3):
    t0=time.perf_counter()
    rc=L.bb_bridge_batched(METH,0,cast(Xp),cast(yp),N,P,C.byref(hyp),nch,123+s,0,0,nsamp,1,0,cast(beta),None,None,None,cast(sig2),cast(tau),None,status.ctypes.data_as(up),C.byref(rt))
    print("call", s, "wall ms", 1e3*(time.perf_counter()-t0), "rc", rc, flush=True)
