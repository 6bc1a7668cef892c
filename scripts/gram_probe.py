"""Times bb_gram (the one-time [X y]'[X y] build on the FP64 tensor cores) at the c3 / c5 / c4 shapes: CUDA events around repeated
calls, result checked against numpy.  BB_GRAM_LDG=1 selects the LDG->STS kernel, default is the TMA-staged one."""
import os, sys, time
sys.path[:0] = ['.']
import numpy as np, torch
import bayesbridge_b200 as bb
L = bb._lib.load()
dev = torch.device("cuda", 0)
for N, P in ((10000, 1000), (100000, 256), (500, 2000), (442, 64)):
    rng = np.random.default_rng(N)
    X = rng.standard_normal((N, P)); y = rng.standard_normal(N)
    Xd = torch.from_numpy(np.asfortranarray(X).ravel(order="F").copy()).to(dev); yd = torch.from_numpy(y).to(dev)
    G = torch.zeros((P + 1) * (P + 1), dtype=torch.float64, device=dev)
    for _ in range(3):
        assert L.bb_gram(Xd.data_ptr(), yd.data_ptr(), N, P, G.data_ptr(), 0) == 0, bb._lib.last_error()
    torch.cuda.synchronize()
    reps = 20
    t0 = time.perf_counter()
    for _ in range(reps):
        L.bb_gram(Xd.data_ptr(), yd.data_ptr(), N, P, G.data_ptr(), 0)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / reps * 1e3
    A = np.column_stack([X, y]); want = A.T @ A
    got = G.cpu().numpy().reshape(P + 1, P + 1)
    err = np.abs(got - want).max() / np.abs(want).max()
    flops = float(N) * (P + 1) * (P + 2)
    print(f"gram {N}x{P}: {ms:.3f} ms per call (host clock, incl. launch + sync)  {flops / ms / 1e9:.2f} TFLOP/s (N (P+1)(P+2) flop)  {8.0 * N * (P + 1) / ms / 1e6:.0f} GB/s of X  err {err:.1e}  kernel {'LDG' if os.environ.get('BB_GRAM_LDG') else 'TMA'}")
