"""CPU, world_size 2 over gloo: the host-side logic of the multi-GPU path (chain sharding, the packed
[X|y] broadcast, the packed Gram all-reduce).  The device kernels are not involved."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    from bayesbridge_b200 import dist as bdist
    r, w, _ = bdist.init("gloo")
    assert (r, w) == (rank, world)
    # chain sharding: contiguous, disjoint, covering
    lo, hi = bdist.chain_range(4097, r, w)
    # broadcast of the packed design
    N, P = 50, 7
    rng = np.random.default_rng(0)
    X = np.asfortranarray(rng.standard_normal((N, P))); y = rng.standard_normal(N)
    Xd, yd, _ = bdist.replicate(X if r == 0 else None, y if r == 0 else None, N, P, torch.device("cpu"))
    ok_bcast = np.allclose(Xd.numpy().reshape(P, N).T, X) and np.allclose(yd.numpy(), y)
    # row-sharded Gram: sum of shard Grams == Gram of the whole (the all-reduce the GPU path performs)
    a, b = bdist.row_range(N, r, w)
    Z = np.column_stack([X[a:b], y[a:b]])
    G = torch.from_numpy(Z.T @ Z)
    dist.all_reduce(G)
    Zf = np.column_stack([X, y])
    ok_gram = np.allclose(G.numpy(), Zf.T @ Zf)
    t = bdist.max_over_ranks(1.0 + r)
    # per-coefficient ESS of sharded chains adds up over the ranks (bench.py's ESS/sec)
    e = bdist.sum_over_ranks(np.arange(5, dtype=np.float64) * (1 + r))
    ok_gram = ok_gram and np.allclose(e, np.arange(5) * 3.0)
    bdist.barrier()
    q.put((rank, lo, hi, ok_bcast, ok_gram, t))
    dist.destroy_process_group()


def test_two_rank_gloo():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0][1] == 0 and res[0][2] == res[1][1] and res[1][2] == 4097
    assert abs((res[0][2] - res[0][1]) - (res[1][2] - res[1][1])) <= 1
    assert all(r[3] and r[4] for r in res)
    assert all(r[5] == 2.0 for r in res)


def test_chain_range_partitions():
    sys.path.insert(0, ROOT)
    from bayesbridge_b200.dist import chain_range
    for total in (1, 7, 512, 4096, 4097):
        for world in (1, 2, 3, 8):
            spans = [chain_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
