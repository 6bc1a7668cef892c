"""Multi-GPU plumbing: one process per GPU, chains sharded, X replicated.

The Gibbs sweep needs no communication: chain c of the job lives on rank c // chains_per_rank and
its Philox stream depends only on (seed, global chain id), so results do not depend on the number
of GPUs.  Two collectives exist, both outside the sampling loop:
  * replicate():      one NCCL broadcast of [X | y] from rank 0 over NVLink;
  * sharded_model():  for very tall X, every rank builds the Gram matrix of its row shard on the
                      FP64 tensor cores, then ONE all-reduce of the packed (P+1)x(P+1) matrix
                      [X'X X'y; y'X y'y] (plus a 2-element all-reduce of the residual sums).
torch is used for device memory and torch.distributed only.
"""
import os

import numpy as np
import torch
import torch.distributed as dist

from . import _lib
from .api import Model


def init(backend=None):
    """Join the torchrun job (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* from the environment).
    Returns (rank, world, device index)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    use_cuda = torch.cuda.is_available()
    if use_cuda:
        torch.cuda.set_device(local)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group(backend or ("nccl" if use_cuda else "gloo"), rank=rank, world_size=world)
    return rank, world, local


def chain_range(total_chains, rank, world):
    """Contiguous block of global chain ids owned by `rank` (sizes differ by at most one)."""
    base, extra = divmod(total_chains, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def row_range(N, rank, world):
    return chain_range(N, rank, world)


def replicate(X, y, N, P, device, src=0):
    """One broadcast of the packed column-major [X | y] from `src`; returns device tensors (Xd, yd).
    X, y are numpy arrays on `src` and may be None elsewhere."""
    buf = torch.empty((P + 1) * N, dtype=torch.float64, device=device)
    if not dist.is_initialized() or dist.get_rank() == src:
        packed = np.concatenate([np.asfortranarray(X, dtype=np.float64).ravel(order="F"), np.asarray(y, dtype=np.float64)])
        buf.copy_(torch.from_numpy(packed))
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.broadcast(buf, src=src)
    return buf[: N * P], buf[N * P:], buf


def replicated_model(X, y, N, P, local_device, src=0):
    """Model on this rank's GPU from data broadcast once over NCCL.  Keeps the device buffer alive."""
    dev = torch.device("cuda", local_device)
    Xd, yd, buf = replicate(X, y, N, P, dev, src)
    torch.cuda.synchronize(dev)
    m = Model.from_device(Xd.data_ptr(), yd.data_ptr(), N, P, local_device)
    m._keep = buf
    return m


def sharded_model(X_rows, y_rows, N_total, P, local_device):
    """Row-sharded construction: X_rows (n_local x P) and y_rows are this rank's rows (numpy)."""
    L = _lib.load()
    dev = torch.device("cuda", local_device)
    n_local = X_rows.shape[0]
    Xd = torch.from_numpy(np.asfortranarray(X_rows, dtype=np.float64).ravel(order="F").copy()).to(dev)
    yd = torch.from_numpy(np.asarray(y_rows, dtype=np.float64).copy()).to(dev)
    G = torch.zeros((P + 1) * (P + 1), dtype=torch.float64, device=dev)
    torch.cuda.synchronize(dev)
    if L.bb_gram(Xd.data_ptr(), yd.data_ptr(), n_local, P, G.data_ptr(), local_device):
        raise RuntimeError(_lib.last_error())
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(G)          # the one all-reduce of the packed Gram matrix
    torch.cuda.synchronize(dev)
    m = Model.from_gram(G.data_ptr(), N_total, P, local_device)
    r = torch.zeros(2, dtype=torch.float64, device=dev)
    if L.bb_model_resid_partial(m.h, Xd.data_ptr(), yd.data_ptr(), n_local, r.data_ptr()):
        raise RuntimeError(_lib.last_error())
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(r)
    r = r.cpu().numpy()
    L.bb_model_set_r0(m.h, float(r[0]), float(r[1]))
    m._keep = (Xd, yd, G)
    return m


def max_over_ranks(x, device=None):
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return float(x)
    t = torch.tensor([float(x)], dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(v, device=None):
    """element-wise sum of a float64 numpy vector over the ranks (the per-coefficient ESS of sharded chains adds up)"""
    import numpy as np
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return np.asarray(v, dtype=np.float64)
    t = torch.as_tensor(np.asarray(v, dtype=np.float64), device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t.cpu().numpy()


def barrier():
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()


def shutdown():
    """tear the process group down (no-op for a single process)"""
    import torch.distributed as td
    if td.is_available() and td.is_initialized():
        td.destroy_process_group()
