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


def _timed(fn, device, timing, key):
    """run fn() between two CUDA events on torch's current stream; add the elapsed ms to timing[key]"""
    if timing is None:
        return fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = fn()
    e1.record()
    torch.cuda.synchronize(device)
    timing[key] = timing.get(key, 0.0) + e0.elapsed_time(e1)
    return out


def replicate(X, y, N, P, device, src=0, timing=None):
    """One broadcast of the packed column-major [X | y] from `src`; returns device tensors (Xd, yd).
    X, y are numpy arrays (or a pinned 1-D torch tensor holding the packed [X | y]) on `src` and may be None elsewhere.
    timing (dict, optional) receives h2d_ms and comm_ms (the NCCL broadcast), timed with CUDA events."""
    buf = torch.empty((P + 1) * N, dtype=torch.float64, device=device)
    if not dist.is_initialized() or dist.get_rank() == src:
        if torch.is_tensor(X):
            host = X
        else:
            host = torch.from_numpy(np.concatenate([np.asfortranarray(X, dtype=np.float64).ravel(order="F"), np.asarray(y, dtype=np.float64)]))
        _timed(lambda: buf.copy_(host, non_blocking=True), device, timing, "h2d_ms")
    if dist.is_initialized() and dist.get_world_size() > 1:
        _timed(lambda: dist.broadcast(buf, src=src), device, timing, "comm_ms")
        if timing is not None:
            timing["collective"] = f"1 ncclBroadcast of [X|y] ({8 * (P + 1) * N / 1e6:.1f} MB) from rank {src}"
    return buf[: N * P], buf[N * P:], buf


def replicated_model(X, y, N, P, local_device, src=0, timing=None):
    """Model on this rank's GPU from data broadcast once over NCCL.  Keeps the device buffer alive."""
    dev = torch.device("cuda", local_device)
    Xd, yd, buf = replicate(X, y, N, P, dev, src, timing)
    torch.cuda.synchronize(dev)
    m = Model.from_device(Xd.data_ptr(), yd.data_ptr(), N, P, local_device)
    m._keep = buf
    return m


def sharded_model(X_rows, y_rows, N_total, P, local_device, timing=None):
    """Row-sharded construction: X_rows (n_local x P) and y_rows are this rank's rows (numpy, or a pinned 1-D torch tensor
    holding the packed column-major [X_rows | y_rows] with n_local = len / (P + 1)).
    timing (dict, optional) receives h2d_ms and comm_ms (the two NCCL all-reduces), timed with CUDA events."""
    L = _lib.load()
    dev = torch.device("cuda", local_device)
    if torch.is_tensor(X_rows):
        host = X_rows
        n_local = host.numel() // (P + 1)
    else:
        n_local = X_rows.shape[0]
        host = torch.from_numpy(np.concatenate([np.asfortranarray(X_rows, dtype=np.float64).ravel(order="F"), np.asarray(y_rows, dtype=np.float64)]))
    buf = torch.empty((P + 1) * n_local, dtype=torch.float64, device=dev)
    _timed(lambda: buf.copy_(host, non_blocking=True), dev, timing, "h2d_ms")
    Xd, yd = buf[: n_local * P], buf[n_local * P:]
    G = torch.zeros((P + 1) * (P + 1), dtype=torch.float64, device=dev)
    torch.cuda.synchronize(dev)
    if L.bb_gram(Xd.data_ptr(), yd.data_ptr(), n_local, P, G.data_ptr(), local_device):
        raise RuntimeError(_lib.last_error())
    multi = dist.is_initialized() and dist.get_world_size() > 1
    if multi:
        _timed(lambda: dist.all_reduce(G), dev, timing, "comm_ms")          # the one all-reduce of the packed Gram matrix
    torch.cuda.synchronize(dev)
    m = Model.from_gram(G.data_ptr(), N_total, P, local_device)
    r = torch.zeros(2, dtype=torch.float64, device=dev)
    if L.bb_model_resid_partial(m.h, Xd.data_ptr(), yd.data_ptr(), n_local, r.data_ptr()):
        raise RuntimeError(_lib.last_error())
    if multi:
        _timed(lambda: dist.all_reduce(r), dev, timing, "comm_ms")
        if timing is not None:
            timing["collective"] = f"ncclAllReduce of the packed Gram [X'X X'y; y'X y'y] ({8 * (P + 1) ** 2 / 1e6:.2f} MB) + 16-byte all-reduce of the residual constants"
    r = r.cpu().numpy()
    L.bb_model_set_r0(m.h, float(r[0]), float(r[1]))
    m._keep = (buf, G)
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
