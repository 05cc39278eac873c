"""Multi-GPU plumbing: one process per GPU (torch.distributed, NCCL over NVLink;
gloo in the CPU tests), chains sharded in contiguous global-id ranges, no
data-path collective inside a transition.  Collectives carry only small
cross-chain reductions: diagnostic moment sums and the dual-averaging accept
statistic.  Replaces the reference's mpi4py gather/bcast launcher
(scripts_old/funnel/funnelmpi.py:140-205)."""
import os

import torch
import torch.distributed as td


def is_dist():
    return td.is_available() and td.is_initialized()


def rank():
    return td.get_rank() if is_dist() else 0


def world_size():
    return td.get_world_size() if is_dist() else 1


def init_from_env(backend=None):
    """Initialise the default process group from torchrun's environment, if any.
    Returns (rank, world_size, local_rank)."""
    ws = int(os.environ.get("WORLD_SIZE", "1"))
    if ws > 1 and not is_dist():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        lr = int(os.environ.get("LOCAL_RANK", "0"))
        if backend == "nccl":
            torch.cuda.set_device(lr)
            td.init_process_group(backend=backend, device_id=torch.device("cuda", lr))
        else:
            td.init_process_group(backend=backend)
    return rank(), world_size(), int(os.environ.get("LOCAL_RANK", "0"))


def shard_range(n_total, r=None, R=None):
    """Contiguous global chain-id range [lo, hi) of rank r of R."""
    r = rank() if r is None else r
    R = world_size() if R is None else R
    base, rem = divmod(int(n_total), R)
    lo = r * base + min(r, rem)
    return lo, lo + base + (1 if r < rem else 0)


def all_reduce_sum(t):
    """In-place SUM all-reduce of a small tensor; identity without a process group."""
    if is_dist() and world_size() > 1:
        td.all_reduce(t, op=td.ReduceOp.SUM)
    return t


def barrier():
    if is_dist():
        td.barrier()
