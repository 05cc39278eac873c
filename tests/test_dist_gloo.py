"""The N>1 host path on CPU: world_size-2 gloo process group, chains sharded by
global id, only small moment vectors all-reduced -- the cross-rank ESS / R-hat /
dual-averaging arithmetic must equal the single-process result."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from hmc_b200 import algorithms, dist, utils
from oracle import samplers


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, x, mean, std, probs, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    r, w, _ = dist.init_from_env(backend="gloo")
    assert (r, w) == (rank, world) and dist.is_dist()
    lo, hi = dist.shard_range(x.shape[1])
    xs = x[:, lo:hi]
    sums, C = utils.all_reduce_sums(utils.local_moment_sums(xs, mean), hi - lo)
    ess = utils.ess_from_sums(sums, C, std)
    rhat = utils.rhat_from_sums(sums, C, float(x.shape[0]))
    m, s = utils.mean_std_from_sums(sums, C, float(x.shape[0]))
    # dual averaging: all ranks feed the same all-reduced statistic -> identical step sizes
    da = algorithms.DualAveragingStepSize(0.1)
    eps = None
    for t in range(probs.shape[0]):
        stat = torch.stack([probs[t, lo:hi].sum(), torch.tensor(float(hi - lo), dtype=torch.float64)])
        dist.all_reduce_sum(stat)
        eps, _ = da.update(float(stat[0] / stat[1]))
    dist.barrier()
    q.put((rank, C, ess.numpy(), rhat.numpy(), m.numpy(), s.numpy(), eps))
    torch.distributed.destroy_process_group()


@pytest.mark.timeout(120)
def test_world_size_2_moment_reductions_match_single_process():
    rng = np.random.default_rng(0)
    T, C, d = 40, 37, 5                      # odd chain count: ragged shards
    x = torch.as_tensor(rng.standard_normal((T, C, d)).cumsum(axis=0) * 0.2 + rng.standard_normal((1, C, d)))
    mean, std = np.linspace(-0.2, 0.2, d), np.linspace(1.0, 2.0, d)
    probs = torch.as_tensor(rng.random((6, C)))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, x, mean, std, probs, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=100) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=30)
        assert p.exitcode == 0
    ess_ref, _ = samplers.get_ess_true(x.numpy(), mean, std)           # oracle == reference src/utils.py:8-27
    da = algorithms.DualAveragingStepSize(0.1)
    for t in range(probs.shape[0]):
        eps_ref, _ = da.update(float(probs[t].mean()))
    xm = x.numpy()
    W = xm.var(axis=0, ddof=1).mean(axis=0)
    B_over_T = xm.mean(axis=0).var(axis=0, ddof=1)
    rhat_np = np.sqrt(((T - 1) / T * W + B_over_T) / W)
    for rank, Ctot, ess, rhat, m, s, eps in res:
        assert Ctot == C
        assert np.allclose(ess, ess_ref, rtol=1e-10)
        assert np.allclose(rhat, rhat_np, rtol=1e-10)
        assert np.allclose(m, xm.mean(axis=(0, 1)), rtol=1e-10, atol=1e-12)
        assert np.allclose(s, xm.reshape(-1, d).std(axis=0), rtol=1e-10)
        assert np.isclose(eps, eps_ref, rtol=1e-12)
    assert res[0][6] == res[1][6]


def _gather_worker(rank, world, port, x, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    from hmc_b200 import results
    dist.init_from_env(backend="gloo")
    lo, hi = dist.shard_range(x.shape[1])
    got = results.gather_to_rank0(x[:, lo:hi].contiguous())
    q.put((rank, None if got is None else np.asarray(got)))
    dist.barrier()
    torch.distributed.destroy_process_group()


@pytest.mark.timeout(120)
def test_world_size_2_gather_restores_global_chain_order():
    """results.gather_to_rank0 stands in for comm.gather + concatenate(axis=1)
    (scripts_old/funnel/funnelmpi.py:193-201); ragged shards."""
    x = torch.arange(6 * 7 * 3, dtype=torch.float64).reshape(6, 7, 3)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gather_worker, args=(r, 2, port, x, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=100) for _ in procs)
    for p in procs:
        p.join(timeout=30)
        assert p.exitcode == 0
    assert res[1] is None
    assert np.array_equal(res[0], x.numpy())
