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


# ------------------------------------------------------------------ HMC.adapt_stepsize across ranks
class _HostEngine:
    """Stands in for hmc_b200.engine.Engine in HMC.adapt_stepsize on a box without a GPU: advances this rank's
    chains with the numpy oracle (draws keyed by global chain id and iteration, as the kernels key Philox) and
    returns what the kernel returns -- the new state and the per-chain statistic prob (clipped to [0,1], 0 for a
    chain that did not move; hmcb_step_out.accept_prob).  Test infrastructure only."""

    def __init__(self, d, chain_offset):
        from oracle import models as om
        self.model, self.dev, self.device, self.dtype = om.Funnel(d), torch.device("cpu"), torch.device("cpu"), torch.float64
        self.off, self.iteration = chain_offset, 0

    def check_dim(self, b):
        assert b.d == self.model.d

    def hmc_step(self, q, n_leap, step_size, rule, p=None, u=None, want_p=True, want_prob=False, q_out=None):
        from oracle import replay
        qn, prob = q.clone(), torch.zeros(q.shape[0], dtype=torch.float64)
        for c in range(q.shape[0]):
            rng = np.random.default_rng([17, self.off + c, self.iteration])
            x = q[c].numpy().copy()
            out = samplers.hmc_step_alg(self.model, x, n_leap, step_size,
                                        replay.Draws([rng.standard_normal(x.size)], [rng.random()]))
            with np.errstate(all="ignore"):
                pr = 0.0 if (out[0] == x).all() else float(np.exp(out[3][0] - out[3][1]))
            prob[c] = 0.0 if np.isnan(pr) else min(pr, 1.0)
            qn[c] = torch.as_tensor(out[0])
        self.iteration += 1
        return {"q": qn, "prob": prob}


def _adapt(q, lo, hi, K):
    from hmc_b200 import models
    s = algorithms.HMC(models.Funnel(q.shape[1]))
    s.engine = _HostEngine(q.shape[1], lo)
    s.step_size = 0.1
    qn = s.adapt_stepsize(q[lo:hi].numpy().copy(), K, nsamples=1, burnin=0, step_size=0.1, nleap=6)   # numpy in -> numpy out
    return s.step_size, list(s.adapt_trace), qn


def _adapt_worker(rank, world, port, q, K, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    dist.init_from_env(backend="gloo")
    lo, hi = dist.shard_range(q.shape[0])
    eps, trace, qn = _adapt(q, lo, hi, K)
    out.put((rank, eps, trace, lo, hi, qn))
    dist.barrier()
    torch.distributed.destroy_process_group()


@pytest.mark.timeout(180)
def test_world_size_2_adapt_stepsize_ends_on_the_single_rank_step_size():
    """HMC.adapt_stepsize (src/algorithms.py:209-230 made batched): the all-reduced statistic makes two ranks with
    half the chains each walk the step-size sequence one rank walks with all of them."""
    C, d, K = 11, 6, 7                     # odd chain count: ragged shards
    q = torch.as_tensor(np.random.default_rng(3).uniform(-1, 1, (C, d)))
    eps1, trace1, q1 = _adapt(q, 0, C, K)
    assert len(trace1) == K + 1 and eps1 != 0.1
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_adapt_worker, args=(r, 2, port, q, K, out)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([out.get(timeout=150) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=30)
        assert p.exitcode == 0
    assert res[0][1] == res[1][1] and res[0][2] == res[1][2]          # both ranks: the very same sequence
    assert np.allclose(res[0][2], trace1, rtol=1e-12) and np.isclose(res[0][1], eps1, rtol=1e-12)
    for rank, eps, trace, lo, hi, qn in res:
        assert np.allclose(qn, q1[lo:hi], rtol=1e-9, atol=1e-12)
