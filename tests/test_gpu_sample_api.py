"""The sample-loop layer of algorithms.py on the GPU against the oracle: ``HMC.sample``
(src/algorithms.py:233-254: accepts for every iteration, samples / Hs / counts only for
i > burnin -> nsamples-1 rows), ``DRHMC.sample`` (:534-555: every row kept) and
``HMC.adapt_stepsize`` (:209-230) with replayed draws.  oracle.samplers.hmc_sample /
drhmc_sample / hmc_adapt_stepsize are pinned bit-exact to the unmodified reference in
tests/test_oracle_vs_reference.py."""
import numpy as np
import pytest
import torch

from hmc_b200 import algorithms, models
from oracle import models as omodels
from oracle import replay, samplers

pytestmark = pytest.mark.gpu


def _dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


def _close(a, b, rtol=1e-10, atol=1e-11):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    assert np.allclose(a, b, rtol=rtol, atol=atol, equal_nan=True), np.abs(a - b).max()


@pytest.mark.parametrize("epsadapt", [0, 7])
def test_hmc_sample_single_chain_vs_oracle(epsadapt):
    d, ns, nb, nleap, eps = 10, 9, 4, 8, 0.1
    om = omodels.Funnel(d)
    T = ns + nb + (epsadapt + 1 if epsadapt else 0)
    P, U = replay.make_draws(31, T, 1, d, 1)
    q0 = np.random.default_rng(8).uniform(-1, 1, d)
    s = algorithms.HMC(models.Funnel(d), device=_dev())
    st = s.sample(q0, nsamples=ns, burnin=nb, step_size=eps, nleap=nleap, epsadapt=epsadapt, draws=(P, U))
    o = samplers.hmc_sample(om, q0, ns, nb, eps, nleap, replay.Draws(list(P[:, 0]), U[:, 0, 0]), epsadapt=epsadapt)
    assert st.i == o["i"] == ns + nb
    assert st.samples.shape == (ns - 1, d) and st.accepts.shape == (ns + nb,)
    assert st.Hs.shape == (ns - 1, 2) and st.counts.shape == (ns - 1, 3)
    assert np.array_equal(st.accepts, o["accepts"]) and np.array_equal(st.counts, o["counts"])
    _close(st.samples, o["samples"])
    _close(st.Hs, o["Hs"], rtol=1e-9, atol=1e-10)
    assert np.isclose(s.step_size, o["step_size"], rtol=1e-9)
    if epsadapt:
        assert s.step_size != eps and len(s.adapt_trace) == epsadapt + 1


def test_hmc_sample_batched_rows_and_values():
    d, C, ns, nb, nleap, eps = 20, 5, 6, 2, 5, 0.07
    om = omodels.Funnel(d)
    P, U = replay.make_draws(32, ns + nb, C, d, 1)
    q0 = np.random.default_rng(9).uniform(-1, 1, (C, d))
    s = algorithms.HMC(models.Funnel(d), device=_dev())
    st = s.sample(q0, nsamples=ns, burnin=nb, step_size=eps, nleap=nleap, draws=(P, U))
    assert st.samples.shape == (ns - 1, C, d) and st.accepts.shape == (ns + nb, C) and st.counts.shape == (ns - 1, C, 3)
    for c in range(C):
        o = samplers.hmc_sample(om, q0[c], ns, nb, eps, nleap, replay.Draws(list(P[:, c]), U[:, c, 0]))
        assert np.array_equal(st.accepts[:, c], o["accepts"])
        _close(st.samples[:, c], o["samples"])


def test_drhmc_sample_vs_oracle_all_rows_kept():
    d, C, ns, nb, nleap, eps, nprops, f = 10, 4, 7, 3, 6, 0.5, 3, 2.0
    om = omodels.Funnel(d)
    P, U = replay.make_draws(33, ns + nb, C, d, nprops)
    q0 = np.random.default_rng(10).uniform(-1, 1, (C, d))
    s = algorithms.DRHMC(models.Funnel(d), device=_dev())
    st = s.sample(q0, nsamples=ns, burnin=nb, step_size=eps, nleap=nleap, nprops=nprops, stepfactor=f, draws=(P, U))
    assert st.i == ns + nb
    assert st.samples.shape == (ns + nb, C, d) and st.accepts.shape == (ns + nb, C)
    assert st.avecs.shape == (ns + nb, C, nprops) and st.counts.shape == (ns + nb, C, 3)
    for c in range(C):
        q = q0[c]
        for t in range(ns + nb):
            q, p, code, avec, cnt = samplers.drhmc_step(om, q, nleap, eps, nprops, f, replay.Draws([P[t, c]], U[t, c]))
            assert st.accepts[t, c] == code and st.counts[t, c].tolist() == cnt
            _close(st.samples[t, c], q)
            _close(st.avecs[t, c], avec, rtol=1e-9, atol=1e-11)
    assert len(set(st.accepts.ravel().tolist())) >= 3


def test_adapt_stepsize_replay_matches_reference_update_for_one_chain():
    d, K, nleap, eps0 = 10, 20, 10, 0.1
    om = omodels.Funnel(d)
    P, U = replay.make_draws(34, K + 1, 1, d, 1)
    q0 = np.random.default_rng(11).uniform(-1, 1, d)
    s = algorithms.HMC(models.Funnel(d), device=_dev())
    s.step_size = eps0
    q = s.adapt_stepsize(q0, K, p=P, u=U, nsamples=1, burnin=0, step_size=eps0, nleap=nleap)
    q_o, eps_o, used = samplers.hmc_adapt_stepsize(om, q0, K, nleap, eps0, replay.Draws(list(P[:, 0]), U[:, 0, 0]))
    assert np.allclose(s.adapt_trace, used, rtol=1e-9)
    assert np.isclose(s.step_size, eps_o, rtol=1e-9)
    _close(q, q_o, rtol=1e-8, atol=1e-9)      # 21 chained transitions, each started from the previous GPU state


def test_adapt_stepsize_many_chains_uses_the_mean_statistic():
    """C chains: one shared step size driven by the chain-mean of the per-chain statistic the reference feeds
    (prob clipped to [0,1], 0 for a chain that did not move)."""
    d, C, K, nleap, eps0 = 10, 64, 8, 10, 0.1
    om = omodels.Funnel(d)
    P, U = replay.make_draws(35, K + 1, C, d, 1)
    q0 = np.random.default_rng(12).uniform(-1, 1, (C, d))
    s = algorithms.HMC(models.Funnel(d), device=_dev())
    s.step_size = eps0
    s.adapt_stepsize(q0, K, p=P, u=U, nsamples=1, burnin=0, step_size=eps0, nleap=nleap)
    da = samplers.DualAveragingStepSize(eps0)
    q, eps = q0.copy(), eps0
    for i in range(K + 1):
        probs = []
        for c in range(C):
            qn, p, acc, Hs, cnt = samplers.hmc_step_alg(om, q[c], nleap, eps, replay.Draws([P[i, c]], U[i, c]))
            with np.errstate(all="ignore"):
                pr = 0.0 if (qn == q[c]).all() else np.exp(Hs[0] - Hs[1])
            probs.append(0.0 if np.isnan(pr) else min(pr, 1.0))
            q[c] = qn
        noisy, avg = da.update(float(np.mean(probs)))
        eps = noisy if i < K else avg
    assert np.isclose(s.step_size, eps, rtol=1e-8)


@pytest.mark.parametrize("cls,extra", [(algorithms.HMC, {}), (algorithms.DRHMC, dict(nprops=3, stepfactor=2.0))])
def test_sample_without_history_equals_the_stored_run(cls, extra):
    """sample(store=False): one fused call per phase -- the final state, accept counts, work counters and running moments
    are those of the stored run over its nsamples kept transitions (both start from the same seed)."""
    d, C, ns, nb = 50, 4500, 6, 3
    q0 = np.random.default_rng(14).uniform(-1, 1, (C, d))
    kw = dict(nsamples=ns, burnin=nb, step_size=0.25, nleap=8, **extra)
    a = cls(models.Funnel(d), device=_dev(), seed=77)
    b = cls(models.Funnel(d), device=_dev(), seed=77)
    if cls is algorithms.HMC:
        # the stored run: burn-in + sampling as single steps, everything kept here for the comparison
        qs, codes, q = [], [], torch.as_tensor(q0, device=_dev())
        for t in range(ns + nb):
            o = a.engine.hmc_step(q, 8, 0.25, a._rule, want_p=False)
            q = o["q"]
            qs.append(q)
            codes.append(o["code"])
        kept, acc = torch.stack(qs[nb:]), torch.stack(codes[nb:]) == 1
        counts = torch.tensor([2 * ns, ns * 9, ns], device=_dev(), dtype=torch.int32).expand(C, 3)
    else:
        st = a.sample(q0, **kw)
        kept = torch.as_tensor(st.samples[nb:], device=_dev())
        acc = torch.as_tensor(st.accepts[nb:], device=_dev()) >= 1
        counts = torch.as_tensor(st.counts[nb:], device=_dev()).sum(dim=0).int()
    f = b.sample(q0, store=False, **kw)
    assert f.samples is None and f.i == ns + nb
    assert np.array_equal(f.q, kept[-1].cpu().numpy())
    assert np.array_equal(f.n_accept, acc.sum(dim=0).cpu().numpy())
    assert np.array_equal(f.counts, counts.cpu().numpy())
    s1 = torch.zeros((C, d), dtype=torch.float64, device=_dev())
    s2 = torch.zeros_like(s1)
    for t in range(ns):
        s1 += kept[t]
        s2 = torch.addcmul(s2, kept[t], kept[t])
    assert np.allclose(f.moment_sum, s1.cpu().numpy(), rtol=1e-13, atol=1e-12)
    assert np.allclose(f.moment_sumsq, s2.cpu().numpy(), rtol=1e-13, atol=1e-12)
