"""GPU parity for MALA (src/mala.py) and AdHMC_eps0 (src/adaptivehmc.py:105-192):
goldens made by the real reference, the numpy oracle on fresh inputs, and
properties of the fused K-step MALA kernel."""
import glob
import json
import os

import numpy as np
import pytest
import torch

from hmc_b200 import adaptivehmc, mala, models
from oracle import models as omodels
from oracle import replay, samplers

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(__file__)


def _dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


def _models(spec):
    k, d = spec["kind"], spec["d"]
    if k == "funnel":
        return models.Funnel(d), omodels.Funnel(d)
    if k == "rosenbrock":
        return models.Rosenbrock(d), omodels.Rosenbrock(d)
    return models.CorrGaussian.ar1(d), omodels.CorrGaussian.ar1(d)


def _close(a, b, rtol=1e-10, atol=1e-11):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    ok = np.isclose(a, b, rtol=rtol, atol=atol, equal_nan=True)
    assert ok.all(), "max rel err %.3e" % np.nanmax(np.abs(a - b) / (np.abs(b) + atol / rtol))


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(HERE, "golden", "mala_*.npz"))),
                         ids=lambda p: os.path.basename(p)[:-4])
def test_mala_golden(path):
    g = np.load(path)
    case = json.loads(str(g["spec"]))
    pm, _ = _models(case["model"])
    s = mala.MALA(pm, device=_dev())
    T = g["P"].shape[0]
    for t in range(T):
        # teacher-forced: start every step from the reference's state and caches
        q_in = g["q0"] if t == 0 else g["q"][t - 1]
        V_in = None if t == 0 else g["V"][t - 1]
        Vg_in = None if t == 0 else g["Vg"][t - 1]
        q, code, V, Vg = s.step(q_in, case["eps"], V_in, Vg_in, z=g["P"][t], u=g["U"][t])
        assert np.array_equal(code, g["code"][t])
        _close(q, g["q"][t])
        _close(V, g["V"][t], rtol=1e-10, atol=1e-10)
        _close(Vg, g["Vg"][t], rtol=1e-10, atol=1e-10)


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(HERE, "golden", "adhmc_*.npz"))),
                         ids=lambda p: os.path.basename(p)[:-4])
def test_adhmc_eps0_golden(path):
    g = np.load(path)
    case = json.loads(str(g["spec"]))
    pm, _ = _models(case["model"])
    a = case["args"]
    s = adaptivehmc.AdHMC_eps0(pm, device=_dev())
    for t in range(g["P"].shape[0]):
        q_in = g["q0"] if t == 0 else g["q"][t - 1]
        q, p, code, probs, aux, counts = s.hmc_step(q_in, a["Nleap"], smin=a["smin"], smax=a["smax"], Tint=a.get("Tint", 0),
                                                    ntry=a["ntry"], nsteps=a["nsteps"], p=g["P"][t], u=g["U"][t])
        assert np.array_equal(code, g["code"][t]), (code, g["code"][t])
        assert np.array_equal(counts, g["counts"][t])
        _close(aux, g["aux"][t], rtol=1e-9, atol=1e-12)
        _close(probs, g["probs"][t], rtol=1e-8, atol=1e-12)
        _close(q, g["q"][t], rtol=1e-9, atol=1e-10)
        _close(p, g["p"][t], rtol=1e-9, atol=1e-10)


@pytest.mark.parametrize("kind,d,eps", [("funnel", 50, 0.05), ("funnel", 100, 0.03), ("rosenbrock", 20, 0.03),
                                        ("gaussian", 50, 0.2), ("funnel", 2, 0.3)])
def test_mala_vs_oracle_fresh(kind, d, eps):
    pm, om = _models(dict(kind=kind, d=d))
    C = 41
    q0 = np.random.default_rng(d).uniform(-1, 1, (C, d))
    if kind == "rosenbrock":
        q0 = 1.0 + 0.1 * q0
    P, U = replay.make_draws(500 + d, 2, C, d, 1)
    s = mala.MALA(pm, device=_dev())
    q, code, V, Vg = s.step(q0, eps, z=P[0], u=U[0])
    ref = [samplers.mala_step(om, q0[c], eps, replay.Draws([P[0, c]], U[0, c])) for c in range(C)]
    assert np.array_equal(code, [r[1] for r in ref])
    _close(q, np.array([r[0] for r in ref]))
    assert np.array_equal(s.Vgcount.cpu().numpy(), [r[4][1] for r in ref])   # 2 gradients on the first step
    # second step reuses the caches: 1 gradient
    q2, code2, V2, Vg2 = s.step(q, eps, V, Vg, z=P[1], u=U[1])
    ref2 = [samplers.mala_step(om, ref[c][0], eps, replay.Draws([P[1, c]], U[1, c]), ref[c][2], ref[c][3]) for c in range(C)]
    assert np.array_equal(code2, [r[1] for r in ref2])
    _close(q2, np.array([r[0] for r in ref2]))
    assert (s.Vgcount == 1).all()


def test_mala_fused_steps_equal_single_steps():
    """K steps in one launch == K launches of one step (same Philox counters)."""
    d, C, K = 50, 4096, 7
    dev = _dev()
    q0 = torch.randn(C, d, dtype=torch.float64, device=dev) * 0.5
    a = mala.MALA(models.Funnel(d), device=dev, seed=9)
    b = mala.MALA(models.Funnel(d), device=dev, seed=9)
    qa, ca, Va, Vga = a.step(q0, 0.05, n_steps=K)
    qb, Vb, Vgb = q0, None, None
    nacc = torch.zeros(C, dtype=torch.int32, device=dev)
    for _ in range(K):
        qb, cb, Vb, Vgb = b.step(qb, 0.05, Vb, Vgb)
        nacc += (cb == 1).int()
    assert torch.equal(qa, qb) and torch.equal(ca, cb) and torch.equal(Va, Vb)
    assert torch.equal(a.n_accept, nacc)
    assert 0.2 < (nacc.double().mean().item() / K) < 1.0


def test_mala_stationarity_gaussian():
    dev = _dev()
    d, C = 20, 1 << 15
    pm = models.CorrGaussian.ar1(d)
    Lc = np.linalg.cholesky(pm.Sigma)
    g = torch.Generator().manual_seed(0)
    q = (torch.randn(C, d, generator=g, dtype=torch.float64) @ torch.as_tensor(Lc).T).to(dev)
    s = mala.MALA(pm, device=dev, seed=1)
    q, code, V, Vg = s.step(q, 0.25, n_steps=40)
    acc = s.n_accept.double().mean().item() / 40
    assert 0.3 < acc < 0.99
    assert abs(q.mean().item()) < 0.02
    assert abs(q.std(dim=0).mean().item() - 1.0) < 0.02


def test_adhmc_vs_oracle_fresh_and_stationarity():
    dev = _dev()
    d, C = 10, 33
    pm, om = _models(dict(kind="funnel", d=d))
    q0 = np.random.default_rng(3).uniform(-1, 1, (C, d))
    P, U = replay.make_draws(77, 1, C, d, 3)
    s = adaptivehmc.AdHMC_eps0(pm, device=dev)
    out = s.hmc_step(q0, 12, smin=0.02, smax=0.6, ntry=10, nsteps=1, p=P[0], u=U[0])
    ref = [samplers.adhmc_eps0_step(om, q0[c], 12, replay.Draws([P[0, c]], U[0, c]), smin=0.02, smax=0.6, ntry=10, nsteps=1)
           for c in range(C)]
    assert np.array_equal(out[2], [r[2] for r in ref])
    assert np.array_equal(out[5], [r[5] for r in ref])
    _close(out[4], np.array([r[4] for r in ref]), rtol=1e-9, atol=1e-12)
    _close(out[0], np.array([r[0] for r in ref]), rtol=1e-9, atol=1e-10)
    # production path keeps the funnel's v marginal
    C2 = 1 << 14
    g = torch.Generator().manual_seed(5)
    v = 3.0 * torch.randn(C2, 1, generator=g, dtype=torch.float64)
    q = torch.cat([v, torch.exp(v / 2) * torch.randn(C2, d - 1, generator=g, dtype=torch.float64)], dim=1).to(dev)
    s2 = adaptivehmc.AdHMC_eps0(pm, device=dev, seed=2)
    for _ in range(10):
        q = s2.hmc_step(q, 10, smin=0.01, smax=0.5, ntry=8, nsteps=1)[0]
    assert abs(q[:, 0].mean().item()) < 6 * 3.0 / np.sqrt(C2)
    assert abs(q[:, 0].std().item() - 3.0) < 0.1


def test_get_stepsize_standalone_vs_oracle():
    """AdHMC_eps0.get_stepsize used on its own (src/adaptivehmc.py:111-134), both of its modes."""
    dev = _dev()
    d, C, ntry = 10, 29, 8
    pm, om = _models(dict(kind="funnel", d=d))
    rng = np.random.default_rng(12)
    q0 = rng.uniform(-1, 1, (C, d))
    p0 = rng.standard_normal((C, d))
    U = rng.random((C, 2))
    s = adaptivehmc.AdHMC_eps0(pm, device=dev)
    ss, prob = s.get_stepsize(q0, p0, smin=0.02, smax=0.6, ntry=ntry, nsteps=2, u=U)
    pe = s.get_stepsize(q0, p0, smin=0.02, smax=0.6, ntry=ntry, nsteps=2, eps=0.1)
    for c in range(C):
        kin, cnt = samplers.Kinetics(None), samplers.Counters()
        steps, pp, cdf = samplers._eps0_trials(om, kin, q0[c], p0[c], 0.02, 0.6, ntry, 2, cnt)
        sx = cdf.min() + (1.0 - cdf.min()) * U[c, 0]
        isx = np.where(sx > cdf)[0][-1]
        want = steps[isx] + (steps[isx + 1] - steps[isx]) * U[c, 1]
        assert np.isclose(ss[c], want, rtol=1e-12) and np.isclose(prob[c], pp[isx + 1], rtol=1e-9)
        assert np.isclose(pe[c], pp[np.where(steps > 0.1)[0][0]], rtol=1e-9)


def test_mala_cached_potential_without_cached_gradient():
    """V_q and V_gq are cached independently in the reference (src/mala.py:49-52): a step given V_q but not V_gq must
    use that V_q (and evaluate the gradient itself) -- the result equals the step given both, or neither."""
    d, C = 10, 33
    pm, om = models.Funnel(d), omodels.Funnel(d)
    q0 = np.random.default_rng(3).uniform(-1, 1, (C, d))
    P, U = replay.make_draws(8, 1, C, d, 1)
    s = mala.MALA(pm, device=_dev())
    V = s.V(q0)
    full = s.step(q0, 0.05, V, s.V_g(q0), z=P[0], u=U[0])
    only_v = s.step(q0, 0.05, V, None, z=P[0], u=U[0])
    none = s.step(q0, 0.05, None, None, z=P[0], u=U[0])
    for a, b, c in zip(full, only_v, none):
        assert np.array_equal(a, b) and np.array_equal(a, c)
    ref = [samplers.mala_step(om, q0[c], 0.05, replay.Draws([P[0, c]], U[0, c]), float(V[c]), None) for c in range(C)]
    assert np.array_equal(only_v[1], [r[1] for r in ref])
    _close(only_v[0], np.array([r[0] for r in ref]))
