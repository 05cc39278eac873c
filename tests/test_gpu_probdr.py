"""Probabilistic delayed rejection on the GPU -- algorithms.Prob_DRHMC_Metrop / _Hastings
(src/algorithms.py:649-818) and pyhmc.PyHMC_multistep_tries / _tries3 (src/pyhmc.py:233-483) --
against goldens made by the real reference classes and against the numpy oracle, with replayed
momenta and uniforms (2m-1 uniforms per transition: u_acc0, u_try1, u_acc1, ...).

Tolerances: float64 trajectories and probabilities within 1e-9 relative (the ghost products
multiply several exp() of energy differences), accept codes (-2 = stage declined, -1, j) and
the work counters identical."""
import glob
import json
import os

import numpy as np
import pytest
import torch

from hmc_b200 import algorithms, models, pyhmc
from oracle import models as omodels
from oracle import replay, samplers

pytestmark = pytest.mark.gpu

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "probdr_*.npz")))


def _dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


def _models(kind, d):
    if kind == "funnel":
        return omodels.Funnel(d), models.Funnel(d)
    return omodels.Rosenbrock(d), models.Rosenbrock(d)


def _close(a, b, rtol=1e-9, atol=1e-11):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    ok = np.isclose(a, b, rtol=rtol, atol=atol, equal_nan=True)
    if not ok.all():
        err = np.abs(a - b) / (np.abs(b) + atol / rtol)
        raise AssertionError("max rel err %.3e (%d of %d entries off)" % (np.nanmax(err), (~ok).sum(), ok.size))


def _cls(name):
    mod, cls = name.split(".")
    return getattr({"algorithms": algorithms, "pyhmc": pyhmc}[mod], cls)


def test_goldens_exist():
    assert len(GOLDEN) == 4


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_golden_from_the_reference(path):
    g = np.load(path)
    spec = json.loads(str(g["spec"]))
    a = spec["args"]
    _, pmodel = _models(spec["model"]["kind"], spec["model"]["d"])
    s = _cls(spec["cls"])(pmodel, device=_dev())
    q = g["q0"]
    for t in range(spec["T"]):
        q, p, code, avec, counts, ptries = s.multi_step(a["m"], q, a["N"], a["ss"], a["fsub"], p=g["P"][t], u=g["U"][t])
        assert np.array_equal(code, g["code"][t]), t
        assert np.array_equal(counts, g["counts"][t]), t
        _close(q, g["q"][t])
        _close(p, g["p"][t])
        _close(avec, g["avec"][t])
        _close(ptries, g["ptries"][t])
        q = g["q"][t]                                       # follow the reference's chain exactly


@pytest.mark.parametrize("kind,d,N,eps", [("funnel", 10, 6, 0.6), ("funnel", 50, 10, 0.4), ("rosenbrock", 10, 8, 0.12)])
@pytest.mark.parametrize("m,fsub", [(1, 2.0), (2, 2.0), (3, 2.0), (4, 1.5)])
@pytest.mark.parametrize("cls,rule", [("algorithms.Prob_DRHMC_Metrop", "metrop"), ("pyhmc.PyHMC_multistep_tries3", "hastings")])
def test_vs_oracle(kind, d, N, eps, m, fsub, cls, rule):
    om, pmodel = _models(kind, d)
    C = 61
    q0 = np.random.default_rng(m * 31 + d).uniform(-1, 1, (C, d))
    if kind == "rosenbrock":
        q0 = 1.0 + 0.1 * q0
    P, U = replay.make_draws(3000 + d + m, 1, C, d, 2 * m - 1)
    s = _cls(cls)(pmodel, device=_dev())
    q, p, code, avec, counts, ptries = s.multi_step(m, q0, N, eps, fsub, p=P[0], u=U[0])
    ref = [samplers.prob_multistep_step(om, rule, m, q0[c], N, eps, fsub, replay.Draws([P[0, c]], U[0, c]))
           for c in range(C)]
    assert np.array_equal(code, [r[2] for r in ref])
    assert np.array_equal(counts, [r[4] for r in ref])
    _close(avec, np.array([r[3] for r in ref]))
    _close(ptries, np.array([r[5] for r in ref]))
    _close(q, np.array([r[0] for r in ref]))
    _close(p, np.array([r[1] for r in ref]))
    if m >= 3 and kind == "funnel":
        assert (code == -2).any() and (code >= 1).any()


def test_declined_chains_keep_state_and_do_no_extra_work():
    """A chain that declines stage 1 returns q0 with the counters of one trajectory only."""
    d, C, m, N = 10, 4096, 3, 6
    s = pyhmc.PyHMC_multistep_tries(models.Funnel(d), device=_dev(), seed=5)
    q0 = torch.as_tensor(np.random.default_rng(1).uniform(-1, 1, (C, d)), device=_dev())
    q, p, code, avec, counts, ptries = s.multi_step(m, q0, N, 0.6, 2.0)
    code, counts = code.cpu().numpy(), counts.cpu().numpy()
    q, avec, ptries = q.cpu().numpy(), avec.cpu().numpy(), ptries.cpu().numpy()
    dec = code == -2
    assert dec.sum() > 10
    first = dec & (avec[:, 1] == 0) & (ptries[:, 1] == 1)           # declined right after stage 0
    assert first.sum() > 5
    assert np.array_equal(counts[first], np.tile([2, N + 1, 1], (first.sum(), 1)))
    assert np.array_equal(q[dec], q0.cpu().numpy()[dec])
    assert ((ptries >= 0) & (ptries <= 1)).all() and ((avec >= 0) & (avec <= 1)).all()
    acc0 = code == 0
    assert np.array_equal(counts[acc0], np.tile([2, N + 1, 1], (acc0.sum(), 1)))


def test_needs_enough_injected_uniforms():
    s = algorithms.Prob_DRHMC_Hastings(models.Funnel(10), device=_dev())
    q0 = np.zeros((3, 10))
    with pytest.raises(ValueError):
        s.multi_step(3, q0, 5, 0.1, 2.0, p=np.zeros((3, 10)), u=np.zeros((3, 3)))


@pytest.mark.parametrize("cls", [algorithms.Prob_DRHMC_Metrop, algorithms.Prob_DRHMC_Hastings])
def test_production_path_keeps_the_funnel_marginal(cls):
    """With the library's own draws (Philox: u_acc and u_try from one stream, consumption order as the
    reference), exact funnel samples stay funnel samples."""
    dev = _dev()
    d, C = 10, 1 << 15
    g = torch.Generator().manual_seed(11)
    v = 3.0 * torch.randn(C, 1, generator=g, dtype=torch.float64)
    q = torch.cat([v, torch.exp(v / 2) * torch.randn(C, d - 1, generator=g, dtype=torch.float64)], dim=1).to(dev)
    s = cls(models.Funnel(d), device=dev, seed=5)
    seen = set()
    for _ in range(30):
        q, p, code, avec, counts, ptries = s.multi_step(3, q, 8, 0.5, 2.0)
        seen |= set(torch.unique(code).tolist())
    assert {-2, -1, 0, 1, 2} <= seen
    assert abs(q[:, 0].mean().item()) < 5 * 3.0 / np.sqrt(C)
    assert abs(q[:, 0].std().item() - 3.0) < 0.08
