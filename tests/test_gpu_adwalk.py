"""adaptivehmc.AdHMC / AdHMC_tint on the GPU (src/adaptivehmc.py:196-354, :362-519): the step
size walks inside the trajectory.  Goldens made by the real reference classes (replayed momenta,
uniforms and np.random.choice draws), the numpy oracle on fresh inputs, and a stationarity check
of the production path.

Tolerances: float64 states within 1e-9, probabilities within 1e-8 relative (products of several
exp() of energy differences and normalised weights); accept codes (1 / 0 / 2 / 100+i) and work
counters identical."""
import glob
import json
import os

import numpy as np
import pytest
import torch

from hmc_b200 import adaptivehmc, models
from oracle import models as omodels
from oracle import replay, samplers

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(__file__)
GOLDEN = sorted(glob.glob(os.path.join(HERE, "golden", "adwalk_*.npz")))


def _dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


def _models(spec):
    if spec["kind"] == "funnel":
        return models.Funnel(spec["d"]), omodels.Funnel(spec["d"])
    return models.Rosenbrock(spec["d"]), omodels.Rosenbrock(spec["d"])


def _close(a, b, rtol, atol):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    ok = np.isclose(a, b, rtol=rtol, atol=atol, equal_nan=True)
    if not ok.all():
        err = np.abs(a - b) / (np.abs(b) + atol / rtol)
        raise AssertionError("max rel err %.3e (%d of %d entries off)" % (np.nanmax(err), (~ok).sum(), ok.size))


def test_goldens_exist():
    assert len(GOLDEN) == 4


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_golden_from_the_reference(path):
    g = np.load(path)
    case = json.loads(str(g["spec"]))
    pm, _ = _models(case["model"])
    s = getattr(adaptivehmc, case["cls"])(pm, device=_dev())
    seen = set()
    for t in range(g["P"].shape[0]):
        q_in = g["q0"] if t == 0 else g["q"][t - 1]
        q, p, code, probs, checks, counts = s.hmc_step(q_in, p=g["P"][t], u=g["U"][t], **case["args"])
        assert np.array_equal(code, g["code"][t]), (code, g["code"][t])
        assert np.array_equal(counts, g["counts"][t])
        _close(checks, g["checks"][t], rtol=1e-8, atol=1e-13)
        _close(probs, g["probs"][t], rtol=1e-8, atol=1e-13)
        _close(q, g["q"][t], rtol=1e-9, atol=1e-10)
        _close(p, g["p"][t], rtol=1e-9, atol=1e-10)
        seen |= set(np.asarray(code).tolist())
    assert {0.0, 1.0} <= seen
    if "wide" in path:
        assert any(c >= 100 for c in seen)                   # the blow-up return (src/adaptivehmc.py:296-301)


@pytest.mark.parametrize("kind,d", [("funnel", 10), ("funnel", 50), ("rosenbrock", 10)])
@pytest.mark.parametrize("tint", [False, True])
def test_vs_oracle_fresh(kind, d, tint):
    pm, om = _models(dict(kind=kind, d=d))
    C = 37
    q0 = np.random.default_rng(d + tint).uniform(-1, 1, (C, d))
    smin, smax = 0.02, 0.8
    if kind == "rosenbrock":
        q0, smin, smax = 1.0 + 0.1 * q0, 0.005, 0.2
    if tint:
        kw = dict(Tint=2.5 if kind == "funnel" else 0.6, ntimes=6, smin=smin, smax=smax, ntry_eps0=8, nsteps_eps0=1)
        N, cls = 6, adaptivehmc.AdHMC_tint
    else:
        kw = dict(Nleap=35, nleap=5, smin=smin, smax=smax, ntry_eps0=8, nsteps_eps0=1)
        N, cls = 7, adaptivehmc.AdHMC
    P, U = replay.make_draws(900 + d, 1, C, d, N + 2)
    out = cls(pm, device=_dev()).hmc_step(q0, p=P[0], u=U[0], **kw)
    ref = [samplers.adhmc_step(om, q0[c], replay.Draws([P[0, c]], U[0, c]), **kw) for c in range(C)]
    assert np.array_equal(out[2], [float(r[2]) for r in ref])
    assert np.array_equal(out[5], [r[5] for r in ref])
    _close(out[4], np.array([r[4] for r in ref]), rtol=1e-8, atol=1e-13)
    _close(out[3], np.array([r[3] for r in ref]), rtol=1e-8, atol=1e-13)
    _close(out[0], np.array([r[0] for r in ref]), rtol=1e-9, atol=1e-10)
    _close(out[1], np.array([r[1] for r in ref]), rtol=1e-9, atol=1e-10)


def test_production_path_keeps_the_funnel_marginal_and_counts_its_work():
    dev = _dev()
    d, C = 10, 1 << 14
    g = torch.Generator().manual_seed(11)
    v = 3.0 * torch.randn(C, 1, generator=g, dtype=torch.float64)
    q = torch.cat([v, torch.exp(v / 2) * torch.randn(C, d - 1, generator=g, dtype=torch.float64)], dim=1).to(dev)
    s = adaptivehmc.AdHMC(models.Funnel(d), device=dev, seed=3)
    acc = []
    for _ in range(8):
        q, p, code, probs, checks, counts = s.hmc_step(q, Nleap=20, nleap=5, smin=0.01, smax=0.5, ntry_eps0=8)
        acc.append((code == 1).double().mean().item())
    assert 0.2 < np.mean(acc) < 0.99
    assert abs(q[:, 0].mean().item()) < 6 * 3.0 / np.sqrt(C)
    assert abs(q[:, 0].std().item() - 3.0) < 0.12
    # N = 4 segments, 2 ratios: 1 + 3*(2 + 1) trajectories of 5 steps + two sweeps of 8 one-step trials
    ok = code < 100
    assert (counts[ok][:, 2] <= 1 + 3 * 3 + 16).all() and (counts[ok][:, 2] >= 1 + 3 * 3 + 8).all()
    eps = checks[:, 2]
    assert ((eps[ok] > 0.01 * 0.7) & (eps[ok] < 0.5 * 1.5)).all()


def test_argument_checks():
    s = adaptivehmc.AdHMC(models.Funnel(10), device=_dev())
    q0 = np.zeros((3, 10))
    with pytest.raises(ValueError):
        s.hmc_step(q0, Nleap=5, nleap=5)                     # one segment only
    with pytest.raises(Exception):
        s.hmc_step(q0, Nleap=20, nleap=5, ratios=[0.5, 3.0])  # not closed under reciprocal
