"""GPU parity: the CUDA path (through the Python classes -> C ABI -> kernels)
against (1) the committed goldens produced by the real reference and (2) the
numpy oracle on freshly seeded inputs, with replayed momentum/uniform draws.

Tolerances (BASELINE.json north_star): fp64 trajectories within 1e-10 relative,
fp32 within 1e-5, accept codes and work counters identical."""
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

RTOL64, ATOL64 = 1e-10, 1e-11
GOLDEN = [p for p in sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))
          if os.path.basename(p).split("_")[0] in ("hmc", "alghmc", "drhmc")]


def _dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


def _product_model(spec):
    k = spec["kind"]
    if k == "funnel":
        return models.Funnel(spec["d"], spec.get("sigma", 3.0))
    if k == "rosenbrock":
        return models.Rosenbrock(spec["d"], spec.get("a", 1.0), spec.get("b", 100.0), spec.get("T", 20.0))
    if k == "gaussian":
        return models.CorrGaussian.ar1(spec["d"], spec.get("rho", 0.9))
    raise ValueError(k)


def _oracle_model(spec):
    k = spec["kind"]
    if k == "funnel":
        return omodels.Funnel(spec["d"], spec.get("sigma", 3.0))
    if k == "rosenbrock":
        return omodels.Rosenbrock(spec["d"], spec.get("a", 1.0), spec.get("b", 100.0), spec.get("T", 20.0))
    if k == "gaussian":
        return omodels.CorrGaussian.ar1(spec["d"], spec.get("rho", 0.9))
    raise ValueError(k)


def _close(a, b, rtol=RTOL64, atol=ATOL64):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    ok = np.isclose(a, b, rtol=rtol, atol=atol, equal_nan=True)
    if not ok.all():
        err = np.abs(a - b) / (np.abs(b) + atol / rtol)
        raise AssertionError("max rel err %.3e at %s (%d of %d entries off)" %
                             (np.nanmax(err), np.unravel_index(np.nanargmax(err), err.shape), (~ok).sum(), ok.size))


# ------------------------------------------------------------------ goldens (made by the real reference)
@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_golden(path):
    g = np.load(path)
    case = json.loads(str(g["spec"]))
    dev = _dev()
    model = _product_model(case["model"])
    inv = g["invmetric"] if g["invmetric"].size else None
    modname, clsname = case["cls"].split(".")
    a = case["args"]
    T = g["P"].shape[0]
    if clsname == "DRHMC":
        s = algorithms.DRHMC(model, invmetric_diag=inv, device=dev)
    elif modname == "algorithms":
        s = algorithms.HMC(model, invmetric_diag=inv, device=dev)
    elif clsname == "PyHMC":
        s = pyhmc.PyHMC(model, invmetric_diag=inv, device=dev)
    else:
        s = pyhmc.HMC(model, invmetric_diag=inv, device=dev)
    for t in range(T):
        q_in = g["q0"] if t == 0 else g["q"][t - 1]  # teacher-forced: each transition starts from the reference state
        if clsname == "DRHMC":
            q, p, code, aux, counts = s.step(q_in, a["nleap"], a["step_size"], a["nprops"], a["stepfactor"],
                                             p=g["P"][t], u=g["U"][t])
        elif clsname == "PyHMC":
            q, p, code, aux, counts = s.hmc_step(q_in, a["N"], a["step_size"], p=g["P"][t], u=g["U"][t])
        else:
            q, p, code, aux, counts = s.step(q_in, a["N"], a["step_size"], p=g["P"][t], u=g["U"][t])
        assert np.array_equal(np.asarray(code, dtype=np.float64), g["code"][t]), "accept codes differ at t=%d" % t
        assert np.array_equal(counts, g["counts"][t]), "work counters differ at t=%d" % t
        _close(q, g["q"][t])
        _close(p, g["p"][t])
        _close(aux, g["aux"][t], rtol=1e-9, atol=1e-12)


# ------------------------------------------------------------------ oracle, fresh inputs, every kernel shape
def _run_oracle_hmc(om, q0, P, U, N, eps, rule, inv=None):
    fn = samplers.hmc_step_pyhmc if rule == "pyhmc" else samplers.hmc_step_alg
    out = [fn(om, q0[c], N, eps, replay.Draws([P[c]], U[c]), inv) for c in range(q0.shape[0])]
    return [np.array([o[i] for o in out], dtype=np.float64) for i in range(5)]


@pytest.mark.parametrize("kind,d", [("funnel", 2), ("funnel", 4), ("funnel", 7), ("funnel", 10), ("funnel", 20),
                                    ("funnel", 33), ("funnel", 50), ("funnel", 100), ("rosenbrock", 2),
                                    ("rosenbrock", 10), ("rosenbrock", 50), ("rosenbrock", 100),
                                    ("gaussian", 3), ("gaussian", 20), ("gaussian", 50), ("gaussian", 100)])
@pytest.mark.parametrize("lpc", [0, 4, 32])
def test_hmc_step_vs_oracle_all_shapes(kind, d, lpc, monkeypatch):
    if lpc:
        monkeypatch.setenv("HMCB_LPC", str(lpc))
    else:
        monkeypatch.delenv("HMCB_LPC", raising=False)
    spec = dict(kind=kind, d=d)
    om, pmodel = _oracle_model(spec), _product_model(spec)
    C = 37  # ragged: not a multiple of any group size
    rng = np.random.default_rng(d * 7 + lpc)
    q0 = rng.uniform(-1, 1, (C, d))
    eps, N = (0.05, 9)
    if kind == "rosenbrock":
        q0, eps = 1.0 + 0.1 * q0, 0.02
    P, U = replay.make_draws(d, 1, C, d, 1)
    for rule, cls in (("pyhmc", pyhmc.HMC), ("alg", algorithms.HMC)):
        s = cls(pmodel, device=_dev())
        q, p, code, Hs, counts = s.step(q0, N, eps, p=P[0], u=U[0])
        rq, rp, rcode, rHs, rcounts = _run_oracle_hmc(om, q0, P[0], U[0], N, eps, rule)
        assert np.array_equal(code, rcode) and np.array_equal(counts, rcounts)
        _close(q, rq)
        _close(p, rp)
        _close(Hs, rHs, rtol=1e-9, atol=1e-10)


@pytest.mark.parametrize("d,lpc", [(2, 0), (10, 8), (18, 4), (50, 2), (100, 0), (100, 32)])
def test_rosenbrock_other_parameters_on_padded_shapes(d, lpc, monkeypatch):
    """a != 1, b, T away from the defaults on lane shapes with padding slots: the pair gradient keeps padding pairs out
    through one masked constant (models.cuh) -- they must stay at q = p = 0 whatever a is.  HMC and DRHMC against the oracle."""
    if lpc:
        monkeypatch.setenv("HMCB_LPC", str(lpc))
    else:
        monkeypatch.delenv("HMCB_LPC", raising=False)
    spec = dict(kind="rosenbrock", d=d, a=1.7, b=40.0, T=7.0)
    om, pmodel = _oracle_model(spec), _product_model(spec)
    C, N, eps = 29, 7, 0.03
    rng = np.random.default_rng(d + lpc)
    q0 = 1.5 + 0.2 * rng.uniform(-1, 1, (C, d))
    P, U = replay.make_draws(500 + d, 1, C, d, 3)
    s = algorithms.HMC(pmodel, device=_dev())
    q, p, code, Hs, counts = s.step(q0, N, eps, p=P[0], u=U[0][:, :1])
    rq, rp, rcode, rHs, rcounts = _run_oracle_hmc(om, q0, P[0], U[0][:, :1], N, eps, "alg")
    assert np.array_equal(code, rcode) and np.array_equal(counts, rcounts)
    _close(q, rq)
    _close(p, rp)
    _close(Hs, rHs, rtol=1e-9, atol=1e-10)
    dr = algorithms.DRHMC(pmodel, device=_dev())
    q, p, code, avec, counts = dr.step(q0, N, 4 * eps, 3, 2.0, p=P[0], u=U[0])
    for c in range(C):
        o = samplers.drhmc_step(om, q0[c], N, 4 * eps, 3, 2.0, replay.Draws([P[0, c]], U[0, c]))
        assert code[c] == o[2] and list(counts[c]) == o[4]
        _close(q[c], o[0])
        _close(avec[c], o[3], rtol=1e-9, atol=1e-11)


@pytest.mark.parametrize("kind,d", [("funnel", 10), ("funnel", 50), ("rosenbrock", 20), ("gaussian", 20)])
def test_hmc_step_diag_metric_vs_oracle(kind, d):
    spec = dict(kind=kind, d=d)
    om, pmodel = _oracle_model(spec), _product_model(spec)
    inv = np.linspace(0.4, 2.5, d)
    C = 19
    q0 = np.random.default_rng(5).uniform(-1, 1, (C, d))
    eps = 0.04
    if kind == "rosenbrock":
        q0, eps = 1.0 + 0.1 * q0, 0.015
    P, U = replay.make_draws(77, 1, C, d, 1)
    s = pyhmc.HMC(pmodel, invmetric_diag=inv, device=_dev())
    q, p, code, Hs, counts = s.step(q0, 11, eps, p=P[0], u=U[0])
    rq, rp, rcode, rHs, rcounts = _run_oracle_hmc(om, q0, P[0], U[0], 11, eps, "pyhmc", inv)
    assert np.array_equal(code, rcode) and np.array_equal(counts, rcounts)
    _close(q, rq)
    _close(p, rp)


@pytest.mark.parametrize("kind,d,nleap,eps", [("funnel", 10, 6, 0.6), ("funnel", 50, 10, 0.35), ("funnel", 50, 4, 0.5),
                                              ("rosenbrock", 10, 8, 0.1), ("gaussian", 12, 6, 0.9)])
@pytest.mark.parametrize("nprops,f", [(1, 2.0), (2, 2.0), (3, 2.0), (3, 1.5), (4, 2.0), (5, 1.3)])
@pytest.mark.parametrize("lpc", [0, 8])
def test_drhmc_step_vs_oracle(kind, d, nleap, eps, nprops, f, lpc, monkeypatch):
    if lpc:
        monkeypatch.setenv("HMCB_LPC", str(lpc))
    else:
        monkeypatch.delenv("HMCB_LPC", raising=False)
    spec = dict(kind=kind, d=d)
    om, pmodel = _oracle_model(spec), _product_model(spec)
    C = 45
    q0 = np.random.default_rng(nprops * 31 + d).uniform(-1, 1, (C, d))
    if kind == "rosenbrock":
        q0 = 1.0 + 0.1 * q0
    P, U = replay.make_draws(1000 + d + nprops, 1, C, d, nprops)
    s = algorithms.DRHMC(pmodel, device=_dev())
    q, p, code, avec, counts = s.step(q0, nleap, eps, nprops, f, p=P[0], u=U[0])
    ref = [samplers.drhmc_step(om, q0[c], nleap, eps, nprops, f, replay.Draws([P[0, c]], U[0, c])) for c in range(C)]
    rcode = np.array([r[2] for r in ref])
    rcounts = np.array([r[4] for r in ref])
    assert np.array_equal(code, rcode), "stage codes differ: %s vs %s" % (code, rcode)
    assert np.array_equal(counts, rcounts)
    _close(avec, np.array([r[3] for r in ref]), rtol=1e-9, atol=1e-12)
    _close(q, np.array([r[0] for r in ref]))      # north_star: 1e-10 relative (atol 1e-11)
    _close(p, np.array([r[1] for r in ref]))
    if kind == "funnel" and nprops >= 3:
        assert len(set(rcode.tolist())) >= 3  # the case really exercises several stages


def test_float32_within_1e5():
    d, C = 50, 64
    om, pmodel = omodels.Funnel(d), models.Funnel(d)
    q0 = np.random.default_rng(2).uniform(-1, 1, (C, d))
    P, U = replay.make_draws(9, 1, C, d, 1)
    s = pyhmc.HMC(pmodel, device=_dev(), dtype=torch.float32)
    q, p, code, Hs, counts = s.step(q0, 10, 0.05, p=P[0], u=U[0])
    rq, rp, rcode, rHs, rcounts = _run_oracle_hmc(om, q0, P[0].astype(np.float32).astype(np.float64), U[0], 10, 0.05, "pyhmc")
    assert (np.asarray(code) != rcode).sum() <= 1 and np.array_equal(counts, rcounts)
    same = np.asarray(code) == rcode
    err = np.abs(np.asarray(q, dtype=np.float64) - rq)[same] / (np.abs(rq[same]) + 1e-2)
    assert err.max() < 1e-5, err.max()


def test_unbatched_shapes_match_reference_convention():
    d = 10
    s = pyhmc.HMC(models.Funnel(d), device=_dev())
    q0 = np.random.default_rng(0).uniform(-1, 1, d)
    P, U = replay.make_draws(3, 1, 1, d, 1)
    q, p, acc, Hs, counts = s.step(q0, 20, 0.1, p=P[0, 0], u=U[0, 0])
    r = samplers.hmc_step_pyhmc(omodels.Funnel(d), q0, 20, 0.1, replay.Draws([P[0, 0]], U[0, 0]))
    assert q.shape == (d,) and isinstance(acc, float) and counts == [2, 21, 1] == r[4]
    _close(q, r[0])
    _close(Hs, r[3], rtol=1e-9)
    assert acc == r[2]


def test_leapfrog_and_energy_api_vs_oracle():
    d, C = 50, 33
    om, s = omodels.Funnel(d), pyhmc.HMC(models.Funnel(d), device=_dev())
    rng = np.random.default_rng(4)
    q0, p0 = rng.uniform(-1, 1, (C, d)), rng.standard_normal((C, d))
    q1, p1 = s.leapfrog(q0, p0, 25, 0.07)
    kin, cnt = samplers.Kinetics(), samplers.Counters()
    ref = [samplers.leapfrog(om, kin, q0[c], p0[c], 25, 0.07, cnt) for c in range(C)]
    _close(q1, np.array([r[0] for r in ref]))
    _close(p1, np.array([r[1] for r in ref]))
    _close(s.V(q0), np.array([-om.log_prob(x) for x in q0]), rtol=1e-12)
    _close(s.V_g(q0), np.array([-om.grad_log_prob(x) for x in q0]), rtol=1e-12, atol=1e-13)
    _close(s.H(q0, p0), np.array([-om.log_prob(x) + 0.5 * (y * y).sum() for x, y in zip(q0, p0)]), rtol=1e-12)
    # standalone metropolis (src/pyhmc.py:50-61) against the rule restated in numpy on the oracle's energies:
    # code 2 if prob is NaN/inf or sum(q0-q1) == 0, else reject iff u > min(1, prob)
    q1n, p1n = np.asarray(q1), np.asarray(p1)
    q1n[3], p1n[3] = q0[3], p0[3]                       # a proposal that did not move
    q0 = q0.copy()
    q0[5, 0] = -800.0                                   # exp(-v) overflows: H0 = inf, prob = inf
    u = np.random.default_rng(6).random(C)
    qa, pa, code, Hs = s.metropolis([q0, p0], [q1n, p1n], u=u)
    with np.errstate(all="ignore"):
        H0 = np.array([-om.log_prob(x) + 0.5 * (y * y).sum() for x, y in zip(q0, p0)])
        H1 = np.array([-om.log_prob(x) + 0.5 * (y * y).sum() for x, y in zip(q1n, p1n)])
        prob = np.exp(H0 - H1)
    bad = np.isnan(prob) | np.isinf(prob) | ((q0 - q1n).sum(axis=1) == 0)
    want = np.where(bad, 2.0, np.where(u > np.minimum(1.0, prob), 0.0, 1.0))
    assert np.array_equal(code, want) and want[3] == 2.0 and want[5] == 2.0 and {0.0, 1.0} <= set(want.tolist())
    acc = want == 1.0
    assert np.array_equal(np.asarray(qa)[acc], q1n[acc]) and np.array_equal(np.asarray(qa)[~acc], q0[~acc])
    _close(np.asarray(Hs)[~bad], np.stack([H0, H1], axis=1)[~bad], rtol=1e-12)


def test_empty_batch_is_a_no_op():
    s = pyhmc.HMC(models.Funnel(10), device=_dev())
    q, p, code, Hs, counts = s.step(np.zeros((0, 10)), 5, 0.1)
    assert q.shape == (0, 10) and code.shape == (0,)


def test_nan_and_no_move_codes():
    d = 10
    om = omodels.Funnel(d)
    q0 = np.zeros((4, d))
    q0[0, 0] = -800.0   # exp(-v) overflows -> NaN energies -> code 2 / -1
    q0[1] = 0.3
    P = np.random.default_rng(1).standard_normal((4, d))
    P[2] = 0.0          # p = 0 at q = 0 for theta, v=0: moves only through v
    U = np.full((4, 1), 0.5)
    for cls, bad in ((pyhmc.HMC, 2.0), (algorithms.HMC, -1.0)):
        s = cls(models.Funnel(d), device=_dev())
        q, p, code, Hs, counts = s.step(q0, 5, 0.1, p=P, u=U)
        rule = "pyhmc" if bad == 2.0 else "alg"
        rq, rp, rcode, rHs, rcounts = _run_oracle_hmc(om, q0, P, U, 5, 0.1, rule)
        assert np.array_equal(code, rcode) and code[0] == bad
        _close(q, rq)


@pytest.mark.parametrize("kind,d,N,eps", [("funnel", 10, 6, 0.6), ("funnel", 50, 10, 0.35), ("rosenbrock", 10, 8, 0.1)])
@pytest.mark.parametrize("m,fsub", [(1, 2.0), (2, 2.0), (3, 2.0), (4, 1.5)])
def test_pyhmc_multistep_vs_oracle(kind, d, N, eps, m, fsub):
    """pyhmc.PyHMC_multistep.multi_step: codes j/-1, first-stage <=, Hcount incl. ghost base energies."""
    spec = dict(kind=kind, d=d)
    om, pmodel = _oracle_model(spec), _product_model(spec)
    C = 53
    q0 = np.random.default_rng(m * 17 + d).uniform(-1, 1, (C, d))
    if kind == "rosenbrock":
        q0 = 1.0 + 0.1 * q0
    P, U = replay.make_draws(2000 + d + m, 1, C, d, m)
    U[0, 0, 0] = 0.0          # u == 0 with alpha_0 == 0 would split "<" from "<="; here it must accept unless alpha_0 < 0
    s = pyhmc.PyHMC_multistep(pmodel, device=_dev())
    q, p, code, avec, counts = s.multi_step(m, q0, N, eps, fsub, p=P[0], u=U[0])
    ref = [samplers.multistep_step(om, m, q0[c], N, eps, fsub, replay.Draws([P[0, c]], U[0, c])) for c in range(C)]
    assert np.array_equal(code, [r[2] for r in ref])
    assert np.array_equal(counts, [r[4] for r in ref])
    _close(avec, np.array([r[3] for r in ref]), rtol=1e-9, atol=1e-12)
    _close(q, np.array([r[0] for r in ref]))
    _close(p, np.array([r[1] for r in ref]))
    assert code[0] == 0       # the u = 0 chain accepted at the first stage (<=)
