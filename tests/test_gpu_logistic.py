"""Bayesian logistic regression: the tcgen05 tensor-core gradient and the HMC transition
built on it, float32 on the GPU against float64 reference arithmetic (golden made by the
real reference's pyhmc.HMC.step; oracle for fresh inputs).  Tolerance 1e-5 relative
(BASELINE.json north_star, fp32)."""
import json
import os

import numpy as np
import pytest
import torch

from hmc_b200 import algorithms, models, pyhmc
from oracle import models as omodels
from oracle import replay, samplers

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(__file__)


def _dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


def _rel(a, b, floor):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.abs(a - b).max() / max(np.abs(b).max(), floor)


def _data(N, d, seed):
    rng = np.random.default_rng(seed)
    X = (rng.standard_normal((N, d)) / np.sqrt(d)).astype(np.float32)
    y = (rng.random(N) < 1.0 / (1.0 + np.exp(-(X.astype(np.float64) @ rng.standard_normal(d))))).astype(np.float32)
    return X, y


@pytest.mark.parametrize("N,d,C", [(1000, 128, 130), (3000, 24, 12), (777, 5, 300), (20000, 100, 64), (128, 16, 1)])
def test_logp_and_grad_vs_float64(N, d, C):
    X, y = _data(N, d, N + d)
    om = omodels.LogisticRegression(X.astype(np.float64), y.astype(np.float64), 1.5)
    s = pyhmc.HMC(models.LogisticRegression(X, y, 1.5), device=_dev())
    th = (0.5 * np.random.default_rng(1).standard_normal((C, d))).astype(np.float32)
    V = s.V(th)
    Vg = s.V_g(th)
    rV = np.array([-om.log_prob(t.astype(np.float64)) for t in th])
    rg = np.array([-om.grad_log_prob(t.astype(np.float64)) for t in th])
    assert _rel(V, rV, 1.0) < 2e-6
    assert _rel(Vg, rg, 1.0) < 2e-5          # relative to the largest gradient entry


def test_logp_and_grad_at_config4_size():
    """BASELINE.json config 4's data size (N=100,000 rows, D=128) on a 256-chain slice: the fp32 accumulator of the
    tensor core truncates, and that error grows with the number of rows summed before a drain -- only this size shows
    whether the drain interval holds it."""
    N, d, C = 100000, 128, 256
    X, y = _data(N, d, 4)
    s = pyhmc.HMC(models.LogisticRegression(X, y, 1.0), device=_dev())
    th = (0.3 * np.random.default_rng(5).standard_normal((C, d))).astype(np.float32)
    V, Vg = np.asarray(s.V(th), dtype=np.float64), np.asarray(s.V_g(th), dtype=np.float64)
    X64, y64, t64 = X.astype(np.float64), y.astype(np.float64), th.astype(np.float64)
    Z = t64 @ X64.T                                                    # [C, N]
    rV = (np.logaddexp(0.0, Z) - Z * y64[None, :]).sum(axis=1) + 0.5 * (t64 * t64).sum(axis=1)
    rg = (1.0 / (1.0 + np.exp(-Z)) - y64[None, :]) @ X64 + t64
    om = omodels.LogisticRegression(X64[:2000], y64[:2000], 1.0)       # the closed forms above are the oracle's
    assert abs(-om.log_prob(t64[0]) - ((np.logaddexp(0.0, Z[0, :2000]) - Z[0, :2000] * y64[:2000]).sum()
                                       + 0.5 * (t64[0] ** 2).sum())) < 1e-8
    eV, eg = _rel(V, rV, 1.0), _rel(Vg, rg, 1.0)
    print("config-4 size: rel err of V %.3e, of grad V %.3e (of the largest entry)" % (eV, eg))
    assert eV < 2e-6
    assert eg < 2e-5


def test_golden_transitions_within_1e5():
    g = np.load(os.path.join(HERE, "golden", "logit_hmc.npz"))
    spec = json.loads(str(g["spec"]))
    a = spec["args"]
    s = pyhmc.HMC(models.LogisticRegression(g["X"], g["y"], spec["tau"]), device=_dev())
    for t in range(g["P"].shape[0]):
        q_in = g["q0"] if t == 0 else g["q"][t - 1]
        q, p, code, Hs, counts = s.step(q_in.astype(np.float32), a["N"], a["step_size"], p=g["P"][t].astype(np.float32),
                                        u=g["U"][t])
        assert np.array_equal(code, g["code"][t])
        assert np.array_equal(counts, g["counts"][t])
        assert _rel(q, g["q"][t], 1e-2) < 1e-5
        assert _rel(p, g["p"][t], 1e-2) < 1e-5
        assert np.abs(np.asarray(Hs, dtype=np.float64) - g["Hs"][t]).max() < 2e-2 * max(1.0, np.abs(g["Hs"][t]).max() * 1e-4)
        # the energy ERROR decides acceptance: it comes from float64 sums and must be tight
        dH = (np.asarray(Hs, dtype=np.float64)[:, 0] - np.asarray(Hs, dtype=np.float64)[:, 1])
        assert np.abs(dH - (g["Hs"][t][:, 0] - g["Hs"][t][:, 1])).max() < 0.05


def test_fresh_inputs_both_rules_and_ragged_sizes():
    N, d, C = 5000, 37, 141
    X, y = _data(N, d, 9)
    om = omodels.LogisticRegression(X.astype(np.float64), y.astype(np.float64), 1.0)
    q0 = (0.2 * np.random.default_rng(2).standard_normal((C, d))).astype(np.float32)
    P, U = replay.make_draws(55, 1, C, d, 1)
    P = P.astype(np.float32)
    for cls, fn in ((pyhmc.HMC, samplers.hmc_step_pyhmc), (algorithms.HMC, samplers.hmc_step_alg)):
        s = cls(models.LogisticRegression(X, y, 1.0), device=_dev())
        q, p, code, Hs, counts = s.step(q0, 8, 0.008, p=P[0], u=U[0])
        ref = [fn(om, q0[c].astype(np.float64), 8, 0.008, replay.Draws([P[0, c].astype(np.float64)], U[0, c])) for c in range(C)]
        assert (np.asarray(code) != np.array([r[2] for r in ref])).sum() <= 1     # a flip needs |u - prob| < fp32 noise
        assert np.array_equal(counts, [r[4] for r in ref])
        same = np.asarray(code) == np.array([r[2] for r in ref])
        assert _rel(np.asarray(q)[same], np.array([r[0] for r in ref])[same], 1e-2) < 1e-5


def test_production_path_samples_the_posterior():
    """Philox path: chains started at the posterior mode stay calibrated -- the mean of the
    chains matches a long float64 oracle run within Monte Carlo error."""
    N, d, C = 2000, 8, 4096
    X, y = _data(N, d, 21)
    om = omodels.LogisticRegression(X.astype(np.float64), y.astype(np.float64), 1.0)
    th = np.zeros(d)
    for _ in range(200):    # Newton iterations to the mode, float64
        z = X.astype(np.float64) @ th
        sg = 1 / (1 + np.exp(-z))
        H = (X.T * (sg * (1 - sg))) @ X + np.eye(d)
        th = th + np.linalg.solve(H, om.grad_log_prob(th))
    cov = np.linalg.inv(H)
    s = algorithms.HMC(models.LogisticRegression(X, y, 1.0), device=_dev(), seed=3)
    q = torch.as_tensor(np.tile(th, (C, 1)).astype(np.float32), device=_dev())
    acc = 0.0
    for _ in range(30):
        q, p, code, Hs, counts = s.step(q, 10, 0.03)
        acc += (code == 1).double().mean().item() / 30
    assert acc > 0.6
    sd = np.sqrt(np.diag(cov))
    assert np.abs(q.mean(0).cpu().numpy() - th).max() < 6 * sd.max() / np.sqrt(C) + 0.02 * sd.max()
    assert np.abs(q.std(0).cpu().numpy() / sd - 1).max() < 0.1


def test_mala_over_the_tensor_core_gradient_vs_oracle():
    """mala.MALA.step on the logistic model (hmcb_logreg_mala_step): float32 state against the float64 oracle with the
    same injected normals and uniforms, cached V / grad V carried from step to step as the reference does."""
    from hmc_b200 import mala
    N, d, C, eps = 3000, 24, 61, 0.15
    X, y = _data(N, d, 31)
    om = omodels.LogisticRegression(X.astype(np.float64), y.astype(np.float64), 1.2)
    s = mala.MALA(models.LogisticRegression(X, y, 1.2), device=_dev())
    q0 = (0.3 * np.random.default_rng(3).standard_normal((C, d))).astype(np.float32)
    P, U = replay.make_draws(91, 3, C, d, 1)
    P = P.astype(np.float32)
    q, V, Vg = q0, None, None
    rq, rV, rVg = [x.astype(np.float64) for x in q0], [None] * C, [None] * C
    seen = set()
    for t in range(3):
        q, code, V, Vg = s.step(q, eps, V, Vg, z=P[t], u=U[t])
        ref = [samplers.mala_step(om, rq[c], eps, replay.Draws([P[t, c].astype(np.float64)], U[t, c]), rV[c], rVg[c]) for c in range(C)]
        rcode = np.array([r[1] for r in ref])
        seen |= set(rcode.tolist())
        assert (np.asarray(code) != rcode).sum() <= 1            # a flip needs |u - prob| within fp32 noise
        same = np.asarray(code) == rcode
        for c in range(C):
            rq[c], rV[c], rVg[c] = ref[c][0], ref[c][2], ref[c][3]
        R = np.array(rq)
        assert _rel(np.asarray(q)[same], R[same], 1e-2) < 1e-5
        assert _rel(np.asarray(V)[same], np.array(rV)[same], 1.0) < 2e-6
        assert _rel(np.asarray(Vg)[same], np.array(rVg)[same], 1.0) < 3e-5
        # continue from the GPU's own state on both sides (a flipped chain would otherwise diverge)
        rq = [np.asarray(q, dtype=np.float64)[c] for c in range(C)]
        rV, rVg = [None] * C, [None] * C
        V, Vg = (V, Vg) if t % 2 == 0 else (None, None)         # with and without the carried caches
    print("codes seen:", seen)
    assert 1.0 in seen


def test_mala_logistic_fused_steps_equal_single_steps():
    from hmc_b200 import mala
    N, d, C, eps, K = 2000, 16, 300, 0.03, 4
    X, y = _data(N, d, 33)
    a = mala.MALA(models.LogisticRegression(X, y, 1.0), device=_dev(), seed=5)
    b = mala.MALA(models.LogisticRegression(X, y, 1.0), device=_dev(), seed=5)
    q0 = torch.as_tensor((0.2 * np.random.default_rng(4).standard_normal((C, d))).astype(np.float32), device=_dev())
    q, V, Vg, nacc = q0, None, None, 0
    for k in range(K):
        q, code, V, Vg = a.step(q, eps, V, Vg)
        nacc = nacc + (code == 1).int()
    qb, codeb, Vb, Vgb = b.step(q0, eps, n_steps=K)
    assert torch.equal(qb, q) and torch.equal(codeb, code) and torch.equal(Vb, V) and torch.equal(Vgb, Vg)
    assert torch.equal(b.n_accept.int(), nacc)
    assert 0.2 < nacc.double().mean().item() / K <= 1.0
