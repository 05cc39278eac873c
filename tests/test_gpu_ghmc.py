"""GPU parity for the generalised-HMC family: algorithms.GHMC,
Persistent_GHMC (src/algorithms.py:259-357) and DRGHMC (:561-646) against the oracle, which is
pinned bit-for-bit to the reference classes (tests/test_oracle_vs_reference.py)."""
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


def _models(kind, d):
    if kind == "funnel":
        return models.Funnel(d), omodels.Funnel(d)
    if kind == "rosenbrock":
        return models.Rosenbrock(d), omodels.Rosenbrock(d)
    return models.CorrGaussian.ar1(d), omodels.CorrGaussian.ar1(d)


def _close(a, b, rtol=1e-10, atol=1e-11):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    ok = np.isclose(a, b, rtol=rtol, atol=atol, equal_nan=True)
    assert ok.all(), "max rel err %.3e" % np.nanmax(np.abs(a - b) / (np.abs(b) + atol / rtol))


@pytest.mark.parametrize("kind,d,eps", [("funnel", 10, 0.1), ("funnel", 50, 0.06), ("rosenbrock", 20, 0.03),
                                        ("gaussian", 12, 0.15)])
@pytest.mark.parametrize("alpha", [0.0, 0.8])
def test_ghmc_three_transitions_with_caches(kind, d, eps, alpha):
    pm, om = _models(kind, d)
    C, N, T = 29, 5, 3
    rng = np.random.default_rng(d)
    q = rng.uniform(-1, 1, (C, d))
    if kind == "rosenbrock":
        q = 1.0 + 0.1 * q
    p = rng.standard_normal((C, d))
    P, U = replay.make_draws(300 + d, T, C, d, 1)
    s = algorithms.GHMC(pm, device=_dev())
    gq, gp, gV, gVg = q, p, None, None
    ref = [(q[c], p[c], None, None) for c in range(C)]
    for t in range(T):
        gq, gp, gV, gVg, acc, Hs, counts = s.step(gq, gp, N, eps, alpha, V_q=gV, V_gq=gVg, pn=P[t], u=U[t])
        out = [samplers.ghmc_step(om, ref[c][0], ref[c][1], N, eps, alpha, replay.Draws([P[t, c]], U[t, c]), ref[c][2], ref[c][3])
               for c in range(C)]
        assert np.array_equal(acc, [o[4] for o in out])
        assert np.array_equal(counts, [o[6] for o in out])        # [1, N+1, 1] first, then [0, N, 1]
        _close(gq, np.array([o[0] for o in out]))
        _close(gp, np.array([o[1] for o in out]))                 # flipped on rejection
        # chains whose first proposal was rejected still have V_q = V_gq = None in the reference: NaN here
        rV = np.array([np.nan if o[2] is None else o[2] for o in out], dtype=np.float64)
        rVg = np.array([np.full(d, np.nan) if o[3] is None else o[3] for o in out], dtype=np.float64)
        _close(gV, rV, rtol=1e-10, atol=1e-10)
        _close(gVg, rVg, rtol=1e-10, atol=1e-10)
        _close(Hs, np.array([o[5] for o in out]), rtol=1e-9, atol=1e-10)
        ref = [o[:4] for o in out]
    assert [0, N, 1] in counts.tolist()


def test_persistent_ghmc_carries_its_uniform():
    pm, om = _models("funnel", 10)
    C, N, T = 31, 4, 4
    rng = np.random.default_rng(4)
    q, p = rng.uniform(-1, 1, (C, 10)), rng.standard_normal((C, 10))
    u = rng.uniform(-1, 1, C)
    P, _ = replay.make_draws(41, T, C, 10, 1)
    s = algorithms.Persistent_GHMC(pm, device=_dev())
    g = (q, p, u, None, None)
    ref = [(q[c], p[c], u[c], None, None) for c in range(C)]
    for t in range(T):
        out_g = s.step(g[0], g[1], g[2], N, 0.1, 0.5, 0.37, V_q=g[3], V_gq=g[4], pn=P[t])
        out = [samplers.pghmc_step(om, *ref[c][:3], N, 0.1, 0.5, 0.37, replay.Draws([P[t, c]], []), ref[c][3], ref[c][4])
               for c in range(C)]
        assert np.array_equal(out_g[5], [o[5] for o in out])
        _close(out_g[0], np.array([o[0] for o in out]))
        _close(out_g[2], np.array([o[2] for o in out]), rtol=1e-9, atol=1e-12)
        g, ref = out_g[:5], [o[:5] for o in out]


@pytest.mark.parametrize("kind,d,nleap,eps", [("funnel", 10, 6, 0.6), ("funnel", 50, 8, 0.35), ("rosenbrock", 10, 8, 0.1)])
def test_drghmc_vs_oracle(kind, d, nleap, eps):
    pm, om = _models(kind, d)
    C = 47
    rng = np.random.default_rng(d + 1)
    q, p = rng.uniform(-1, 1, (C, d)), rng.standard_normal((C, d))
    if kind == "rosenbrock":
        q = 1.0 + 0.1 * q
    P, U = replay.make_draws(500 + d, 1, C, d, 3)
    s = algorithms.DRGHMC(pm, device=_dev())
    gq, gp, code, avec, counts = s.step(q, p, nleap, eps, 0.6, 3, 2.0, pn=P[0], u=U[0])
    ref = [samplers.drghmc_step(om, q[c], p[c], nleap, eps, 0.6, 3, 2.0, replay.Draws([P[0, c]], U[0, c])) for c in range(C)]
    assert np.array_equal(code, [r[2] for r in ref])
    assert np.array_equal(counts, [r[4] for r in ref])
    _close(gq, np.array([r[0] for r in ref]), rtol=1e-9, atol=1e-10)
    _close(gp, np.array([r[1] for r in ref]), rtol=1e-9, atol=1e-10)   # p0 (unflipped) on rejection
    _close(avec, np.array([r[3] for r in ref]), rtol=1e-8, atol=1e-12)


def test_ghmc_sample_keeps_the_gaussian_invariant():
    dev = _dev()
    d, C = 20, 1 << 14
    pm = models.CorrGaussian.ar1(d)
    Lc = np.linalg.cholesky(pm.Sigma)
    g = torch.Generator().manual_seed(0)
    q0 = (torch.randn(C, d, generator=g, dtype=torch.float64) @ torch.as_tensor(Lc).T).to(dev)
    for name, kw in (("ghmc", dict(alpha=0.7)), ("pghmc", dict(alpha=0.7, delta=0.3)), ("drghmc", dict(alpha=0.7, nprops=2, stepfactor=2.0))):
        s = algorithms.alg_dict[name](pm, device=dev, seed=3)
        st = s.sample(q0, nsamples=12, burnin=0, step_size=0.25, nleap=3, **kw)
        q = st.samples[-1]
        assert abs(q.mean().item()) < 0.03, name
        assert abs(q.std(dim=0).mean().item() - 1.0) < 0.03, name


def test_leapfrog_Vgq_and_V_vandg_standalone():
    """algorithms.HMC.leapfrog_Vgq / V_vandg called on their own (src/algorithms.py:109-115,150-172)."""
    import torch
    from hmc_b200 import algorithms, models
    from oracle import models as omodels
    from oracle import samplers
    dev = torch.device("cuda:0")
    d, C, N, eps = 12, 17, 9, 0.07
    pm, om = models.Rosenbrock(d), omodels.Rosenbrock(d)
    rng = np.random.default_rng(4)
    q0 = 1.0 + 0.1 * rng.standard_normal((C, d))
    p0 = rng.standard_normal((C, d))
    s = algorithms.HMC(pm, device=dev)
    V, Vg = s.V_vandg(q0)
    for c in range(C):
        v, g = om.log_prob_and_grad(q0[c])
        assert np.isclose(V[c], -v, rtol=1e-12) and np.allclose(Vg[c], -g, rtol=1e-12, atol=1e-13)
    s.Vgcount = s.leapcount = 0
    out_cached = s.leapfrog_Vgq(q0, p0, N, eps, V, Vg)
    assert s.Vgcount == N and s.leapcount == 1
    out_plain = s.leapfrog_Vgq(q0, p0, N, eps)
    assert s.Vgcount == 2 * N + 1
    for c in range(C):
        kin, cnt = samplers.Kinetics(None), samplers.Counters()
        ref = samplers._leapfrog_Vgq(om, kin, q0[c], p0[c], N, eps, cnt, -om.log_prob(q0[c]), -om.grad_log_prob(q0[c]))
        for out in (out_cached, out_plain):
            for a, b in zip(out, ref):
                assert np.allclose(np.asarray(a[c], dtype=float), np.asarray(b, dtype=float), rtol=1e-10, atol=1e-11)
    assert np.array_equal(s.unit_norm_KE(p0), 0.5 * (p0 ** 2).sum(axis=1)) or np.allclose(s.unit_norm_KE(p0), 0.5 * (p0 ** 2).sum(axis=1), rtol=1e-14)


def test_gaussian_partial_refresh_reaches_the_generic_kernel(monkeypatch):
    """hmcb_hmc_step with rng->p_prev on the f64 dense Gaussian: the tensor-path kernel only draws a full refresh,
    so the call must be served by the generic kernel -- same result as with HMCB_GAUSS_GENERIC=1, and the momentum
    really is alpha*p_prev + sqrt(1-alpha^2)*draw."""
    import ctypes as C
    from hmc_b200 import _lib
    d, n, alpha = 20, 257, 0.8
    dev = torch.device("cuda:0")
    s = algorithms.HMC(models.CorrGaussian.ar1(d), device=dev, seed=5)
    e = s.engine
    m = e._ensure()
    g = torch.Generator().manual_seed(0)
    q = torch.randn(n, d, generator=g, dtype=torch.float64).to(dev)
    pv = torch.randn(n, d, generator=g, dtype=torch.float64).to(dev)
    res = []
    for generic in ("", "1"):
        if generic:
            monkeypatch.setenv("HMCB_GAUSS_GENERIC", "1")
        else:
            monkeypatch.delenv("HMCB_GAUSS_GENERIC", raising=False)
        qo, p0 = torch.empty_like(q), torch.empty_like(q)
        so = _lib.StepOut()
        so.p0_out = _lib.ptr(p0)
        r, keep = e._rng(None, None, 1, n, p_prev=pv, alpha=alpha)
        _lib.check(e.lib.hmcb_hmc_step(C.byref(m), C.byref(e._chains(q, qo)), C.byref(r), 5, 0.1, _lib.RULE_ALG,
                                       C.byref(so), e._stream()))
        # the same call with a full refresh gives the plain draw
        z, so2 = torch.empty_like(q), _lib.StepOut()
        so2.p0_out = _lib.ptr(z)
        r2, _ = e._rng(None, None, 1, n)
        _lib.check(e.lib.hmcb_hmc_step(C.byref(m), C.byref(e._chains(q, torch.empty_like(q))), C.byref(r2), 5, 0.1,
                                       _lib.RULE_ALG, C.byref(so2), e._stream()))
        torch.cuda.synchronize()
        assert torch.equal(p0, alpha * pv + np.sqrt(1 - alpha ** 2) * z)
        res.append((qo, p0))
    assert torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][1], res[1][1])


@pytest.mark.parametrize("persistent", [False, True])
def test_sample_loop_vs_oracle(persistent):
    """GHMC.sample / Persistent_GHMC.sample (src/algorithms.py:284-303, :336-357) with replayed draws against
    oracle.samplers.ghmc_sample / pghmc_sample (pinned to the reference loops in tests/test_oracle_vs_reference.py):
    every iteration stored, caches starting as None, u carried."""
    d, C, ns, nb, N, alpha = 10, 6, 7, 3, 5, 0.6
    eps = 0.4 if persistent else 0.25      # (u in (-1, 1) accepts easily: a longer step for some rejections)
    pm, om = _models("funnel", d)
    T = ns + nb
    P, U = replay.make_draws(410, T, C, d, 1)
    rng = np.random.default_rng(2)
    q0, p0, u0 = rng.uniform(-1, 1, (C, d)), rng.standard_normal((C, d)), rng.standard_normal(C)
    kw = dict(nsamples=ns, burnin=nb, step_size=eps, nleap=N, alpha=alpha)
    if persistent:
        st = algorithms.Persistent_GHMC(pm, device=_dev()).sample(q0, p=p0, u=u0, draws=P, delta=0.37, **kw)
    else:
        st = algorithms.GHMC(pm, device=_dev()).sample(q0, p=p0, draws=(P, U), **kw)
    assert st.i == T and st.samples.shape == (T, C, d) and st.accepts.shape == (T, C) and st.counts.shape == (T, C, 3)
    for c in range(C):
        if persistent:
            o = samplers.pghmc_sample(om, q0[c], ns, nb, eps, N, alpha, 0.37,
                                      replay.Draws([np.array([u0[c]])] + list(P[:, c]), []), p=p0[c])
            _close(st.u[:, c], o["u"], rtol=1e-8, atol=1e-11)
        else:
            o = samplers.ghmc_sample(om, q0[c], ns, nb, eps, N, alpha, replay.Draws(list(P[:, c]), U[:, c, 0]), p=p0[c])
        assert np.array_equal(st.accepts[:, c], o["accepts"]) and np.array_equal(st.counts[:, c], o["counts"])
        _close(st.samples[:, c], o["samples"])
        _close(st.Hs[:, c], o["Hs"], rtol=1e-9, atol=1e-10)
    assert len(set(st.accepts.ravel().tolist())) >= 2
