"""K delayed-rejection transitions per call (hmcb_drhmc_sample: the stage kernels of several chain
ranges pipelined over streams, running moments and summed counters folded in) against single
hmcb_drhmc_step calls, which the oracle and golden tests pin to algorithms.DRHMC.step / get_num
(src/algorithms.py:474-530) and PyHMC_multistep / the probabilistic classes: every per-chain output
must be IDENTICAL, bit for bit, for every convention, kernel shape and number of stages, with
replayed and with Philox draws; K transitions in one call must equal K single calls; and the folded
running moments must equal hmcb_moments_accumulate's."""
import numpy as np
import pytest
import torch

from hmc_b200 import _lib, algorithms, models
from oracle import models as omodels
from oracle import replay, samplers

pytestmark = pytest.mark.gpu


def _dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


def _model(kind, d):
    return {"funnel": lambda: models.Funnel(d), "rosenbrock": lambda: models.Rosenbrock(d),
            "gaussian": lambda: models.CorrGaussian.ar1(d)}[kind]()


def _q0(kind, C, d, seed):
    rng = np.random.default_rng(seed)
    if kind == "funnel":
        v = 1.5 * rng.standard_normal((C, 1))
        return np.concatenate([v, np.exp(v / 2) * rng.standard_normal((C, d - 1))], axis=1)
    q = rng.uniform(-1, 1, (C, d))
    return 1.0 + 0.1 * q if kind == "rosenbrock" else q


def _same(a, b, keys):
    for k in keys:
        if a.get(k) is None and b.get(k) is None:
            continue
        assert torch.equal(a[k], b[k]), "%s differs: %d of %d entries" % (k, (a[k] != b[k]).sum().item(), a[k].numel())


@pytest.mark.parametrize("kind,d,nleap,eps", [("funnel", 10, 6, 0.6), ("funnel", 50, 10, 0.35), ("rosenbrock", 10, 8, 0.1),
                                              ("gaussian", 12, 6, 0.9)])
@pytest.mark.parametrize("nprops,f", [(1, 2.0), (2, 2.0), (3, 2.0), (4, 1.5), (5, 1.3)])
@pytest.mark.parametrize("conv", [_lib.DR_ALGORITHMS, _lib.DR_MULTISTEP, _lib.DR_PROB_METROP, _lib.DR_PROB_HASTINGS])
@pytest.mark.parametrize("lpc", [0, 8])
def test_sample_call_equals_step_call_with_replayed_draws(kind, d, nleap, eps, nprops, f, conv, lpc, monkeypatch):
    if lpc:
        monkeypatch.setenv("HMCB_LPC", str(lpc))
    else:
        monkeypatch.delenv("HMCB_LPC", raising=False)
    C = 77
    q0 = torch.as_tensor(_q0(kind, C, d, 3 * nprops + d), device=_dev())
    P, U = replay.make_draws(500 + d + nprops, 1, C, d, 2 * nprops - 1)
    e = algorithms.DRHMC(_model(kind, d), device=_dev()).engine
    a = e.drhmc_step(q0, nleap, eps, nprops, f, p=P[0], u=U[0], convention=conv, want_p0=True)
    b = e.drhmc_sample(q0, nleap, eps, nprops, f, 1, p=P[0], u=U[0], convention=conv, want_p=True)
    _same(a, b, ("q", "p", "code", "avec", "counts", "ptries"))
    if nprops >= 3 and kind == "funnel":
        assert len(set(a["code"].tolist())) >= 3


@pytest.mark.parametrize("conv", [_lib.DR_ALGORITHMS, _lib.DR_PROB_HASTINGS])
@pytest.mark.parametrize("C", [1, 31, 4096 + 5, 70001, 140003])
def test_sample_call_equals_step_call_with_philox_draws(conv, C):
    d = 50
    q0 = torch.as_tensor(_q0("funnel", C, d, 5), device=_dev())
    a = algorithms.DRHMC(models.Funnel(d), device=_dev(), seed=9, chain_offset=123).engine
    b = algorithms.DRHMC(models.Funnel(d), device=_dev(), seed=9, chain_offset=123).engine
    ra = a.drhmc_step(q0, 12, 0.3, 3, 2.0, convention=conv)
    rb = b.drhmc_sample(q0, 12, 0.3, 3, 2.0, 1, convention=conv, want_p=True)
    _same(ra, rb, ("q", "p", "code", "avec", "counts", "ptries"))


def test_sample_call_matches_the_oracle_directly():
    """Not only equal to hmcb_drhmc_step: the sample call against oracle.samplers.drhmc_step itself."""
    d, C, nleap, eps, nprops, f = 50, 200, 10, 0.35, 3, 2.0
    om = omodels.Funnel(d)
    q0 = _q0("funnel", C, d, 8)
    P, U = replay.make_draws(77, 1, C, d, nprops)
    e = algorithms.DRHMC(models.Funnel(d), device=_dev()).engine
    out = e.drhmc_sample(torch.as_tensor(q0, device=_dev()), nleap, eps, nprops, f, 1, p=P[0], u=U[0], want_p=True)
    ref = [samplers.drhmc_step(om, q0[c], nleap, eps, nprops, f, replay.Draws([P[0, c]], U[0, c])) for c in range(C)]
    assert np.array_equal(out["code"].cpu().numpy(), [r[2] for r in ref])
    assert np.array_equal(out["counts"].cpu().numpy(), [r[4] for r in ref])
    assert np.allclose(out["q"].cpu().numpy(), np.array([r[0] for r in ref]), rtol=1e-10, atol=1e-11)
    assert np.allclose(out["p"].cpu().numpy(), np.array([r[1] for r in ref]), rtol=1e-10, atol=1e-11)


@pytest.mark.parametrize("kind,d,lpc", [("funnel", 50, 0), ("funnel", 10, 0), ("funnel", 50, 4), ("rosenbrock", 20, 0)])
@pytest.mark.parametrize("conv", [_lib.DR_ALGORITHMS, _lib.DR_PROB_METROP])
def test_k_transitions_in_one_call_equal_k_single_calls(kind, d, lpc, conv, monkeypatch):
    if lpc:
        monkeypatch.setenv("HMCB_LPC", str(lpc))
    else:
        monkeypatch.delenv("HMCB_LPC", raising=False)
    C, K, nleap, eps, nprops, f = 5000, 6, 7, (0.4 if kind == "funnel" else 0.08), 3, 2.0
    dev = _dev()
    q0 = torch.as_tensor(_q0(kind, C, d, 11), device=dev)
    a = algorithms.DRHMC(_model(kind, d), device=dev, seed=21, chain_offset=7).engine
    b = algorithms.DRHMC(_model(kind, d), device=dev, seed=21, chain_offset=7).engine
    a.iteration = b.iteration = 40
    q = q0
    s1, s2 = torch.zeros_like(q0), torch.zeros_like(q0)
    counts = torch.zeros((C, 3), dtype=torch.int32, device=dev)
    nacc = torch.zeros(C, dtype=torch.int32, device=dev)
    for k in range(K):
        r = a.drhmc_step(q, nleap, eps, nprops, f, convention=conv)
        q = r["q"]
        a.moments_accumulate(q, s1, s2)
        counts += r["counts"]
        nacc += (r["code"] >= (0 if conv else 1)).int()
    t1, t2 = torch.zeros_like(q0), torch.zeros_like(q0)
    macc = torch.zeros(C, dtype=torch.int32, device=dev)
    rb = b.drhmc_sample(q0, nleap, eps, nprops, f, K, msum=t1, msumsq=t2, n_accept=macc, convention=conv, want_p=True)
    assert b.iteration == a.iteration == 40 + K
    assert torch.equal(rb["q"], q) and torch.equal(rb["code"], r["code"]) and torch.equal(rb["avec"], r["avec"])
    assert torch.equal(rb["p"], r["p"])
    assert torch.equal(rb["counts"], counts) and torch.equal(macc, nacc)
    assert torch.equal(t1, s1) and torch.equal(t2, s2)
    assert (q != q0).any(dim=1).float().mean().item() > 0.5


def test_in_place_and_repeated_launches():
    """q_out == q_in, and the scratch of one call does not leak into the next."""
    d, C = 50, 9000
    dev = _dev()
    q0 = torch.as_tensor(_q0("funnel", C, d, 13), device=dev)
    a = algorithms.DRHMC(models.Funnel(d), device=dev, seed=1).engine
    b = algorithms.DRHMC(models.Funnel(d), device=dev, seed=1).engine
    ref = a.drhmc_sample(q0, 9, 0.3, 3, 2.0, 4)
    ref2 = a.drhmc_sample(ref["q"], 9, 0.3, 3, 2.0, 3)
    q = q0.clone()
    b.drhmc_sample(q, 9, 0.3, 3, 2.0, 4, q_out=q)
    assert torch.equal(q, ref["q"])
    out = b.drhmc_sample(q, 9, 0.3, 3, 2.0, 3, q_out=q)
    assert torch.equal(q, ref2["q"]) and torch.equal(out["counts"], ref2["counts"])


# ------------------------------------------------------------------ K vanilla HMC transitions per call
@pytest.mark.parametrize("kind,d,C", [("funnel", 50, 4096 + 300), ("funnel", 50, 70000), ("funnel", 20, 5000), ("funnel", 10, 5000),
                                      ("rosenbrock", 20, 5000), ("gaussian", 20, 3000)])
@pytest.mark.parametrize("rule", [_lib.RULE_PYHMC, _lib.RULE_ALG])
def test_k_hmc_transitions_in_one_call_equal_k_single_calls(kind, d, C, rule):
    """hmcb_hmc_sample: one fused launch for float64 thread-per-chain shapes (d=50, d=20 funnel: k_hmc_multi_ws), K launches
    inside the call for the others -- either way identical to K hmcb_hmc_step calls."""
    from hmc_b200 import pyhmc
    K, L, eps = 7, 9, (0.25 if kind == "funnel" else 0.05)
    dev = _dev()
    q0 = torch.as_tensor(_q0(kind, C, d, 17), device=dev)
    a = pyhmc.HMC(_model(kind, d), device=dev, seed=31, chain_offset=11).engine
    b = pyhmc.HMC(_model(kind, d), device=dev, seed=31, chain_offset=11).engine
    a.iteration = b.iteration = 3
    q = q0
    nacc = torch.zeros(C, dtype=torch.int32, device=dev)
    for k in range(K):
        r = a.hmc_step(q, L, eps, rule)
        q = r["q"]
        nacc += (r["code"] == 1).int()
    macc = torch.zeros(C, dtype=torch.int32, device=dev)
    rb = b.hmc_sample(q0, L, eps, rule, K, n_accept=macc, want_p=True)
    assert torch.equal(rb["q"], q), "%d chains differ" % (rb["q"] != q).any(dim=1).sum().item()
    assert torch.equal(rb["code"], r["code"]) and torch.equal(rb["Hs"], r["Hs"]) and torch.equal(rb["p"], r["p"])
    assert torch.equal(macc, nacc)
    assert (rb["counts"] == torch.tensor([2 * K, K * (L + 1), K], dtype=torch.int32, device=dev)).all()
    assert 0.3 < nacc.double().mean().item() / K < 1.0


def test_k_hmc_transitions_float32_and_diagonal_metric():
    """The generic fused kernel (k_hmc_multi) on a float32 shape and with a diagonal metric."""
    from hmc_b200 import pyhmc
    d, C, K, L, eps = 10, 3000, 5, 8, 0.1
    dev = _dev()
    q0 = torch.as_tensor(_q0("funnel", C, d, 23), device=dev)
    for kw in (dict(dtype=torch.float32), dict(invmetric_diag=np.linspace(0.5, 2.0, d))):
        a = pyhmc.HMC(models.Funnel(d), device=dev, seed=8, **kw).engine
        b = pyhmc.HMC(models.Funnel(d), device=dev, seed=8, **kw).engine
        q = q0.to(a.dtype)
        for k in range(K):
            r = a.hmc_step(q, L, eps, _lib.RULE_PYHMC)
            q = r["q"]
        rb = b.hmc_sample(q0.to(b.dtype), L, eps, _lib.RULE_PYHMC, K, want_p=True)
        assert torch.equal(rb["q"], q) and torch.equal(rb["code"], r["code"]) and torch.equal(rb["p"], r["p"])


def test_k_hmc_transitions_with_moments_and_in_place():
    from hmc_b200 import pyhmc
    d, C, K, L, eps = 50, 6000, 5, 6, 0.2
    dev = _dev()
    q0 = torch.as_tensor(_q0("funnel", C, d, 19), device=dev)
    a = pyhmc.HMC(models.Funnel(d), device=dev, seed=2).engine
    b = pyhmc.HMC(models.Funnel(d), device=dev, seed=2).engine
    q = q0
    s1, s2 = torch.zeros_like(q0), torch.zeros_like(q0)
    for k in range(K):
        q = a.hmc_step(q, L, eps, _lib.RULE_PYHMC)["q"]
        a.moments_accumulate(q, s1, s2)
    t1, t2 = torch.zeros_like(q0), torch.zeros_like(q0)
    qb = q0.clone()
    b.hmc_sample(qb, L, eps, _lib.RULE_PYHMC, K, msum=t1, msumsq=t2, q_out=qb)
    assert torch.equal(qb, q) and torch.equal(t1, s1) and torch.equal(t2, s2)
    qc = q0.clone()
    c = pyhmc.HMC(models.Funnel(d), device=dev, seed=2).engine
    c.hmc_sample(qc, L, eps, _lib.RULE_PYHMC, K, q_out=qc)          # fused kernel, in place
    assert torch.equal(qc, q)


@pytest.mark.parametrize("conv", [_lib.DR_ALGORITHMS, _lib.DR_MULTISTEP, _lib.DR_PROB_METROP, _lib.DR_PROB_HASTINGS])
@pytest.mark.parametrize("nprops,eps,K", [(3, 0.25, 19), (3, 0.6, 9), (2, 0.45, 8), (1, 0.3, 5), (4, 0.5, 11)])
def test_short_trajectory_sample_mode_equals_single_calls(conv, nprops, eps, K, monkeypatch):
    """hmcb_drhmc_sample at short trajectories (nleap < 48, no moments, d=50 funnel: k_dr0_multi_ws -- chains run in
    registers while the first stage accepts, the stage kernels finish the transitions that do not, pass after pass):
    every output equals K single hmcb_drhmc_step calls, at high and at low acceptance, several chunks of 8."""
    d, C, nleap, f = 50, 6000, 6, 2.0
    monkeypatch.setenv("HMCB_DR_SHORT_MIN", "4096")      # the library uses this form from 2^19 chains on
    dev = _dev()
    q0 = torch.as_tensor(_q0("funnel", C, d, 29), device=dev)
    a = algorithms.DRHMC(models.Funnel(d), device=dev, seed=41, chain_offset=3).engine
    b = algorithms.DRHMC(models.Funnel(d), device=dev, seed=41, chain_offset=3).engine
    a.iteration = b.iteration = 100
    q = q0
    counts = torch.zeros((C, 3), dtype=torch.int32, device=dev)
    nacc = torch.zeros(C, dtype=torch.int32, device=dev)
    stage0 = 0.0
    for k in range(K):
        r = a.drhmc_step(q, nleap, eps, nprops, f, convention=conv)
        q = r["q"]
        counts += r["counts"]
        nacc += (r["code"] >= (0 if conv else 1)).int()
        stage0 += (r["code"] == (0 if conv else 1)).double().mean().item() / K
    macc = torch.zeros(C, dtype=torch.int32, device=dev)
    rb = b.drhmc_sample(q0, nleap, eps, nprops, f, K, n_accept=macc, convention=conv, want_p=True)
    assert torch.equal(rb["q"], q), "%d chains differ" % (rb["q"] != q).any(dim=1).sum().item()
    assert torch.equal(rb["code"], r["code"]) and torch.equal(rb["avec"], r["avec"]) and torch.equal(rb["p"], r["p"])
    if "ptries" in r:
        assert torch.equal(rb["ptries"], r["ptries"])
    assert torch.equal(rb["counts"], counts) and torch.equal(macc, nacc)
    print("first-stage acceptance %.2f" % stage0)


@pytest.mark.parametrize("nleap,conv", [(7, _lib.DR_ALGORITHMS), (60, _lib.DR_PROB_HASTINGS)])
def test_running_moments_of_a_four_range_sample_call(nleap, conv):
    """>= 65,536 chains: several chain ranges (8 here) on as many streams, plain stage-0 kernel, one moments / counters pass per range and
    transition -- the sums must be hmcb_moments_accumulate's, bit for bit (nleap=60 also takes stage 2 through its two
    kernels)."""
    d, C, K = 50, 140003, 3
    dev = _dev()
    q0 = torch.as_tensor(_q0("funnel", C, d, 37), device=dev)
    a = algorithms.DRHMC(models.Funnel(d), device=dev, seed=6).engine
    b = algorithms.DRHMC(models.Funnel(d), device=dev, seed=6).engine
    q = q0
    s1, s2 = torch.zeros_like(q0), torch.zeros_like(q0)
    counts = torch.zeros((C, 3), dtype=torch.int32, device=dev)
    for k in range(K):
        r = a.drhmc_step(q, nleap, 0.3, 3, 2.0, convention=conv)
        q = r["q"]
        a.moments_accumulate(q, s1, s2)
        counts += r["counts"]
    t1, t2 = torch.zeros_like(q0), torch.zeros_like(q0)
    rb = b.drhmc_sample(q0, nleap, 0.3, 3, 2.0, K, msum=t1, msumsq=t2, convention=conv)
    assert torch.equal(rb["q"], q) and torch.equal(rb["counts"], counts) and torch.equal(rb["code"], r["code"])
    assert torch.equal(t1, s1), "%d rows of the running sum differ" % (t1 != s1).any(dim=1).sum().item()
    assert torch.equal(t2, s2)
    codes = set(r["code"].tolist())
    assert len(codes) >= 3


@pytest.mark.parametrize("C,which", [(5003, "hmc"), (1237, "hmc"), (5003, "hmcm"), (5003, "dr"), (140003, "dr")])
def test_sample_calls_stay_inside_caller_buffers(C, which):
    """The caller-supplied outputs of the sample calls (state, running moments, accept counts, and for DR the code and
    counters) placed inside larger sentinel-filled allocations, ragged chain counts (a partial last warp; 5003 takes the
    warp-specialised fused HMC kernel, 1237 the generic one, "hmcm" the per-transition path with moments, 140003 the
    range-pipelined DR path): nothing outside the
    windows may change, and the windows equal an unpadded call."""
    d, K, G = 50, 3, 211
    dev = _dev()
    q0 = torch.as_tensor(_q0("funnel", C, d, 5), device=dev)
    S = -777.0

    def padded(shape, dtype, fill=None):
        n = int(np.prod(shape))
        sent = S if dtype.is_floating_point else -777
        buf = torch.full((n + 2 * G,), sent, dtype=dtype, device=dev)
        win = buf[G:G + n].view(shape)
        if fill is not None:
            win.fill_(fill)
        return buf, win, sent
    P = {"q": padded((C, d), torch.float64), "s1": padded((C, d), torch.float64, 0.0), "s2": padded((C, d), torch.float64, 0.0),
         "na": padded((C,), torch.int32, 0)}
    if which == "hmc":
        a = algorithms.HMC(models.Funnel(d), device=dev, seed=3).engine
        b = algorithms.HMC(models.Funnel(d), device=dev, seed=3).engine
        ref = a.hmc_sample(q0, 6, 0.2, _lib.RULE_ALG, K)
        out = b.hmc_sample(q0, 6, 0.2, _lib.RULE_ALG, K, n_accept=P["na"][1], q_out=P["q"][1])
    elif which == "hmcm":       # with running moments: one launch per transition plus the moments / counters pass
        a = algorithms.HMC(models.Funnel(d), device=dev, seed=3).engine
        b = algorithms.HMC(models.Funnel(d), device=dev, seed=3).engine
        ref = a.hmc_sample(q0, 6, 0.2, _lib.RULE_ALG, K, msum=torch.zeros_like(q0), msumsq=torch.zeros_like(q0))
        out = b.hmc_sample(q0, 6, 0.2, _lib.RULE_ALG, K, msum=P["s1"][1], msumsq=P["s2"][1], n_accept=P["na"][1], q_out=P["q"][1])
    else:
        P["code"] = padded((C,), torch.int32)
        P["counts"] = padded((C, 3), torch.int32)
        a = algorithms.DRHMC(models.Funnel(d), device=dev, seed=3).engine
        b = algorithms.DRHMC(models.Funnel(d), device=dev, seed=3).engine
        ref = a.drhmc_sample(q0, 7, 0.3, 3, 2.0, K)
        out = b.drhmc_sample(q0, 7, 0.3, 3, 2.0, K, msum=P["s1"][1], msumsq=P["s2"][1], n_accept=P["na"][1], q_out=P["q"][1],
                             code_out=P["code"][1], counts_out=P["counts"][1])
        assert torch.equal(out["code"], ref["code"]) and torch.equal(out["counts"], ref["counts"])
    torch.cuda.synchronize()
    assert torch.equal(out["q"], ref["q"]) and torch.equal(P["q"][1], ref["q"])
    for k, (buf, win, sent) in P.items():
        assert bool((buf[:G] == sent).all()) and bool((buf[-G:] == sent).all()), "wrote outside " + k
    assert int(P["na"][1].sum()) > 0 and bool((P["s2"][1] >= 0).all())
