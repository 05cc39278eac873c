"""Size-independent properties of the CUDA path at benchmark-scale batches:
leapfrog reversibility, work counters, determinism and sharding invariance of
the counter-based RNG, DR bookkeeping identities, and stationarity of the
funnel under the production (Philox) path."""
import numpy as np
import pytest
import torch

from hmc_b200 import _lib, algorithms, models, pyhmc

pytestmark = pytest.mark.gpu


def _dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


def _exact_funnel(C, d, seed, dev):
    g = torch.Generator(device="cpu").manual_seed(seed)
    v = 3.0 * torch.randn(C, 1, generator=g, dtype=torch.float64)
    th = torch.exp(v / 2) * torch.randn(C, d - 1, generator=g, dtype=torch.float64)
    return torch.cat([v, th], dim=1).to(dev)


@pytest.mark.parametrize("kind,d", [("funnel", 50), ("gaussian", 100), ("rosenbrock", 20)])
def test_leapfrog_reversibility_65536_chains(kind, d):
    dev = _dev()
    model = {"funnel": lambda: models.Funnel(d), "gaussian": lambda: models.CorrGaussian.ar1(d),
             "rosenbrock": lambda: models.Rosenbrock(d)}[kind]()
    s = pyhmc.HMC(model, device=dev)
    C = 65536
    g = torch.Generator(device="cpu").manual_seed(1)
    q0 = (torch.rand(C, d, generator=g, dtype=torch.float64) * 2 - 1).to(dev)
    if kind == "rosenbrock":
        q0 = 1.0 + 0.1 * q0
    p0 = torch.randn(C, d, generator=g, dtype=torch.float64).to(dev)
    eps = 0.02 if kind != "gaussian" else 0.05
    q1, p1 = s.leapfrog(q0, p0, 20, eps)
    q2, p2 = s.leapfrog(q1, -p1, 20, eps)
    # flip and integrate back: the property delayed rejection relies on (src/algorithms.py:522)
    assert (q2 - q0).abs().max().item() < 1e-9
    assert (p2 + p0).abs().max().item() < 1e-9
    assert (q1 - q0).abs().max().item() > 1e-3


def test_counts_codes_and_determinism_1m_chains():
    dev = _dev()
    d, C, N = 50, 1 << 20, 5
    q0 = _exact_funnel(C, d, 3, dev)
    a = pyhmc.HMC(models.Funnel(d), device=dev, seed=11)
    b = pyhmc.HMC(models.Funnel(d), device=dev, seed=11)
    qa, pa, ca, Ha, na = a.step(q0, N, 0.05)
    qb, pb, cb, Hb, nb = b.step(q0, N, 0.05)
    assert torch.equal(qa, qb) and torch.equal(ca, cb) and torch.equal(pa, pb)
    assert (na == torch.tensor([2, N + 1, 1], device=dev, dtype=na.dtype)).all()
    assert set(torch.unique(ca).tolist()) <= {0.0, 1.0, 2.0}
    acc = (ca == 1.0)
    assert torch.equal(qa[~acc], q0[~acc])            # rejected chains keep their state
    assert (qa[acc] != q0[acc]).any(dim=1).all()      # accepted chains moved
    # a different seed gives different momenta
    c2 = pyhmc.HMC(models.Funnel(d), device=dev, seed=12)
    assert not torch.equal(c2.step(q0, N, 0.05)[1], pa)
    # second transition uses fresh draws
    assert not torch.equal(a.step(q0, N, 0.05)[1], pa)


def test_sharding_invariance_of_rng():
    """Chains keyed by global id: one call over [0,C) == two calls over halves with chain_offset."""
    dev = _dev()
    d, C = 50, 4096
    q0 = _exact_funnel(C, d, 5, dev)
    full = algorithms.DRHMC(models.Funnel(d), device=dev, seed=7)
    qf, pf, cf, af, nf = full.step(q0, 8, 0.3, 3, 2.0)
    h = C // 2 + 37
    lo = algorithms.DRHMC(models.Funnel(d), device=dev, seed=7, chain_offset=0)
    hi = algorithms.DRHMC(models.Funnel(d), device=dev, seed=7, chain_offset=h)
    ql, pl, cl, al, nl = lo.step(q0[:h], 8, 0.3, 3, 2.0)
    qh, ph, chh, ah, nh = hi.step(q0[h:], 8, 0.3, 3, 2.0)
    assert torch.equal(torch.cat([ql, qh]), qf)
    assert torch.equal(torch.cat([cl, chh]), cf)
    assert torch.equal(torch.cat([nl, nh]), nf)
    assert torch.equal(torch.cat([al, ah]), af)


def test_lane_shape_does_not_change_the_draws(monkeypatch):
    """The momentum of element i is a function of (seed, chain, iteration, i) only."""
    dev = _dev()
    d, C = 50, 512
    q0 = _exact_funnel(C, d, 6, dev)
    ps = []
    for lpc in ("1", "2", "8", "32"):
        monkeypatch.setenv("HMCB_LPC", lpc)
        s = pyhmc.HMC(models.Funnel(d), device=dev, seed=3)
        q, p, code, Hs, n = s.step(q0, 1, 1e-9)   # a tiny step: p_out is p0 up to 1e-9
        ps.append(p)
    for p in ps[1:]:
        assert (p - ps[0]).abs().max().item() < 1e-6


def test_drhmc_bookkeeping_identities():
    dev = _dev()
    d, C, nleap, nprops, f = 50, 1 << 17, 10, 3, 2.0
    q0 = _exact_funnel(C, d, 8, dev)
    s = algorithms.DRHMC(models.Funnel(d), device=dev, seed=2)
    q, p, code, avec, counts = s.step(q0, nleap, 0.35, nprops, f)
    L = [int(nleap * f ** j) for j in range(nprops)]
    # stage-0 acceptances did exactly one trajectory, two energy evaluations
    m0 = code == 1
    assert m0.any() and (counts[m0] == torch.tensor([2, L[0] + 1, 1], device=dev, dtype=counts.dtype)).all()
    # stage-1 acceptances: proposals L0, L1 and one ghost of L0 -> 3 leapfrogs, 4 energies
    m1 = code == 2
    assert m1.any() and (counts[m1] == torch.tensor([4, 2 * (L[0] + 1) + L[1] + 1, 3], device=dev, dtype=counts.dtype)).all()
    # later stages: 5 or 7 leapfrog calls depending on the ghost's early exit (src/algorithms.py:494-496)
    m2 = (code == 3) | (code == 0)
    assert m2.any() and set(torch.unique(counts[m2][:, 2]).tolist()) <= {5, 7}
    assert (counts[:, 0] == counts[:, 2] + 1).all()     # one energy per trajectory plus H0
    # avec: probabilities in [0,1]; entries after the accepting stage stay 0; accepted stage has avec > 0
    assert (avec >= 0).all() and (avec <= 1).all()
    for j in range(nprops):
        acc_j = code == j + 1
        if acc_j.any():
            assert (avec[acc_j][:, j] > 0).all() and (avec[acc_j][:, j + 1:] == 0).all()
    rej = code == 0
    assert torch.equal(q[rej], q0[rej])
    assert (avec[rej] < 1).all()


def test_funnel_stationarity_production_path():
    """Start from exact funnel draws; HMC and DRHMC must leave the distribution of v
    (mean 0, std 3) unchanged (scripts_old/tests/test_drhmc_funnel.py:244-254)."""
    dev = _dev()
    d, C = 20, 1 << 16
    for make, args in ((lambda: pyhmc.HMC(models.Funnel(d), device=dev, seed=5), (10, 0.1)),
                       (lambda: algorithms.DRHMC(models.Funnel(d), device=dev, seed=6), (10, 0.2, 3, 2.0))):
        s = make()
        q = _exact_funnel(C, d, 9, dev)
        nacc = 0.0
        for _ in range(15):
            out = s.step(q, *args)
            q = out[0]
            nacc += (out[2] > 0).double().mean().item()
        v = q[:, 0]
        se = 3.0 / np.sqrt(C)
        assert abs(v.mean().item()) < 6 * se, v.mean().item()
        assert abs(v.std().item() - 3.0) < 0.06, v.std().item()
        th = q[:, 1:] * torch.exp(-v / 2)[:, None]   # should be N(0,1)
        assert abs(th.std().item() - 1.0) < 0.02
        assert nacc / 15 > 0.5


def test_dual_averaging_batched_adapts_to_target():
    dev = _dev()
    d, C = 20, 8192
    s = algorithms.HMC(models.CorrGaussian.ar1(d), device=dev, seed=4)
    q0 = torch.randn(C, d, dtype=torch.float64, device=dev)
    s.step_size = 0.05
    q = s.adapt_stepsize(q0, 150, nsamples=1, burnin=0, step_size=0.05, nleap=10)
    assert 0.05 < s.step_size < 2.0
    out = s.engine.hmc_step(q, 10, s.step_size, s._rule, want_p=False, want_prob=True)
    assert abs(out["prob"].mean().item() - 0.65) < 0.1


def test_host_pipeline_equals_one_shot_call():
    """DRHMC.step on a pinned host batch (chunked H2D/compute/D2H pipeline) gives exactly
    what the device-resident call gives: chains are independent and draws are keyed by
    global chain id."""
    dev = _dev()
    d, C = 50, 50_000
    q0 = _exact_funnel(C, d, 11, dev)
    a = algorithms.DRHMC(models.Funnel(d), device=dev, seed=21)
    b = algorithms.DRHMC(models.Funnel(d), device=dev, seed=21)
    qa, pa, ca, aa, na = a.step(q0, 8, 0.3, 3, 2.0)
    qh = q0.cpu().pin_memory()
    qb, pb, cb, ab, nb = b.step(qh, 8, 0.3, 3, 2.0)
    assert qb.device.type == "cpu" and qb.is_pinned()
    assert torch.equal(qa.cpu(), qb) and torch.equal(pa.cpu(), pb) and torch.equal(ca.cpu(), cb)
    assert torch.equal(aa.cpu(), ab) and torch.equal(na.cpu(), nb)
    # second transition: iteration counter advanced once per call, not once per chunk
    qa2 = a.step(qa, 8, 0.3, 3, 2.0)[0]
    qb2 = b.step(qb, 8, 0.3, 3, 2.0)[0]
    assert torch.equal(qa2.cpu(), qb2)
    # without the momentum (what a sample loop needs): same chain, p is None
    qa3 = a.step(qa2, 8, 0.3, 3, 2.0)[0]
    qb3, pb3, cb3, _, _ = b.step(qb2, 8, 0.3, 3, 2.0, want_p=False)
    assert pb3 is None and torch.equal(qa3.cpu(), qb3)


@pytest.mark.parametrize("d,C", [(50, 4097), (26, 1000), (100, 777)])
def test_state_io_paths_agree_and_work_in_place(d, C, monkeypatch):
    """The single-trajectory kernels read and write row-major chain state either lane-per-row or
    through a per-warp shared-memory tile (coalesced), and the thread-per-chain float64 shapes have a
    warp-specialised form: same transition, same draws -> identical results, also when q_out is q_in
    and on a ragged last warp."""
    dev = _dev()
    q0 = _exact_funnel(C, d, 5, dev)
    outs = []
    for env in ({}, {"HMCB_NO_TILE": "1"}, {"HMCB_NO_WS": "1"}, {"HMCB_NO_WS": "1", "HMCB_NO_TILE": "1"}):
        for k in ("HMCB_NO_TILE", "HMCB_NO_WS"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        s = pyhmc.HMC(models.Funnel(d), device=dev, seed=9)
        o = s.engine.hmc_step(q0, 7, 0.25, _lib.RULE_PYHMC)
        outs.append((o["q"].clone(), o["p"].clone(), o["code"].clone()))
        # in place
        s2 = pyhmc.HMC(models.Funnel(d), device=dev, seed=9)
        qi = q0.clone()
        o2 = s2.engine.hmc_step(qi, 7, 0.25, _lib.RULE_PYHMC, q_out=qi)
        assert torch.equal(qi, o["q"]) and torch.equal(o2["code"], o["code"])
        # delayed rejection, stage 0 through the same paths
        dr = algorithms.DRHMC(models.Funnel(d), device=dev, seed=9)
        od = dr.engine.drhmc_step(q0, 7, 0.6, 3, 2.0)
        outs[-1] += (od["q"].clone(), od["code"].clone(), od["counts"].clone())
    for o in outs[1:]:
        for a, b in zip(outs[0], o):
            assert torch.equal(a, b)
    codes = outs[0][2]
    assert (codes == 1).any() and (codes != 1).any()         # accepted and rejected rows both went through the tile


@pytest.mark.parametrize("d", [10, 50])
def test_structure_of_arrays_state_gives_the_same_transitions(d):
    """The C ABI addresses chain state through (stride_chain, stride_dim): a [d][C] structure-of-arrays
    buffer, seen as a [C, d] view with strides (1, C), goes through the same kernels as row-major state
    and must give the same chains (vanilla step and all delayed-rejection stages)."""
    dev = _dev()
    C = 5000
    q0 = _exact_funnel(C, d, 6, dev)
    soa = q0.t().contiguous()                    # [d, C]
    qs = soa.t()                                  # [C, d] view, strides (1, C)
    assert qs.stride() == (1, C)
    a = pyhmc.HMC(models.Funnel(d), device=dev, seed=4)
    b = pyhmc.HMC(models.Funnel(d), device=dev, seed=4)
    oa = a.engine.hmc_step(q0, 9, 0.2, _lib.RULE_PYHMC)
    ob = b.engine.hmc_step(qs, 9, 0.2, _lib.RULE_PYHMC)
    assert ob["q"].stride() == (1, C)
    for k in ("q", "p", "code", "counts"):
        assert torch.equal(oa[k], ob[k]), k
    da = algorithms.DRHMC(models.Funnel(d), device=dev, seed=4)
    db = algorithms.DRHMC(models.Funnel(d), device=dev, seed=4)
    ra = da.engine.drhmc_step(q0, 9, 0.6, 3, 2.0)
    rb = db.engine.drhmc_step(qs, 9, 0.6, 3, 2.0)
    for k in ("q", "p", "code", "avec", "counts"):
        assert torch.equal(ra[k], rb[k]), k
    assert len(torch.unique(ra["code"])) >= 3                 # later stages ran too
